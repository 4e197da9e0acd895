/* dcv.h — C-ABI of the B200-native deconvnet / DeepDream hot path.
 *
 * The reference (rashanarshad/deconv_api) is pure Python on Keras/TensorFlow-CPU; it has no FFI of
 * its own.  Each entry point below replaces one piece of reference app/deepdream.py (file:line
 * cited per function) and is what a ctypes stub inside that file would bind (INTEGRATION.md shows
 * the stubs).  Conventions:
 *   - plain pointers and sizes only; tensors are dense, C-contiguous, NHWC float32 unless stated;
 *     conv kernels are HWIO (kh,kw,Cin,Cout) exactly as Keras `layer.get_weights()` returns them;
 *   - `*_host` / unsuffixed data pointers are HOST memory borrowed for the duration of the call;
 *     `*_dev` entry points take DEVICE pointers on the context's device;
 *   - every function returns 0 on success or a negative dcv_status; `dcv_last_error` gives the text;
 *   - one context per GPU; a context is thread-compatible, not thread-safe; all work runs on the
 *     context's own CUDA stream and every host-pointer call returns after the result is complete;
 *   - there is NO CPU fallback: without a CUDA device `dcv_create` fails.
 */
#ifndef DCV_H_
#define DCV_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define DCV_API __attribute__((visibility("default")))
#else
#define DCV_API
#endif

typedef struct dcv_ctx dcv_ctx;

enum dcv_status {
  DCV_OK = 0,
  DCV_ERR_ARG = -1,      /* bad argument (null pointer, negative size, unknown layer, odd unpool size) */
  DCV_ERR_CUDA = -2,     /* CUDA runtime / driver error, text in dcv_last_error */
  DCV_ERR_STATE = -3,    /* call order (e.g. project before forward) */
  DCV_ERR_NOMEM = -4,
  DCV_ERR_UNSUPPORTED = -5
};

/* visualize_mode of visualize_all_layers (deepdream.py:383,454-460) */
enum dcv_vis_mode { DCV_VIS_ALL = 0, DCV_VIS_MAX = 1 };

/* arithmetic of the backward projection chain.  FP32 (parity mode, rel-L2 <= 1e-5): fp32 data between the layers;
 * the convs run on the tcgen05 tensor cores with every projection scaled by a power of two into fp16's range and both
 * operands split into two fp16 planes (22 significand bits), the three significant cross products accumulated in
 * fp32 registers (DCV_FP32_SCHEME=0: three bf16 planes / six products as in the forward pass; DCV_FP32_TC=0 or
 * dcv_set_fp32_mode(ctx, 0): the CUDA-core FFMA kernels).  BF16 (throughput mode, rel-L2 <= 2e-2): tcgen05 chain on
 * bf16 operands with fp32 accumulation in TMEM.  The forward tape and the switches are fp32-accurate in both. */
enum dcv_precision { DCV_FP32 = 0, DCV_BF16 = 1 };

/* element type of the projections returned by dcv_project_fmt.  U8 = deprocess_image (deepdream.py:483-498)
 * applied to every projection separately, which is what the reference's only consumer does next (main.py:72). */
enum dcv_out_format { DCV_OUT_F32 = 0, DCV_OUT_BF16 = 1, DCV_OUT_U8 = 2 };

/* activation ids for dcv_net_add_dense */
enum dcv_act { DCV_ACT_LINEAR = 0, DCV_ACT_RELU = 1, DCV_ACT_SOFTMAX = 2 };

/* ---- life cycle ------------------------------------------------------------------------------ */
DCV_API int dcv_version(void);
/* Creates a context on CUDA device `device` (stream, workspace).  Replaces nothing in the reference
 * (TensorFlow session creation, main.py:17). */
DCV_API int dcv_create(int device, dcv_ctx** out);
DCV_API void dcv_destroy(dcv_ctx* ctx);
/* Text of the last error on this context (or of the last failed dcv_create when ctx == NULL). */
DCV_API const char* dcv_last_error(const dcv_ctx* ctx);
DCV_API int dcv_device(const dcv_ctx* ctx);
/* Blocks until all work queued on the context's stream has finished. */
DCV_API int dcv_sync(dcv_ctx* ctx);
/* Number of kernels this context has launched since creation (bench.py's gpu_launches). */
DCV_API long long dcv_launch_count(const dcv_ctx* ctx);
/* The context's cudaStream_t, as an opaque pointer (for event timing by the caller). */
DCV_API void* dcv_stream(const dcv_ctx* ctx);

/* Page-locked host memory for result buffers (no counterpart in the reference): a device-to-host copy into ordinary
 * pageable memory is staged by the driver at a fraction of the PCIe rate, into these buffers it runs at full rate and
 * overlaps the kernels.  Free with dcv_host_free before dcv_destroy. */
DCV_API int dcv_host_alloc(dcv_ctx* ctx, size_t bytes, void** out);
DCV_API int dcv_host_free(dcv_ctx* ctx, void* p);

/* Kernel classes for dcv_profile_read. */
enum dcv_kernel_class {
  DCV_K_CONV_FP32 = 0,  /* FFMA 3x3 conv / backward conv */
  DCV_K_CONV_TC = 1,    /* tcgen05 implicit-GEMM backward conv */
  DCV_K_POOL = 2,       /* max-pool + switch */
  DCV_K_UNPOOL = 3,     /* switch-driven unpool (tensor and single-map forms) */
  DCV_K_RANK1 = 4,      /* rank-1 first backward conv */
  DCV_K_OTHER = 5,      /* seed gather, layout/precision conversion, sums, de-process, dense */
  DCV_K_CONV_FWD = 6,   /* tcgen05 forward conv with split operands (6 MMA products per algorithmic flop) */
  DCV_K_NUM = 7
};
/* enable != 0: record a CUDA event pair on the context's stream around every kernel launch (cleared on
 * every call).  dcv_profile_read synchronises and returns, for one class, the summed device time in
 * ms, the number of launches and the algorithmic flops / bytes those launches were given. */
DCV_API int dcv_profile(dcv_ctx* ctx, int enable);
DCV_API int dcv_profile_read(dcv_ctx* ctx, int kernel_class, double* ms, long long* launches, double* flops, double* bytes);
/* The index-th recorded launch since dcv_profile(ctx, 1), in launch order; DCV_ERR_ARG past the end. */
DCV_API int dcv_profile_launch(dcv_ctx* ctx, int index, int* kernel_class, double* ms, double* flops, double* bytes);

/* ---- network description: replaces the wrapper stack of visualize_all_layers (deepdream.py:401-423)
 * and DConvolution2D.__init__ (deepdream.py:58-89), which the reference rebuilds on every request.
 * Layers are appended in model order; the returned value (>= 0) is the layer index. ------------- */
DCV_API int dcv_net_clear(dcv_ctx* ctx);
/* 3x3, stride 1, padding 'same'.  Does the one-time weight preparation of deepdream.py:80-88: keeps W
 * for the forward pass and builds the transposed + spatially flipped kernel (zero bias) for the
 * backward pass, in fp32 and in the bf16 tensor-core packing.  relu != 0: activation='relu' in both
 * directions (the reference's down model inherits the forward activation, deepdream.py:69,86). */
DCV_API int dcv_net_add_conv(dcv_ctx* ctx, const float* w_hwio, const float* bias, int cin, int cout, int relu);
/* 2x2 stride-2 max pool with switches (DPooling, deepdream.py:114-209). */
DCV_API int dcv_net_add_pool(dcv_ctx* ctx);
/* Flatten (DFlatten, deepdream.py:324-366). */
DCV_API int dcv_net_add_flatten(dcv_ctx* ctx);
/* Dense layer, W is (nin, nout) row-major as Keras stores it (DDense, deepdream.py:264-321): forward
 * act(x W + b); backward x W^T with no bias and no activation (deepdream.py:289-297). */
DCV_API int dcv_net_add_dense(dcv_ctx* ctx, const float* w, const float* bias, int nin, int nout, int act);
DCV_API int dcv_net_num_layers(const dcv_ctx* ctx);

/* ---- forward tape: replaces the forward pass of visualize_all_layers (deepdream.py:426-428), i.e.
 * DConvolution2D.up (:91-100), DActivation.up (:238-248) and DPooling.up (:129-139,152-188).
 * Runs layers 0..upto_layer on a batch of B images (B,H,W,C0) and keeps every layer's output (and
 * every pool's switches) on the device until the next dcv_forward.
 * dcv_forward returns as soon as the host image has been copied (it may then be reused); the tape's kernels are queued on
 * the context's stream and every later call on the context is ordered behind them (a kernel error surfaces there). */
DCV_API int dcv_forward(dcv_ctx* ctx, const float* img_nhwc, int B, int H, int W, int upto_layer);
/* Arithmetic of the forward convs (no counterpart in the reference).  1 (default): tcgen05 with both operands
 * split into three bf16 planes and the six significant cross products (x*w = hh + hm + mh + hl + lh + mm) summed in
 * fp32 REGISTERS - every pipeline stage starts a fresh TMEM accumulator, because the tensor core truncates when it
 * adds into one: measured rel-L2 1e-7 .. 1e-6 per layer against float64, the same as the FFMA kernel, at 3-10x its
 * speed for batch 1; the first conv (K = 27) stays FFMA.  0: FFMA fp32 everywhere. */
DCV_API int dcv_set_forward_mode(dcv_ctx* ctx, int mode);
/* Arithmetic of the backward convs of the DCV_FP32 chain (projection, DeepDream, dcv_deconv3x3); no counterpart in
 * the reference.  1 (default, or DCV_FP32_TC in the environment): split-operand tcgen05 kernels with fp32 register
 * accumulation (two fp16 planes per operand and a per-projection power-of-two scale, see dcv_precision); 0: CUDA-core FFMA.  Both meet rel-L2 <= 1e-5; the switch exists for A/B parity tests. */
DCV_API int dcv_set_fp32_mode(dcv_ctx* ctx, int mode);
DCV_API int dcv_forward_dev(dcv_ctx* ctx, const float* img_nhwc_dev, int B, int H, int W, int upto_layer);
/* Shape of layer `layer`'s output in the current tape: dims[0..3] = B,H,W,C (H=W=1 after flatten). */
DCV_API int dcv_layer_shape(const dcv_ctx* ctx, int layer, int dims[4]);
/* Copy a layer's output (`up_data` of the reference wrappers) to the host: B*H*W*C floats. */
DCV_API int dcv_read_activation(dcv_ctx* ctx, int layer, float* out);
/* Copy a pool layer's switches to the host: one byte per POOLED element, (B,H/2,W/2,C), value
 * r*2+c = position of the maximum inside its 2x2 patch (first maximum in row-major order, the rule of
 * deepdream.py:180-183).  The reference's float64 one-hot mask (deepdream.py:162,184-187) is the
 * expansion of this index. */
DCV_API int dcv_read_switch(dcv_ctx* ctx, int layer, uint8_t* out);
/* Per-channel sums over (B,H,W) of a layer's output, accumulated in fp64: the quantity
 * find_top_filters ranks (deepdream.py:369-380).  out: C doubles. */
DCV_API int dcv_channel_sums(dcv_ctx* ctx, int layer, double* out);
/* The same reduction on a caller-supplied host matrix of `rows` rows x C channels (what
 * find_top_filters receives when it is called on an ndarray, deepdream.py:369). */
DCV_API int dcv_channel_sums_buf(dcv_ctx* ctx, const float* x, size_t rows, int C, double* out);

/* ---- projection: replaces the masking + backward chain of visualize_all_layers (deepdream.py:448-473),
 * i.e. DConvolution2D.down (:102-111), DActivation.down (:251-261), DPooling.down (:141-150,191-209).
 * For each k < n: take feature map feat[k] of layer `layer` for ALL B images of the tape (in
 * DCV_VIS_MAX mode keep only the positions equal to the map's maximum, ties kept, :454-457), zero
 * every other map, and run the chain down to the input.  out: n * B*H0*W0*C0 floats.
 * The zeroed tensor is never materialised: the first backward conv reads the one map. ---------- */
DCV_API int dcv_project(dcv_ctx* ctx, int layer, const int* feat, int n, int vis_mode, int precision, float* out);
/* Same with the result left in device memory (n*B*H0*W0*C0 floats at out_dev). */
DCV_API int dcv_project_dev(dcv_ctx* ctx, int layer, const int* feat, int n, int vis_mode, int precision, float* out_dev);

/* The same projections in a compact result format (dcv_out_format): BF16 halves and U8 quarters the bytes that leave
 * the GPU.  U8: every (feature, image) projection is de-processed with its own mean / std, exactly deprocess_image
 * (deepdream.py:483-498) called on that projection, as main.py:72 does with the reference's result.  out: n*B*H0*W0*C0
 * elements of the chosen type (bf16 = the upper 16 bits of the float, round to nearest even). */
DCV_API int dcv_project_fmt(dcv_ctx* ctx, int layer, const int* feat, int n, int vis_mode, int precision, int out_format,
                            void* out);
DCV_API int dcv_project_fmt_dev(dcv_ctx* ctx, int layer, const int* feat, int n, int vis_mode, int precision, int out_format,
                                void* out_dev);

/* ---- request tail: replaces what return_deconv does between cv2.resize and cv2.imencode (main.py:60-72) in ONE call:
 * img_to_array + preprocess_input 'caffe' (channel flip, mean subtraction; main.py:60-61) of the uint8 (H,W,3) image
 * that cv2 decoded, forward tape up to `layer`, find_top_filters (deepdream.py:369-380), projection of the four
 * strongest maps (main.py:64,67-68 use [0..3] of the top-8 list), 2x2 mosaic (main.py:67-69: maps 0|1 over 2|3) and
 * deprocess_image over the whole mosaic (main.py:72).  out_mosaic: (2H,2W,3) uint8; feats_out (optional): the four
 * feature indices.  DCV_ERR_STATE when fewer than four maps have a positive sum (the reference raises IndexError).
 * Only 3*H*W bytes go to the GPU and 12*H*W come back; the tape stays valid for later dcv_project calls. */
DCV_API int dcv_request_mosaic(dcv_ctx* ctx, const uint8_t* img_bgr, int H, int W, int layer, int vis_mode, int precision,
                               uint8_t* out_mosaic, int* feats_out);

/* ---- image tail: replaces deprocess_image (deepdream.py:483-498): (x-mean)/(std+1e-7)*0.1+0.5,
 * clip to [0,1], *255, clip, truncate to uint8; population std.  n elements in, n bytes out.
 * The reference mutates its argument in place up to the "+= 0.5" step (:485-490); with mutate != 0
 * the host buffer x is overwritten with those normalised values, as a drop-in caller would see. --- */
DCV_API int dcv_deprocess(dcv_ctx* ctx, float* x, size_t n, uint8_t* out, int mutate);
DCV_API int dcv_deprocess_dev(dcv_ctx* ctx, const float* x_dev, size_t n, uint8_t* out_dev);

/* De-process without any GPU allocation (no counterpart in the reference, which never touches a GPU).
 * For callers that cannot call cudaMalloc: a server whose memory pool owns the whole device, a request
 * path that cannot afford the device-wide synchronisation cudaMalloc/cudaFree imply, a sandbox that only
 * admits pre-declared buffers.  Same arithmetic and same result as dcv_deprocess_dev, but the reduction
 * partials live in a fixed 16 KiB array inside the library image (one copy per GPU, held until the
 * process exits) instead of the context's growable scratch buffer.
 *   - x_dev (16-byte aligned) and out_dev (4-byte aligned) are device pointers supplied by the caller;
 *     n <= DCV_DEPROCESS_NOALLOC_MAX_ELEMS, larger inputs are refused with DCV_ERR_ARG;
 *   - the first call on each device looks up the array's device address (and may make the CUDA runtime load
 *     the library's device code onto that GPU); make that call before any stream capture begins.  From the
 *     second call on nothing is allocated.  The address is cached for the life of the process: do not call
 *     cudaDeviceReset while the library is in use;
 *   - all calls (from every context on that device) share the one array: two calls whose kernels can
 *     overlap on the GPU race on it — issue them on one stream or order them with events;
 *   - the call is asynchronous on the context's stream, like dcv_deprocess_dev. */
#define DCV_DEPROCESS_NOALLOC_MAX_PARTIALS 1024
#define DCV_DEPROCESS_NOALLOC_MAX_ELEMS (1ull << 28)
DCV_API int dcv_deprocess_dev_noalloc(dcv_ctx* ctx, const float* x_dev, size_t n, uint8_t* out_dev);

/* ---- standalone layer operators on host buffers (unit parity; used by the D* wrapper classes) --- */
/* DPooling.up (deepdream.py:129-139,152-188).  x (B,H,W,C) -> pooled (B,H/2,W/2,C) float32 (the
 * reference's float64 holds the same values), sw (B,H/2,W/2,C) bytes as in dcv_read_switch. */
DCV_API int dcv_maxpool_switch(dcv_ctx* ctx, const float* x, int B, int H, int W, int C, float* pooled, uint8_t* sw);
/* DPooling.down (deepdream.py:141-150,191-209).  x (B,H/2,W/2,C), sw same shape -> out (B,H,W,C).
 * H and W must be even (the reference raises on odd sizes). */
DCV_API int dcv_unpool_switch(dcv_ctx* ctx, const float* x, const uint8_t* sw, int B, int H, int W, int C, float* out);
/* DConvolution2D.up (deepdream.py:91-100): out = act(conv3x3_same(x, W) + b).  Runs the kernel the forward tape
 * would use for this shape (dcv_set_forward_mode; tensor cores from 16 input channels up). */
DCV_API int dcv_conv3x3(dcv_ctx* ctx, const float* x, const float* w_hwio, const float* bias,
                int B, int H, int W, int cin, int cout, int relu, float* out);
/* DConvolution2D.down (deepdream.py:102-111): out = act(conv3x3_same(x, flipT(W))), zero bias;
 * x has `cout` channels, out has `cin`.  precision: DCV_BF16 = the bf16 tcgen05 kernels, DCV_FP32 = the kernels of
 * the fp32 chain (dcv_set_fp32_mode: split-operand tcgen05 by default, FFMA otherwise). */
DCV_API int dcv_deconv3x3(dcv_ctx* ctx, const float* x, const float* w_hwio,
                  int B, int H, int W, int cin, int cout, int relu, int precision, float* out);

/* The 64-channel tail of a DCV_BF16 projection as ONE kernel (unit-parity hook of csrc/tc_strip.cuh): optional
 * DPooling.down (deepdream.py:141-150,191-209; sw != NULL: x is the pooled tensor (B,H/2,W/2,c2) and sw its switches),
 * DConvolution2D.down (deepdream.py:102-111) of a conv with forward kernel w2 (3,3,cmid,c2), and optionally the same of
 * the conv below it, w1 (3,3,c0,cmid) (w1 == NULL: out = (B,H,W,cmid), else (B,H,W,c0)).  bf16 operands, fp32
 * accumulation; the intermediate between the two convs is rounded to bf16 as in the unfused chain.  Needs
 * cmid == c2 == 64 and c0 <= 3 (VGG16 block1), DCV_ERR_UNSUPPORTED otherwise. */
DCV_API int dcv_deconv_tail(dcv_ctx* ctx, const float* x, const uint8_t* sw, int B, int H, int W, const float* w2_hwio, int cmid,
                            int c2, int relu2, const float* w1_hwio, int c0, int relu1, float* out);

/* DActivation.up/.down (deepdream.py:212-261): the layer's activation (dcv_act id) applied to `rows`
 * rows of `n` values; softmax runs along each row. */
DCV_API int dcv_activation(dcv_ctx* ctx, const float* x, int rows, int n, int act, float* out);
/* DDense.up (deepdream.py:300-309): out = act(x W + b), x (B,nin), W (nin,nout). */
DCV_API int dcv_dense(dcv_ctx* ctx, const float* x, const float* w, const float* bias, int B, int nin, int nout,
                      int act, float* out);
/* DDense.down (deepdream.py:311-321): out = x W^T, x (B,nout) -> (B,nin); zero bias, no activation
 * (deepdream.py:289-297). */
DCV_API int dcv_dense_transpose(dcv_ctx* ctx, const float* x, const float* w, int B, int nin, int nout, float* out);

/* ---- DeepDream (BASELINE.json north_star; NO counterpart in the reference: app/deepdream.py holds only
 * deconvnet code).  Algorithm = the derived spec of SURVEY.md Appendix A, restated in
 * csrc/dream.cu.  The network must consist of conv/pool layers (include_top = False), any image size.
 *   loss = sum_k coeffs[k] * mean(act_{loss_layers[k]}^2);   per iteration: x = roll(img, (oy,ox)),
 *   g = dloss/dx, img += step * unroll(g) / max(rms(g), 1e-7)  (rms per image).
 * shapes: n_oct (h,w) pairs, smallest first, the last one equal to (H,W); jitter: n_oct*iterations (oy,ox)
 * pairs drawn by the caller (so that an oracle can use the same stream); losses (optional, host):
 * n_oct*iterations loss values BEFORE each update (costs a device sync per iteration).
 * precision: DCV_FP32 = fp32-accurate forward and backward (split-operand tcgen05 kernels; FFMA with DCV_FP32_TC=0);
 * DCV_BF16 = forward AND backward on tcgen05 with a bf16 tape in the padded-flat layout (layer 0, K = 27, runs as
 * im2col rows + one K = 64 GEMM); the image and its update stay fp32. ---- */
DCV_API int dcv_dream(dcv_ctx* ctx, const float* img, int B, int H, int W, const int* loss_layers, const float* coeffs, int nl,
                      const int* shapes, int n_oct, int iterations, float step, const int* jitter, int precision, float* out,
                      double* losses);
/* Same on device buffers, asynchronous on the context's stream. */
DCV_API int dcv_dream_dev(dcv_ctx* ctx, const float* img_dev, int B, int H, int W, const int* loss_layers, const float* coeffs,
                          int nl, const int* shapes, int n_oct, int iterations, float step, const int* jitter, int precision,
                          float* out_dev);
/* One forward + backward: grad_out (B,H,W,C0) = dloss/dimg (un-normalised), loss_out (optional) = loss. */
DCV_API int dcv_dream_gradient(dcv_ctx* ctx, const float* img, int B, int H, int W, const int* loss_layers, const float* coeffs,
                               int nl, int precision, float* grad_out, double* loss_out);
/* Bilinear resize NHWC, half-pixel centres (align_corners = False), no antialiasing: the octave resize. */
DCV_API int dcv_resize_bilinear(dcv_ctx* ctx, const float* x, int B, int H, int W, int C, int OH, int OW, float* out);
/* numpy.roll over the two spatial axes by (oy, ox): the jitter. */
DCV_API int dcv_roll(dcv_ctx* ctx, const float* x, int B, int H, int W, int C, int oy, int ox, float* out);

/* ---- device-resident forms of the streaming operators above: x_dev / out_dev / sw_dev are device pointers owned
 * by the caller (16-byte aligned for pool / unpool), nothing is copied or allocated, and the call is asynchronous
 * on the context's stream (dcv_sync to wait).  Same arithmetic as the host-buffer forms; for pipelines that keep
 * their images on the GPU between DPooling.up / DPooling.down (deepdream.py:129-150) or DeepDream steps.
 * dcv_resize_bilinear_dev moves its source rows with 16-byte bulk copies: it may read the (at most 12) bytes between
 * a row segment and the 16-byte boundaries around it, so x_dev must lie in an allocation that starts and ends on
 * 16-byte boundaries (true of every cudaMalloc / caching-allocator block); x_dev itself needs only 4-byte alignment. --- */
DCV_API int dcv_maxpool_switch_dev(dcv_ctx* ctx, const float* x_dev, int B, int H, int W, int C, float* pooled_dev, uint8_t* sw_dev);
DCV_API int dcv_unpool_switch_dev(dcv_ctx* ctx, const float* x_dev, const uint8_t* sw_dev, int B, int H, int W, int C, float* out_dev);
DCV_API int dcv_resize_bilinear_dev(dcv_ctx* ctx, const float* x_dev, int B, int H, int W, int C, int OH, int OW, float* out_dev);
DCV_API int dcv_roll_dev(dcv_ctx* ctx, const float* x_dev, int B, int H, int W, int C, int oy, int ox, float* out_dev);

#ifdef __cplusplus
}
#endif
#endif /* DCV_H_ */
