// Internal host-side types of the dcv library.
#pragma once
#include "common.cuh"

enum LayerKind { LK_CONV = 0, LK_POOL = 1, LK_FLATTEN = 2, LK_DENSE = 3 };

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(dcv_ctx* ctx, size_t bytes);   // grow-only; contents are NOT preserved on growth
  void release();
};

// tensor-core side of a conv layer (conv_tc.cu)
struct TcWeights {
  __nv_bfloat16* wd_bf16 = nullptr;   // backward weights packed [9][n_pad][k_pad] bf16 (K-major rows)
  int n_pad = 0, k_pad = 0;           // N = cin, K = cout of the forward layer, rounded up to 64
  __nv_bfloat16* wf_bf16 = nullptr;   // forward weights [9][fn_pad][fk_pad] (N = cout, K = cin); built on first bf16 DeepDream
  int fn_pad = 0, fk_pad = 0;
  __nv_bfloat16* wf0_bf16 = nullptr;  // first layer (cin <= 4) as a GEMM over im2col rows: [n_pad][64], taps x channels along K, twice (for the hi / lo planes of the input)
  __nv_bfloat16* wall3_bf16 = nullptr; // all-tap weights of a narrow last layer split into three planes: [3][32][k_pad] (exact GEMM + col2im)
  __nv_bfloat16* wd3_bf16 = nullptr;  // backward (flipped, transposed) weights split the same way: [9][3][n_pad][k_pad] (exact projection)
  __nv_bfloat16* wf3_bf16 = nullptr;  // forward weights split into (hi, mid, lo) bf16 planes: [9][3][fn_pad][fk_pad] (exact forward)
  __nv_bfloat16* wall_bf16 = nullptr; // narrow last layer (9*cin <= 32): all taps stacked along N, [32][k_pad]
  // fp16 two-plane scheme of the DCV_FP32 backward convs (tc_exact.cuh, SCHEME 1); 16-bit storage, fp16 contents
  __nv_bfloat16* wd2_f16 = nullptr;   // backward weights scaled by a power of two and split into (hi, lo): [9][2][n_pad][k_pad]
  __nv_bfloat16* wall2_f16 = nullptr; // the same for the all-tap weights of a narrow last layer: [2][32][k_pad]
  uint32_t* wd_amax = nullptr;        // float bit pattern of the largest |weight| (device scalar)
};

struct Layer {
  int kind = LK_CONV;
  // conv: cin/cout; dense: nin/nout
  int cin = 0, cout = 0, relu = 0, act_kind = 0;
  float* w = nullptr;        // conv: HWIO == [9][cin][cout]; dense: (nin, nout)
  float* bias = nullptr;
  float* wd = nullptr;       // conv backward kernel [9][cout][cin] (deepdream.py:80-81)
  TcWeights tc;
  // tape (valid for layers <= tape_upto)
  int B = 0, H = 0, W = 0, C = 0;      // output shape
  int in_H = 0, in_W = 0, in_C = 0;    // pool / flatten: input shape
  DevBuf act;                          // output, fp32
  DevBuf sw;                           // pool: switches, 1 byte per pooled element
  DevBuf act_h;                        // bf16 padded-flat copy of the output (bf16 DeepDream tape)
};

struct ProfRec {
  cudaEvent_t a, b;
  int cat;
  double flops, bytes;
};

struct dcv_ctx {
  int device = 0;
  int num_sms = dcv::kNumSMs;
  cudaStream_t stream = nullptr;
  std::string err;
  long long launches = 0;
  size_t workspace_budget = 0;
  std::vector<Layer> net;
  int tape_upto = -1;
  int latency_tiles = 0;   // 1 while DeepDream runs: N tiles chosen by the wave/cycle model (conv_tc.cu)
  int pdl_hold = 0;        // 1: a kernel that writes weights was queued; the next launch through launch_kc is a plain one
  int fwd_mode = 1;    // dcv_set_forward_mode: 0 = FFMA fp32, 1 = tcgen05 with split operands (default)
  int fp32_tc = 1;     // dcv_set_fp32_mode: backward convs of the DCV_FP32 chain, 0 = FFMA, 1 = split-operand tcgen05 (DCV_FP32_TC)
  int in_B = 0, in_H = 0, in_W = 0, in_C = 0;
  DevBuf input;
  DevBuf ws[6];        // projection workspace: [0],[1] ping-pong tensors, [2],[3] maps, [4],[5] bf16 ping-pong
  DevBuf feat_dev;
  DevBuf scratch;      // reduction partials
  DevBuf amax;         // fp16 scheme: per-item float bit patterns: two arrays of largest |activation| (read / filled), two of fused-split scale sources
  int amax_cur = -1;   // which of the two holds the maxima of the last exact conv's output (-1: none)
  int split_ready = -1; // 0 / 1: the last exact conv wrote its result as the next conv's planes into ws[4 + split_ready]
  int scale_cur = 0;    // which of the two scale-source arrays those planes were scaled by
  DevBuf stage;        // staging for host-buffer entry points
  // host-output projections: chunks are copied back on a second stream while the next chunk computes
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_ready[2] = {nullptr, nullptr}, ev_copied[2] = {nullptr, nullptr};
  DevBuf out_stage[2];
  DevBuf dream[6];     // DeepDream: image, original, shrunk original, rolled input / resize temp, 2 gradient buffers
  void* dream_graphs = nullptr;   // cached CUDA graphs of DeepDream iterations (dream.cu)
  DevBuf jit_dev;      // DeepDream: {iteration counter, (oy, ox) per iteration}
  unsigned long long alloc_epoch = 0;   // bumped whenever a DevBuf is (re)allocated: invalidates captured graphs
  // optional per-kernel-class timing (dcv_profile): CUDA events around every launch of a class
  bool prof_on = false;
  std::vector<ProfRec> prof;
  std::vector<cudaEvent_t> ev_pool;
  void prof_begin(int cat, double flops, double bytes);
  void prof_end();
  int fail(int code, const char* fmt, ...);
};

struct ProfScope {
  dcv_ctx* c;
  bool on;
  // enabled = false: a caller up the stack already has a scope open around this launch
  ProfScope(dcv_ctx* ctx, int cat, double flops, double bytes, bool enabled = true) : c(ctx), on(enabled && ctx->prof_on) {
    if (on) c->prof_begin(cat, flops, bytes);
  }
  ~ProfScope() { if (on) c->prof_end(); }
};

// api.cu
int forward_impl(dcv_ctx* ctx, const float* img, bool on_device, int B, int H, int W, int upto);

// dream.cu
void dream_release_graphs(dcv_ctx* ctx);

// conv_tc.cu — bf16 tensor-core chain in the padded-flat layout (see the header comment there)
int tc_prepare_layer(dcv_ctx* ctx, Layer& l);
void tc_free_layer(Layer& l);
int tc_cpad(int c);
size_t tc_padded_elems(int h, int w, int c);         // bf16 elements of one padded item
int tc_pad_convert(dcv_ctx* ctx, const float* x, int P, int h, int w, int C, __nv_bfloat16* out);
int tc_rank1(dcv_ctx* ctx, const float* maps, const Layer& l, const int* feat, int n, int B, int h, int w, __nv_bfloat16* out);
// (ph,pw) pooled size, (h,w) unpooled size with h in {2ph, 2ph+1}
// DeepDream gradient prep of the layer the result flows into, folded into the producing kernel:
//   out = mask(act > 0 if relu) * (value + scale * act), act in the output's padded-flat layout
struct TcPrep { const __nv_bfloat16* act; float scale; int relu; };
int tc_unpool(dcv_ctx* ctx, const __nv_bfloat16* x, const uint8_t* sw, int P, int Bsw, int ph, int pw, int h, int w, int C,
              __nv_bfloat16* out, const TcPrep* prep = nullptr);
int tc_pool(dcv_ctx* ctx, const __nv_bfloat16* x, int P, int h, int w, int C, __nv_bfloat16* out, uint8_t* sw);
int tc_dream_prep(dcv_ctx* ctx, __nv_bfloat16* g, const __nv_bfloat16* act, size_t elems, float scale, int has_g, int relu);
int tc_sumsq_partials(size_t elems);
int tc_sumsq(dcv_ctx* ctx, const __nv_bfloat16* x, size_t elems, double* partial, int np);
int tc_prepare_forward(dcv_ctx* ctx, Layer& l);
// first conv layer (cin <= 4) of the bf16 DeepDream forward: fp32 image (P,h,w,cin) -> bf16 padded-flat activations;
// im2col rows (9*cin values as bf16 hi + lo planes, K = 64) + one un-shifted GEMM with bias / ReLU epilogue
bool tc_first_layer_ok(const Layer& l);
int tc_conv_first(dcv_ctx* ctx, Layer& l, const float* x, int P, int h, int w, __nv_bfloat16* out_bf16,
                  const int* jit = nullptr);      // jit: device jitter table, the image is read rolled (DeepDream)
// fp32-accurate forward conv on tcgen05 (bf16 x 3 operand split, six products); fp32 compact in and out
int tc_prepare_forward_exact(dcv_ctx* ctx, Layer& l);
int tc_conv_fwd_exact(dcv_ctx* ctx, const Layer& l, const float* in, int P, int h, int w, float* out, int planes_in = -1,
                      const Layer* fuse_next = nullptr, int* planes_out = nullptr);
// the same for the backward (deconvnet) conv of the fp32 projection mode: in fp32 compact (P,h,w,cout) ->
// out fp32 compact (P,h,w,cin) = relu?(conv(in, flipT(W))); scratch ctx->ws[4] sized by tc_split_elems()
int tc_prepare_backward_exact(dcv_ctx* ctx, Layer& l);
// chain: `in` is the output of the previous tc_conv_bwd_exact call on this context (possibly unpooled / rectified since):
// its per-item maxima are already known
// fuse_next (two-plane scheme): the layer whose backward conv consumes this result next, with nothing in between; the
// result is then written as THAT conv's fp16 operand planes (no fp32 copy in `out`, no split pass in the next call,
// which must be the next tc_conv_bwd_exact on this context, with chain = true)
int tc_conv_bwd_exact(dcv_ctx* ctx, const Layer& l, const float* in, int P, int h, int w, float* out, int relu, bool chain = false,
                      const Layer* fuse_next = nullptr);
bool tc_bwd_exact_can_fuse_split();
// the same with DPooling.down in front: `pooled` is (P,h/2,w/2,cout), unpooled by the switches inside the split pass
bool tc_bwd_exact_fuses_unpool();
int tc_conv_bwd_exact_unpool(dcv_ctx* ctx, const Layer& l, const float* pooled, const uint8_t* sw, int Bsw, int P, int h, int w,
                             float* out, int relu, bool chain, const Layer* fuse_next = nullptr);
size_t tc_split_elems(int h, int w, int c);          // bf16 elements of one item split into three planes
// backward conv; exactly one of out_bf16 (padded-flat, pitch l.tc.n_pad) / out_f32 (compact (P,h,w,l.cin)) is
// non-null; relu < 0: the layer's own activation (deconvnet rule), 0/1: explicit (DeepDream uses 0)
// up (optional, bf16 output only): fuse the switch-driven unpool of the pool layer below into the epilogue; out_bf16 is
// then the padded-flat tensor at the unpooled size (up->h, up->w) and the pooled tensor is never written
struct TcUnpool { const uint8_t* sw; int B, C, h, w; };
int tc_conv(dcv_ctx* ctx, const Layer& l, const __nv_bfloat16* in, int P, int h, int w, __nv_bfloat16* out_bf16, float* out_f32,
            int relu, const TcUnpool* up = nullptr, const TcPrep* prep = nullptr);
// forward conv + bias + activation, bf16 padded-flat in and out
int tc_conv_fwd(dcv_ctx* ctx, const Layer& l, const __nv_bfloat16* in, int P, int h, int w, __nv_bfloat16* out_bf16);
int tc_deconv(dcv_ctx* ctx, const Layer& l, const float* in, int P, int h, int w, float* out);
// strip kernel (tc_strip.cuh): 64 -> 64 backward conv with optional fused unpool (sw != null: `in` is the pooled tensor)
// and optional fused last conv (last != null: fp32 compact result of the whole tail)
bool tc_strip_conv_ok(const Layer& l);
bool tc_strip_last_ok(const Layer& conv, const Layer& last);
int tc_strip(dcv_ctx* ctx, const Layer& conv, const Layer* last, const __nv_bfloat16* in, const uint8_t* sw, int Bsw, int P, int h,
             int w, __nv_bfloat16* out_bf16, float* out_f32, int relu);
