// C-ABI of the dcv library (include/dcv.h): context, network description, forward tape,
// projection chain and the standalone operators.  Host-side orchestration only; kernels live in
// kernels_stream.cu, conv_fp32.cu and conv_tc.cu.
#include "common.cuh"
#include "internal.cuh"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

using namespace dcv;

namespace {
thread_local std::string g_create_error;
}

// ---- error helpers ------------------------------------------------------------------------------
int dcv_ctx::fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  err = buf;
  return code;
}

#define CU(ctx, expr)                                                                              \
  do {                                                                                             \
    cudaError_t e_ = (expr);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return (ctx)->fail(e_ == cudaErrorMemoryAllocation ? DCV_ERR_NOMEM : DCV_ERR_CUDA, "%s: %s", #expr, \
                         cudaGetErrorString(e_));                                                  \
  } while (0)

#define CHECK_LAUNCH(ctx) CU(ctx, cudaGetLastError())

int DevBuf::ensure(dcv_ctx* ctx, size_t bytes) {
  if (bytes <= cap) return 0;
  if (p) { cudaFree(p); p = nullptr; cap = 0; }
  size_t want = bytes + bytes / 8;       // slack so a slightly larger request does not reallocate
  cudaError_t e = cudaMalloc(&p, want);
  if (e != cudaSuccess) {
    p = nullptr;
    return ctx->fail(DCV_ERR_NOMEM, "cudaMalloc(%zu bytes): %s", want, cudaGetErrorString(e));
  }
  cap = want;
  ctx->alloc_epoch++;
  return 0;
}

void DevBuf::release() {
  if (p) cudaFree(p);
  p = nullptr;
  cap = 0;
}

static size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }

static int set_device(dcv_ctx* ctx) {
  CU(ctx, cudaSetDevice(ctx->device));
  return 0;
}

#define ENTER(ctx)                                   \
  if (!(ctx)) return DCV_ERR_ARG;                    \
  { int r_ = set_device(ctx); if (r_) return r_; }

#define TRY(expr) do { int r_ = (expr); if (r_) return r_; } while (0)

// ---- life cycle ---------------------------------------------------------------------------------
extern "C" int dcv_version(void) { return 100; }

extern "C" int dcv_create(int device, dcv_ctx** out) {
  if (!out) return DCV_ERR_ARG;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e);
    return DCV_ERR_CUDA;
  }
  if (device < 0 || device >= ndev) { g_create_error = "device index out of range"; return DCV_ERR_ARG; }
  e = cudaSetDevice(device);
  if (e != cudaSuccess) { g_create_error = cudaGetErrorString(e); return DCV_ERR_CUDA; }
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) { g_create_error = cudaGetErrorString(e); return DCV_ERR_CUDA; }
  if (prop.major != 10) {
    g_create_error = "this library is built for sm_100a (B200) only; device is sm_" + std::to_string(prop.major) +
                     std::to_string(prop.minor);
    return DCV_ERR_UNSUPPORTED;
  }
  dcv_ctx* ctx = new (std::nothrow) dcv_ctx();
  if (!ctx) return DCV_ERR_NOMEM;
  ctx->device = device;
  ctx->num_sms = prop.multiProcessorCount;
  e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { g_create_error = cudaGetErrorString(e); delete ctx; return DCV_ERR_CUDA; }
  e = cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking);
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
    e = cudaEventCreateWithFlags(&ctx->ev_ready[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_copied[i], cudaEventDisableTiming);
  }
  if (e != cudaSuccess) { g_create_error = cudaGetErrorString(e); delete ctx; return DCV_ERR_CUDA; }
  if (const char* t = getenv("DCV_FP32_TC")) ctx->fp32_tc = atoi(t) != 0;
  const char* ws = getenv("DCV_WORKSPACE_MB");
  ctx->workspace_budget = (ws ? (size_t)atoll(ws) : (size_t)12288) << 20;
  *out = ctx;
  return DCV_OK;
}

static void free_layer(Layer& l) {
  if (l.w) cudaFree(l.w);
  if (l.bias) cudaFree(l.bias);
  if (l.wd) cudaFree(l.wd);
  tc_free_layer(l);
  l.act.release();
  l.sw.release();
  l.act_h.release();
}

extern "C" void dcv_destroy(dcv_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  dream_release_graphs(ctx);
  for (auto& l : ctx->net) free_layer(l);
  ctx->input.release();
  for (auto& b : ctx->ws) b.release();
  ctx->feat_dev.release();
  ctx->scratch.release();
  ctx->amax.release();
  ctx->stage.release();
  for (auto& b : ctx->dream) b.release();
  ctx->jit_dev.release();
  for (auto& b : ctx->out_stage) b.release();
  cudaStreamSynchronize(ctx->copy_stream);
  for (int i = 0; i < 2; ++i) { cudaEventDestroy(ctx->ev_ready[i]); cudaEventDestroy(ctx->ev_copied[i]); }
  cudaStreamDestroy(ctx->copy_stream);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
}

extern "C" const char* dcv_last_error(const dcv_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }
extern "C" int dcv_device(const dcv_ctx* ctx) { return ctx ? ctx->device : DCV_ERR_ARG; }
extern "C" long long dcv_launch_count(const dcv_ctx* ctx) { return ctx ? ctx->launches : 0; }
extern "C" void* dcv_stream(const dcv_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

extern "C" int dcv_sync(dcv_ctx* ctx) {
  ENTER(ctx);
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

// ---- page-locked host memory for result buffers (device-to-host copies into pageable memory are staged by the driver
// at a fraction of the PCIe rate) ---------------------------------------------------------------------
extern "C" int dcv_host_alloc(dcv_ctx* ctx, size_t bytes, void** out) {
  ENTER(ctx);
  if (!out) return ctx->fail(DCV_ERR_ARG, "dcv_host_alloc: null output");
  *out = nullptr;
  cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault);
  if (e != cudaSuccess) { *out = nullptr; return ctx->fail(DCV_ERR_NOMEM, "cudaHostAlloc(%zu bytes): %s", bytes, cudaGetErrorString(e)); }
  return 0;
}

extern "C" int dcv_host_free(dcv_ctx* ctx, void* p) {
  ENTER(ctx);
  if (p) CU(ctx, cudaFreeHost(p));
  return 0;
}

// ---- per-class kernel timing -------------------------------------------------------------------------
void dcv_ctx::prof_begin(int cat, double flops, double bytes) {
  ProfRec r;
  r.cat = cat; r.flops = flops; r.bytes = bytes;
  for (cudaEvent_t* e : {&r.a, &r.b}) {
    if (!ev_pool.empty()) { *e = ev_pool.back(); ev_pool.pop_back(); }
    else cudaEventCreate(e);
  }
  cudaEventRecord(r.a, stream);
  prof.push_back(r);
}

void dcv_ctx::prof_end() { cudaEventRecord(prof.back().b, stream); }

extern "C" int dcv_profile(dcv_ctx* ctx, int enable) {
  ENTER(ctx);
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  for (auto& r : ctx->prof) { ctx->ev_pool.push_back(r.a); ctx->ev_pool.push_back(r.b); }
  ctx->prof.clear();
  ctx->prof_on = enable != 0;
  return 0;
}

extern "C" int dcv_profile_read(dcv_ctx* ctx, int cat, double* ms, long long* launches, double* flops, double* bytes) {
  ENTER(ctx);
  if (cat < 0 || cat >= DCV_K_NUM) return ctx->fail(DCV_ERR_ARG, "dcv_profile_read: unknown kernel class %d", cat);
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  double t = 0, f = 0, by = 0;
  long long n = 0;
  for (auto& r : ctx->prof) {
    if (r.cat != cat) continue;
    float e = 0.f;
    CU(ctx, cudaEventElapsedTime(&e, r.a, r.b));
    t += e; f += r.flops; by += r.bytes; ++n;
  }
  if (ms) *ms = t;
  if (launches) *launches = n;
  if (flops) *flops = f;
  if (bytes) *bytes = by;
  return 0;
}

extern "C" int dcv_profile_launch(dcv_ctx* ctx, int index, int* cat, double* ms, double* flops, double* bytes) {
  ENTER(ctx);
  if (index < 0 || index >= (int)ctx->prof.size()) return DCV_ERR_ARG;      // end of the list (no error text)
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  const ProfRec& r = ctx->prof[index];
  float e = 0.f;
  CU(ctx, cudaEventElapsedTime(&e, r.a, r.b));
  if (cat) *cat = r.cat;
  if (ms) *ms = e;
  if (flops) *flops = r.flops;
  if (bytes) *bytes = r.bytes;
  return 0;
}

// ---- network description --------------------------------------------------------------------------
extern "C" int dcv_net_clear(dcv_ctx* ctx) {
  ENTER(ctx);
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  // captured DeepDream graphs hold raw pointers into the tape and weight buffers freed below
  dream_release_graphs(ctx);
  ctx->alloc_epoch++;
  for (auto& l : ctx->net) free_layer(l);
  ctx->net.clear();
  ctx->tape_upto = -1;
  return 0;
}

static int upload(dcv_ctx* ctx, float** dst, const float* src, size_t n) {
  CU(ctx, cudaMalloc(dst, std::max<size_t>(n, 1) * sizeof(float)));
  if (src) CU(ctx, cudaMemcpyAsync(*dst, src, n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
  else CU(ctx, cudaMemsetAsync(*dst, 0, n * sizeof(float), ctx->stream));
  return 0;
}

extern "C" int dcv_net_add_conv(dcv_ctx* ctx, const float* w_hwio, const float* bias, int cin, int cout, int relu) {
  ENTER(ctx);
  if (!w_hwio || cin <= 0 || cout <= 0) return ctx->fail(DCV_ERR_ARG, "dcv_net_add_conv: bad arguments");
  Layer l;
  l.kind = LK_CONV; l.cin = cin; l.cout = cout; l.relu = relu ? 1 : 0;
  const size_t nw = (size_t)9 * cin * cout;
  TRY(upload(ctx, &l.w, w_hwio, nw));
  TRY(upload(ctx, &l.bias, bias, cout));
  CU(ctx, cudaMalloc(&l.wd, nw * sizeof(float)));
  launch_flip_transpose(l.w, cin, cout, l.wd, ctx->stream); ctx->launches++;
  CHECK_LAUNCH(ctx);
  TRY(tc_prepare_layer(ctx, l));
  CU(ctx, cudaStreamSynchronize(ctx->stream));      // host pointers are borrowed for the call only
  ctx->net.push_back(l);
  return (int)ctx->net.size() - 1;
}

extern "C" int dcv_net_add_pool(dcv_ctx* ctx) {
  ENTER(ctx);
  Layer l;
  l.kind = LK_POOL;
  ctx->net.push_back(l);
  return (int)ctx->net.size() - 1;
}

extern "C" int dcv_net_add_flatten(dcv_ctx* ctx) {
  ENTER(ctx);
  Layer l;
  l.kind = LK_FLATTEN;
  ctx->net.push_back(l);
  return (int)ctx->net.size() - 1;
}

extern "C" int dcv_net_add_dense(dcv_ctx* ctx, const float* w, const float* bias, int nin, int nout, int act) {
  ENTER(ctx);
  if (!w || nin <= 0 || nout <= 0 || act < 0 || act > 2) return ctx->fail(DCV_ERR_ARG, "dcv_net_add_dense: bad arguments");
  Layer l;
  l.kind = LK_DENSE; l.cin = nin; l.cout = nout; l.act_kind = act;
  TRY(upload(ctx, &l.w, w, (size_t)nin * nout));
  TRY(upload(ctx, &l.bias, bias, nout));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->net.push_back(l);
  return (int)ctx->net.size() - 1;
}

extern "C" int dcv_net_num_layers(const dcv_ctx* ctx) { return ctx ? (int)ctx->net.size() : DCV_ERR_ARG; }

// ---- forward tape -----------------------------------------------------------------------------------
int forward_impl(dcv_ctx* ctx, const float* img, bool on_device, int B, int H, int W, int upto) {
  if (!img || B <= 0 || H <= 0 || W <= 0) return ctx->fail(DCV_ERR_ARG, "dcv_forward: bad arguments");
  if (ctx->net.empty()) return ctx->fail(DCV_ERR_STATE, "dcv_forward: empty network");
  if (upto < 0 || upto >= (int)ctx->net.size()) return ctx->fail(DCV_ERR_ARG, "dcv_forward: layer %d out of range", upto);
  int C = -1;
  for (auto& l : ctx->net) if (l.kind == LK_CONV) { C = l.cin; break; }
  if (C < 0) return ctx->fail(DCV_ERR_UNSUPPORTED, "dcv_forward: network has no conv layer to define the input channels");
  ctx->tape_upto = -1;
  const size_t n_in = (size_t)B * H * W * C;
  TRY(ctx->input.ensure(ctx, n_in * sizeof(float)));
  CU(ctx, cudaMemcpyAsync(ctx->input.p, img, n_in * sizeof(float),
                          on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, ctx->stream));
  if (!on_device) CU(ctx, cudaEventRecord(ctx->ev_ready[0], ctx->stream));    // the host image may be released after this point
  ctx->in_B = B; ctx->in_H = H; ctx->in_W = W; ctx->in_C = C;
  const float* cur = (const float*)ctx->input.p;
  int cb = B, ch = H, cw = W, cc = C;
  bool flat = false;
  int fwd_planes = -1;               // workspace index holding the current tensor as split operand planes (-1: none)
  for (int j = 0; j <= upto; ++j) {
    Layer& l = ctx->net[j];
    if (l.kind != LK_CONV) fwd_planes = -1;
    switch (l.kind) {
      case LK_CONV: {
        if (flat) return ctx->fail(DCV_ERR_UNSUPPORTED, "conv after flatten");
        if (l.cin != cc) return ctx->fail(DCV_ERR_ARG, "layer %d: expects %d input channels, got %d", j, l.cin, cc);
        TRY(l.act.ensure(ctx, (size_t)cb * ch * cw * l.cout * sizeof(float)));
        // fp32-accurate forward: tcgen05 with split operands (conv_tc.cu) for every layer but the first (K = 27),
        // FFMA otherwise / when disabled (DCV_FWD_TC=0)
        if (ctx->fwd_mode == 1 && l.cin >= 16) {
          TRY(tc_prepare_forward_exact(ctx, l));
          // a conv directly above (no pool in between) takes this result as ready-made operand planes
          const Layer* nx = (j + 1 <= upto && ctx->net[j + 1].kind == LK_CONV && ctx->net[j + 1].cin >= 16) ? &ctx->net[j + 1] : nullptr;
          int planes_out = -1;
          TRY(tc_conv_fwd_exact(ctx, l, cur, cb, ch, cw, (float*)l.act.p, fwd_planes, nx, &planes_out));
          fwd_planes = planes_out;
          cc = l.cout;
          break;
        }
        ConvEpilogue e; e.bias = l.bias; e.relu = l.relu;
        {
          ProfScope ps(ctx, DCV_K_CONV_FP32, 2.0 * 9 * cb * ch * cw * (double)l.cin * l.cout,
                       4.0 * cb * ch * cw * ((double)l.cin + l.cout));
          launch_conv3x3_fp32(cur, l.w, cb, ch, cw, l.cin, l.cout, e, (float*)l.act.p, ctx->stream); ctx->launches++;
        }
        fwd_planes = -1;
        cc = l.cout;
        break;
      }
      case LK_POOL: {
        if (flat) return ctx->fail(DCV_ERR_UNSUPPORTED, "pool after flatten");
        const int oh = ch / 2, ow = cw / 2;
        if (oh == 0 || ow == 0) return ctx->fail(DCV_ERR_ARG, "layer %d: input %dx%d too small to pool", j, ch, cw);
        TRY(l.act.ensure(ctx, (size_t)cb * oh * ow * cc * sizeof(float)));
        TRY(l.sw.ensure(ctx, (size_t)cb * oh * ow * cc));
        {
          // algorithmic bytes: 4*HWC read + 5*HWC/4 written
          ProfScope ps(ctx, DCV_K_POOL, 0.0, (double)cb * ch * cw * cc * (4.0 + 5.0 / 4.0));
          launch_maxpool_switch(cur, cb, ch, cw, cc, (float*)l.act.p, (uint8_t*)l.sw.p, ctx->stream); ctx->launches++;
        }
        l.in_H = ch; l.in_W = cw;
        ch = oh; cw = ow;
        break;
      }
      case LK_FLATTEN: {
        // NHWC row-major flatten is a no-op on the data (DFlatten.up, deepdream.py:342-352)
        l.in_H = ch; l.in_W = cw; l.in_C = cc;
        TRY(l.act.ensure(ctx, (size_t)cb * ch * cw * cc * sizeof(float)));
        CU(ctx, cudaMemcpyAsync(l.act.p, cur, (size_t)cb * ch * cw * cc * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
        cc = ch * cw * cc; ch = 1; cw = 1;
        flat = true;
        break;
      }
      case LK_DENSE: {
        if (!flat && !(ch == 1 && cw == 1)) return ctx->fail(DCV_ERR_UNSUPPORTED, "dense needs a flattened input");
        if (l.cin != cc) return ctx->fail(DCV_ERR_ARG, "layer %d: dense expects %d inputs, got %d", j, l.cin, cc);
        TRY(l.act.ensure(ctx, (size_t)cb * l.cout * sizeof(float)));
        TRY(ctx->scratch.ensure(ctx, (size_t)dense_splits(l.cin) * cb * l.cout * sizeof(float)));
        launch_dense_fwd(cur, l.w, l.bias, cb, l.cin, l.cout, l.act_kind, (float*)ctx->scratch.p, (float*)l.act.p, ctx->stream);
        ctx->launches += 2;
        if (l.act_kind == DCV_ACT_SOFTMAX) {
          launch_softmax_rows((float*)l.act.p, cb, l.cout, (float*)l.act.p, ctx->stream); ctx->launches++;
        }
        cc = l.cout;
        break;
      }
    }
    l.B = cb; l.H = ch; l.W = cw; l.C = cc;
    cur = (const float*)l.act.p;
  }
  CHECK_LAUNCH(ctx);
  ctx->tape_upto = upto;
  return 0;
}

extern "C" int dcv_set_forward_mode(dcv_ctx* ctx, int mode) {
  ENTER(ctx);
  if (mode != 0 && mode != 1) return ctx->fail(DCV_ERR_ARG, "dcv_set_forward_mode: mode must be 0 (FFMA) or 1 (tensor cores)");
  ctx->fwd_mode = mode;
  return 0;
}

extern "C" int dcv_set_fp32_mode(dcv_ctx* ctx, int mode) {
  ENTER(ctx);
  if (mode != 0 && mode != 1) return ctx->fail(DCV_ERR_ARG, "dcv_set_fp32_mode: mode must be 0 (FFMA) or 1 (tensor cores)");
  if (ctx->fp32_tc != mode) {
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    dream_release_graphs(ctx);          // captured fp32 iterations hold the other kernels
    ctx->fp32_tc = mode;
  }
  return 0;
}

extern "C" int dcv_forward(dcv_ctx* ctx, const float* img, int B, int H, int W, int upto) {
  ENTER(ctx);
  TRY(forward_impl(ctx, img, false, B, H, W, upto));
  // The borrowed host image is only needed until its copy has finished; the tape's kernels stay queued on the context's
  // stream, where every later call (dcv_project, dcv_read_activation, ...) is ordered behind them.  Draining the stream
  // here cost the end-to-end step ~0.5 ms of idle GPU per request.
  CU(ctx, cudaEventSynchronize(ctx->ev_ready[0]));
  return 0;
}

extern "C" int dcv_forward_dev(dcv_ctx* ctx, const float* img, int B, int H, int W, int upto) {
  ENTER(ctx);
  return forward_impl(ctx, img, true, B, H, W, upto);
}

static int check_tape(const dcv_ctx* c, dcv_ctx* ctx, int layer, const char* who) {
  (void)c;
  if (layer < 0 || layer >= (int)ctx->net.size()) return ctx->fail(DCV_ERR_ARG, "%s: layer %d out of range", who, layer);
  if (layer > ctx->tape_upto) return ctx->fail(DCV_ERR_STATE, "%s: layer %d is not in the current tape (run dcv_forward first)", who, layer);
  return 0;
}

extern "C" int dcv_layer_shape(const dcv_ctx* ctx, int layer, int dims[4]) {
  if (!ctx || !dims) return DCV_ERR_ARG;
  dcv_ctx* m = const_cast<dcv_ctx*>(ctx);
  TRY(check_tape(ctx, m, layer, "dcv_layer_shape"));
  const Layer& l = ctx->net[layer];
  dims[0] = l.B; dims[1] = l.H; dims[2] = l.W; dims[3] = l.C;
  return 0;
}

extern "C" int dcv_read_activation(dcv_ctx* ctx, int layer, float* out) {
  ENTER(ctx);
  if (!out) return ctx->fail(DCV_ERR_ARG, "dcv_read_activation: null output");
  TRY(check_tape(ctx, ctx, layer, "dcv_read_activation"));
  const Layer& l = ctx->net[layer];
  CU(ctx, cudaMemcpyAsync(out, l.act.p, (size_t)l.B * l.H * l.W * l.C * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_read_switch(dcv_ctx* ctx, int layer, uint8_t* out) {
  ENTER(ctx);
  if (!out) return ctx->fail(DCV_ERR_ARG, "dcv_read_switch: null output");
  TRY(check_tape(ctx, ctx, layer, "dcv_read_switch"));
  const Layer& l = ctx->net[layer];
  if (l.kind != LK_POOL) return ctx->fail(DCV_ERR_ARG, "dcv_read_switch: layer %d is not a pool layer", layer);
  CU(ctx, cudaMemcpyAsync(out, l.sw.p, (size_t)l.B * l.H * l.W * l.C, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_channel_sums(dcv_ctx* ctx, int layer, double* out) {
  ENTER(ctx);
  if (!out) return ctx->fail(DCV_ERR_ARG, "dcv_channel_sums: null output");
  TRY(check_tape(ctx, ctx, layer, "dcv_channel_sums"));
  const Layer& l = ctx->net[layer];
  const size_t rows = (size_t)l.B * l.H * l.W;
  const int np = channel_sums_partials(rows, l.C);
  TRY(ctx->scratch.ensure(ctx, ((size_t)np + 1) * l.C * sizeof(double)));
  double* partial = (double*)ctx->scratch.p;
  double* res = partial + (size_t)np * l.C;
  launch_channel_sums((const float*)l.act.p, rows, l.C, partial, np, res, ctx->stream); ctx->launches += 2;
  CHECK_LAUNCH(ctx);
  CU(ctx, cudaMemcpyAsync(out, res, (size_t)l.C * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_channel_sums_buf(dcv_ctx* ctx, const float* x, size_t rows, int C, double* out) {
  ENTER(ctx);
  if (!x || !out || C <= 0) return ctx->fail(DCV_ERR_ARG, "dcv_channel_sums_buf: bad arguments");
  if (rows == 0) { memset(out, 0, (size_t)C * sizeof(double)); return 0; }
  const size_t nbytes = (rows * C * sizeof(float) + 255) & ~(size_t)255;
  TRY(ctx->stage.ensure(ctx, nbytes));
  CU(ctx, cudaMemcpyAsync(ctx->stage.p, x, rows * C * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
  const int np = channel_sums_partials(rows, C);
  TRY(ctx->scratch.ensure(ctx, ((size_t)np + 1) * C * sizeof(double)));
  double* partial = (double*)ctx->scratch.p;
  double* res = partial + (size_t)np * C;
  launch_channel_sums((const float*)ctx->stage.p, rows, C, partial, np, res, ctx->stream); ctx->launches += 2;
  CHECK_LAUNCH(ctx);
  CU(ctx, cudaMemcpyAsync(out, res, (size_t)C * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

// ---- projection chain -------------------------------------------------------------------------------
// Largest per-item fp32 element count any intermediate of a chain starting at `layer` can have.
static size_t chain_max_elems(const dcv_ctx* ctx, int layer) {
  size_t m = (size_t)ctx->in_H * ctx->in_W * ctx->in_C;
  for (int j = 0; j <= layer; ++j) {
    const Layer& l = ctx->net[j];
    m = std::max(m, (size_t)l.H * l.W * l.C);
  }
  return m;
}

static int ensure_nonneg(dcv_ctx* ctx, bool& nonneg, float* data, size_t n) {
  if (nonneg) return 0;
  launch_relu(data, n, data, ctx->stream); ctx->launches++;
  nonneg = true;
  return 0;
}

// Runs the chain for `n` feature maps (already on the device in ctx->feat_dev at `foff`) and leaves the
// result, (n*B, H0, W0, C0) fp32, at *result (a pointer into the workspace).
// The data is in one of three forms while it travels down:
//   MAP    one single-channel map per item (the projected tensor has ONE non-zero channel, so it is
//          never materialised; pools unpool the map, the first conv below runs as a rank-1 update)
//   TENSOR full (P,h,w,C) NHWC tensors
//   VEC    (P,C) rows, above a Flatten (dense layers)
//
// With DCV_BF16 the TENSOR form lives in the padded-flat bf16 layout of conv_tc.cu between the first
// (rank-1) and the last backward conv; only that last conv writes fp32 again.  `final_dst`, when not
// null, is where the last kernel of the chain writes directly (saves a device-to-device copy).
// DCV_FP32 projections run their backward convs on the tensor cores too (operands split into three bf16 planes,
// fp32 accumulation in registers: tc_exact.cuh) unless DCV_FP32_TC=0 asks for the FFMA kernels.  Measured on the
// block5_conv3 sweep against the reference's golden projections: rel-L2 1.8e-6 (FFMA 6.3e-6), 3.6x the throughput.
static bool fp32_on_tc(const dcv_ctx* ctx, const Layer& l) { return ctx->fp32_tc != 0 && l.wd != nullptr; }

// fp32 mode, two-plane scheme: the conv directly below conv layer j (no pool in between) can take this conv's result as
// ready-made operand planes - the split pass between the two disappears (conv_tc.cu, tc_conv_bwd_exact).  Not when the
// lower conv rectifies its input and this one does not (the ReLU pass in between works on the fp32 tensor).
static const Layer* exact_fuse_target(dcv_ctx* ctx, int j, bool) {
  if (j < 1 || !tc_bwd_exact_can_fuse_split()) return nullptr;
  const Layer& l = ctx->net[j];
  const Layer& nx = ctx->net[j - 1];
  if (nx.kind != LK_CONV || nx.cout != l.cin || !fp32_on_tc(ctx, nx)) return nullptr;
  if (nx.relu && !l.relu) return nullptr;
  return &nx;
}

static int project_chunk(dcv_ctx* ctx, int layer, int foff, int n, int vis_mode, int precision, float* final_dst,
                         float** result) {
  enum { M_MAP, M_TENSOR, M_VEC };
  __nv_bfloat16* hb[2] = {(__nv_bfloat16*)ctx->ws[4].p, (__nv_bfloat16*)ctx->ws[5].p};
  int hcur = 0;
  bool rep_bf16 = false;             // TENSOR data currently lives in hb[hcur] (padded-flat bf16)
  bool exact_chain = false;          // fp32 mode: the current tensor was written by tc_conv_bwd_exact (its per-item maxima are known)
  ctx->split_ready = -1;
  int lowest_conv = -1;
  for (int j = 0; j <= layer; ++j) if (ctx->net[j].kind == LK_CONV) { lowest_conv = j; break; }
  const int B = ctx->in_B;
  const int P = n * B;
  const int* feat = (const int*)ctx->feat_dev.p + foff;
  const Layer& top = ctx->net[layer];
  float* bufs[2] = {(float*)ctx->ws[0].p, (float*)ctx->ws[1].p};
  int cur = 0;                       // bufs[cur] holds the current data
  float* maps = (float*)ctx->ws[2].p;
  float* maps2 = (float*)ctx->ws[3].p;

  // seed (deepdream.py:450-465): feature map feat[k] of the layer's output, all other maps zero
  int h = top.H, w = top.W, C = top.C;
  launch_gather_map((const float*)top.act.p, B, h, w, C, feat, n, maps, ctx->stream); ctx->launches++;
  if (vis_mode == DCV_VIS_MAX) {
    launch_keep_max(maps, n, (size_t)B * h * w, nullptr, ctx->stream); ctx->launches++;
  }
  int mode = M_MAP;
  if (top.kind == LK_DENSE || top.kind == LK_FLATTEN) {
    launch_scatter_map(maps, n, B, 1, 1, C, feat, bufs[cur], ctx->stream); ctx->launches++;
    mode = M_VEC;
  }
  // Data is non-negative whenever the op that produced it ended in a ReLU; tracking this lets the
  // '<name>_activation'.down ReLUs (deepdream.py:260) run only where they can change a value.
  bool nonneg = false;
  for (int j = layer; j >= 0; --j) {
    const Layer& l = ctx->net[j];
    if (l.kind == LK_CONV) { nonneg = l.relu != 0; break; }
    if (l.kind == LK_DENSE) { nonneg = l.act_kind != DCV_ACT_LINEAR; break; }
  }

  for (int j = layer; j >= 0; --j) {
    const Layer& l = ctx->net[j];
    if (l.kind == LK_DENSE) {
      if (mode != M_VEC) return ctx->fail(DCV_ERR_STATE, "chain: dense layer %d below a spatial layer", j);
      if (l.cout != C) return ctx->fail(DCV_ERR_STATE, "chain: dense layer %d has %d outputs, data has %d", j, l.cout, C);
      if (j != layer) {              // '<name>_activation'.down of this layer comes first
        if (l.act_kind == DCV_ACT_RELU) TRY(ensure_nonneg(ctx, nonneg, bufs[cur], (size_t)P * C));
        if (l.act_kind == DCV_ACT_SOFTMAX) {
          launch_softmax_rows(bufs[cur], P, C, bufs[cur], ctx->stream); ctx->launches++;
        }
      }
      launch_dense_bwd(bufs[cur], l.w, P, l.cin, l.cout, bufs[cur ^ 1], ctx->stream); ctx->launches++;
      cur ^= 1;
      C = l.cin;
      nonneg = false;
    } else if (l.kind == LK_FLATTEN) {
      if (mode != M_VEC) return ctx->fail(DCV_ERR_STATE, "chain: flatten layer %d below a spatial layer", j);
      h = l.in_H; w = l.in_W; C = l.in_C;      // DFlatten.down is a reshape (deepdream.py:355-366)
      mode = M_TENSOR;
    } else if (l.kind == LK_POOL) {
      if (mode == M_VEC) return ctx->fail(DCV_ERR_STATE, "chain: pool layer %d above a flatten", j);
      if ((l.in_H & 1) || (l.in_W & 1))
        return ctx->fail(DCV_ERR_ARG, "unpool of layer %d: odd input size %dx%d (the reference raises here too)", j, l.in_H, l.in_W);
      if (mode == M_MAP) {
        ProfScope ps(ctx, DCV_K_UNPOOL, 0.0, (double)P * h * w * (5.0 + 16.0));
        launch_unpool_map(maps, (const uint8_t*)l.sw.p, B, h, w, C, feat, n, maps2, ctx->stream); ctx->launches++;
        std::swap(maps, maps2);
      } else if (rep_bf16) {
        // 64-channel tail (block1 of VGG16): unpool -> 64 -> 64 conv [-> last conv] as ONE kernel (tc_strip.cuh); the
        // unpooled tensor and the conv's 64-channel output never touch HBM.  DCV_TC_TAIL: bit 0 = fuse the unpool, bit 1 =
        // fuse the last conv (default 3).
        static const int tail = getenv("DCV_TC_TAIL") ? atoi(getenv("DCV_TC_TAIL")) : 3;
        const bool big = (long long)P * (l.in_H + 1) * (l.in_W + 1) / 120 >= ctx->num_sms;
        if ((tail & 1) && big && j >= 1 && tc_strip_conv_ok(ctx->net[j - 1]) && ctx->net[j - 1].cout == 64 && C == 64 &&
            (nonneg || !ctx->net[j - 1].relu)) {
          const Layer& c2 = ctx->net[j - 1];
          const Layer* c1 = ((tail & 2) && j >= 2 && j - 2 == lowest_conv && tc_strip_last_ok(c2, ctx->net[j - 2])) ? &ctx->net[j - 2] : nullptr;
          if (c1) {
            if (j - 2 == 0 && final_dst) bufs[cur ^ 1] = final_dst;
            TRY(tc_strip(ctx, c2, c1, hb[hcur], (const uint8_t*)l.sw.p, B, P, l.in_H, l.in_W, nullptr, bufs[cur ^ 1], -1));
            cur ^= 1;
            rep_bf16 = false;
            nonneg = c1->relu != 0;
            C = c1->cin;
            j -= 2;
          } else {
            TRY(tc_strip(ctx, c2, nullptr, hb[hcur], (const uint8_t*)l.sw.p, B, P, l.in_H, l.in_W, hb[hcur ^ 1], nullptr, -1));
            hcur ^= 1;
            nonneg = c2.relu != 0;
            C = c2.cin;
            j -= 1;
          }
          h = l.in_H; w = l.in_W;
          mode = M_TENSOR;
          continue;
        }
        TRY(tc_unpool(ctx, hb[hcur], (const uint8_t*)l.sw.p, P, B, h, w, l.in_H, l.in_W, C, hb[hcur ^ 1]));
        hcur ^= 1;
      } else if (precision == DCV_FP32 && mode == M_TENSOR && j >= 1 && ctx->net[j - 1].kind == LK_CONV && ctx->net[j - 1].cout == C &&
                 fp32_on_tc(ctx, ctx->net[j - 1]) && tc_bwd_exact_fuses_unpool() && (nonneg || !ctx->net[j - 1].relu)) {
        // fp32 mode: the unpool is done by the operand split of the conv below (the unpooled fp32 tensor is never written)
        Layer& cl = ctx->net[j - 1];
        if (j - 1 == 0 && final_dst) bufs[cur ^ 1] = final_dst;
        TRY(tc_prepare_backward_exact(ctx, cl));
        const Layer* nx = exact_fuse_target(ctx, j - 1, final_dst != nullptr);
        if (nx) TRY(tc_prepare_backward_exact(ctx, ctx->net[j - 2]));
        TRY(tc_conv_bwd_exact_unpool(ctx, cl, bufs[cur], (const uint8_t*)l.sw.p, B, P, l.in_H, l.in_W, bufs[cur ^ 1], cl.relu, exact_chain, nx));
        exact_chain = true;
        cur ^= 1;
        h = l.in_H; w = l.in_W;
        nonneg = cl.relu != 0;
        C = cl.cin;
        --j;                          // the conv layer is done
        continue;
      } else {
        // algorithmic bytes: 5*HWC/4 read + 4*HWC written, (H,W) = unpooled size
        ProfScope ps(ctx, DCV_K_UNPOOL, 0.0, (double)P * l.in_H * l.in_W * C * (4.0 + 5.0 / 4.0));
        launch_unpool_switch_items(bufs[cur], (const uint8_t*)l.sw.p, P, B, l.in_H, l.in_W, C, 0, bufs[cur ^ 1], ctx->stream);
        ctx->launches++;
        cur ^= 1;
      }
      h = l.in_H; w = l.in_W;
    } else {   // LK_CONV
      if (mode == M_VEC) return ctx->fail(DCV_ERR_STATE, "chain: conv layer %d above a flatten", j);
      if (l.cout != C) return ctx->fail(DCV_ERR_STATE, "chain: layer %d has %d outputs, data has %d channels", j, l.cout, C);
      const bool more_convs = j > lowest_conv;      // another backward conv follows
      float* fdst = (j == 0 && final_dst) ? final_dst : nullptr;
      if (mode == M_MAP) {
        if (j != layer && l.relu) TRY(ensure_nonneg(ctx, nonneg, maps, (size_t)P * h * w));
        if (precision == DCV_BF16 && more_convs) {
          TRY(tc_rank1(ctx, maps, l, feat, n, B, h, w, hb[hcur]));
          rep_bf16 = true;
        } else {
          if (fdst) { bufs[cur] = fdst; }
          ProfScope ps(ctx, DCV_K_RANK1, 2.0 * 9 * P * h * w * (double)l.cin, 4.0 * P * h * w * (1.0 + l.cin));
          launch_rank1_deconv(maps, l.wd, l.cout, l.cin, feat, n, B, h, w, l.relu, bufs[cur], ctx->stream); ctx->launches++;
        }
        mode = M_TENSOR;
      } else {
        if (j != layer && l.relu && !nonneg) {
          if (rep_bf16) return ctx->fail(DCV_ERR_UNSUPPORTED, "DCV_BF16 chain: a linear conv above a ReLU conv is not supported");
          TRY(ensure_nonneg(ctx, nonneg, bufs[cur], (size_t)P * h * w * C));
        }
        const double fl = 2.0 * 9 * P * h * w * (double)l.cin * l.cout;
        if (precision == DCV_BF16) {
          if (!rep_bf16) {              // entering from the fp32 side (below a Flatten)
            TRY(tc_pad_convert(ctx, bufs[cur], P, h, w, C, hb[hcur]));
            rep_bf16 = true;
          }
          if (more_convs) {
            // a pool layer directly below: its switch-driven unpool is fused into this conv's epilogue
            static const int fuse_unpool = getenv("DCV_TC_FUSE_UNPOOL") ? atoi(getenv("DCV_TC_FUSE_UNPOOL")) : 1;
            // Measured: the 4x wider stores make the epilogue the bottleneck for N tiles <= 128 (block3_conv1 +0.14 ms,
            // block2_conv1 +0.48 ms net), so only layers that run with the N-tile-256 kernels fuse (DCV_TC_FUSE_UNPOOL=2: all)
            const bool wide = l.tc.n_pad % 256 == 0 && (long long)P * (h + 1) * (w + 1) >= 256LL * ctx->num_sms / 2;
            const Layer* pl = (fuse_unpool && (wide || fuse_unpool == 2) && j > 0 && ctx->net[j - 1].kind == LK_POOL) ? &ctx->net[j - 1] : nullptr;
            if (pl) {
              if ((pl->in_H & 1) || (pl->in_W & 1))
                return ctx->fail(DCV_ERR_ARG, "unpool of layer %d: odd input size %dx%d (the reference raises here too)", j - 1, pl->in_H, pl->in_W);
              TcUnpool up{(const uint8_t*)pl->sw.p, B, l.cin, pl->in_H, pl->in_W};
              TRY(tc_conv(ctx, l, hb[hcur], P, h, w, hb[hcur ^ 1], nullptr, -1, &up));
              h = pl->in_H; w = pl->in_W;
              nonneg = l.relu != 0;
              C = l.cin;
              hcur ^= 1;
              --j;                      // the pool layer is done
              continue;
            }
            static const int tail2 = getenv("DCV_TC_TAIL") ? atoi(getenv("DCV_TC_TAIL")) : 3;
            if ((tail2 & 2) && j - 1 == lowest_conv && tc_strip_conv_ok(l) && tc_strip_last_ok(l, ctx->net[j - 1]) &&
                (long long)P * (h + 1) * (w + 1) / 120 >= ctx->num_sms) {
              // 64 -> 64 conv with the last conv fused behind it (no pool in between, or unpool fusion switched off)
              const Layer& c1 = ctx->net[j - 1];
              if (j - 1 == 0 && final_dst) bufs[cur ^ 1] = final_dst;
              TRY(tc_strip(ctx, l, &c1, hb[hcur], nullptr, B, P, h, w, nullptr, bufs[cur ^ 1], -1));
              cur ^= 1;
              rep_bf16 = false;
              nonneg = c1.relu != 0;
              C = c1.cin;
              --j;
              continue;
            }
            TRY(tc_conv(ctx, l, hb[hcur], P, h, w, hb[hcur ^ 1], nullptr, -1));
            hcur ^= 1;
            cur ^= 1;                   // keeps the fp32 ping-pong bookkeeping below uniform
          } else {
            if (fdst) bufs[cur ^ 1] = fdst;
            TRY(tc_conv(ctx, l, hb[hcur], P, h, w, nullptr, bufs[cur ^ 1], -1));
            rep_bf16 = false;
          }
        } else if (fp32_on_tc(ctx, l)) {
          // fp32 mode on the tensor cores: operands split into three bf16 planes, six products (conv_tc.cu)
          if (fdst) bufs[cur ^ 1] = fdst;
          TRY(tc_prepare_backward_exact(ctx, ctx->net[j]));
          const Layer* nx = exact_fuse_target(ctx, j, final_dst != nullptr);
          if (nx) TRY(tc_prepare_backward_exact(ctx, ctx->net[j - 1]));
          TRY(tc_conv_bwd_exact(ctx, l, bufs[cur], P, h, w, bufs[cur ^ 1], l.relu, exact_chain, nx));
          exact_chain = true;           // the next conv's input maxima come out of this one's epilogue
        } else {
          if (fdst) bufs[cur ^ 1] = fdst;
          ConvEpilogue e; e.relu = l.relu;
          ProfScope ps(ctx, DCV_K_CONV_FP32, fl, 4.0 * P * h * w * ((double)l.cin + l.cout));
          // backward conv: input channels = l.cout, output channels = l.cin, weights wd [9][cout][cin]
          launch_conv3x3_fp32(bufs[cur], l.wd, P, h, w, l.cout, l.cin, e, bufs[cur ^ 1], ctx->stream); ctx->launches++;
        }
        cur ^= 1;
      }
      nonneg = l.relu != 0;
      C = l.cin;
    }
  }
  if (mode == M_MAP) {   // no conv below the projected layer: the result is the map in channel f of zeros
    launch_scatter_map(maps, n, B, h, w, C, feat, bufs[cur], ctx->stream); ctx->launches++;
  }
  CHECK_LAUNCH(ctx);
  *result = bufs[cur];
  return 0;
}

// Converts `items` fp32 projections of `len` floats at `src` into the requested result format at `dst` (device).
static int convert_result(dcv_ctx* ctx, const float* src, int items, size_t len, int fmt, void* dst) {
  if (fmt == DCV_OUT_U8) {
    TRY(ctx->scratch.ensure(ctx, (size_t)items * kItemPartials * 2 * sizeof(double)));
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 9.0 * items * len);      // two reads of the projection + one byte written
    launch_deprocess_items(src, items, len, (double*)ctx->scratch.p, (uint8_t*)dst, ctx->stream); ctx->launches += 2;
  } else {
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 6.0 * items * len);
    launch_convert_bf16(src, (size_t)items * len, (__nv_bfloat16*)dst, ctx->stream); ctx->launches++;
  }
  CHECK_LAUNCH(ctx);
  return 0;
}

// feat == nullptr: the n feature indices are already in ctx->feat_dev (written by a kernel, dcv_request_mosaic)
static int project_impl(dcv_ctx* ctx, int layer, const int* feat, int n, int vis_mode, int precision, void* out_v,
                        bool out_on_device, int fmt = DCV_OUT_F32) {
  if (!out_v || n < 0) return ctx->fail(DCV_ERR_ARG, "dcv_project: bad arguments");
  if (vis_mode != DCV_VIS_ALL && vis_mode != DCV_VIS_MAX) return ctx->fail(DCV_ERR_ARG, "Illegal visualize mode");
  if (precision != DCV_FP32 && precision != DCV_BF16) return ctx->fail(DCV_ERR_ARG, "dcv_project: unknown precision");
  if (fmt != DCV_OUT_F32 && fmt != DCV_OUT_BF16 && fmt != DCV_OUT_U8) return ctx->fail(DCV_ERR_ARG, "dcv_project: unknown output format %d", fmt);
  TRY(check_tape(ctx, ctx, layer, "dcv_project"));
  const Layer& top = ctx->net[layer];
  for (int k = 0; feat && k < n; ++k)
    if (feat[k] < 0 || feat[k] >= top.C) return ctx->fail(DCV_ERR_ARG, "dcv_project: feature %d out of range [0,%d)", feat[k], top.C);
  if (n == 0) return 0;
  const size_t esz = fmt == DCV_OUT_F32 ? sizeof(float) : (fmt == DCV_OUT_BF16 ? 2 : 1);
  char* out = (char*)out_v;
  const int B = ctx->in_B;
  const size_t out_elems = (size_t)B * ctx->in_H * ctx->in_W * ctx->in_C;   // per feature
  size_t max_elems = chain_max_elems(ctx, layer);                           // fp32 elements per item
  if (precision == DCV_BF16 && (top.kind == LK_CONV || top.kind == LK_POOL)) {
    // between the first and the last backward conv the data is bf16 (ws[4], ws[5]); fp32 buffers only
    // hold what the last conv (or anything below it) writes
    int lowest_conv = -1;
    for (int j = 0; j <= layer; ++j) if (ctx->net[j].kind == LK_CONV) { lowest_conv = j; break; }
    if (lowest_conv >= 0) {
      max_elems = (size_t)ctx->in_H * ctx->in_W * ctx->in_C;
      for (int j = 0; j < lowest_conv; ++j) max_elems = std::max(max_elems, (size_t)ctx->net[j].H * ctx->net[j].W * ctx->net[j].C);
      const Layer& lc = ctx->net[lowest_conv];
      max_elems = std::max(max_elems, (size_t)lc.H * lc.W * lc.cin);
    }
  }
  // padded-flat bf16 elements per item (DCV_BF16 only)
  size_t max_pad = 0;
  if (precision == DCV_BF16) {
    max_pad = tc_padded_elems(ctx->in_H, ctx->in_W, ctx->in_C);
    for (int j = 0; j <= layer; ++j) {
      const Layer& l = ctx->net[j];
      if (l.kind == LK_CONV || l.kind == LK_POOL) max_pad = std::max(max_pad, tc_padded_elems(l.H, l.W, l.C));
    }
  }
  // fp32 mode on the tensor cores: one split (three-plane) copy of a conv's input in ws[4]
  size_t max_split = 0;
  if (precision == DCV_FP32) {
    for (int j = 0; j <= layer; ++j) {
      const Layer& l = ctx->net[j];
      if (l.kind == LK_CONV && fp32_on_tc(ctx, l)) max_split = std::max(max_split, tc_split_elems(l.H, l.W, l.cout));
    }
  }
  // chunk so that the ping-pong buffers fit the workspace budget
  size_t per_feat = 2 * (size_t)B * (max_elems * sizeof(float) + max_pad * sizeof(__nv_bfloat16)) +
                    (size_t)B * max_split * sizeof(__nv_bfloat16);
  int chunk = (int)std::max<size_t>(1, ctx->workspace_budget / per_feat);
  chunk = std::min(chunk, n);
  // Host output: split into chunks so that the device-to-host copy of chunk i (second stream) overlaps the kernels
  // of chunk i+1; each chunk still carries >= 32 feature maps, far more rows than the GPU needs.  Measured with the
  // tapered tail below (512 maps): 2 base chunks (256,128,64,32,32) 40.2 k proj/s end to end, 3: 38.7, 4: 39.6, 5: 38.7.
  const bool pipelined = !out_on_device && n >= 64;
  // Compact results (uint8: a quarter of the bytes) want fewer, larger chunks: one base chunk halved down to 64 maps
  // (256,128,64,64) measured 44.9 k against 43.7 k end to end with the float32 schedule.
  static const int pipe_chunks_env = getenv("DCV_PIPE_CHUNKS") ? std::max(1, atoi(getenv("DCV_PIPE_CHUNKS"))) : 0;
  const int pipe_chunks = pipe_chunks_env ? pipe_chunks_env : (fmt == DCV_OUT_U8 ? 1 : 2);
  if (pipelined) chunk = std::min(chunk, std::max(32, (n + pipe_chunks - 1) / pipe_chunks));
  TRY(ctx->ws[0].ensure(ctx, (size_t)chunk * B * max_elems * sizeof(float)));
  TRY(ctx->ws[1].ensure(ctx, (size_t)chunk * B * max_elems * sizeof(float)));
  if (max_split) TRY(ctx->ws[4].ensure(ctx, (size_t)chunk * B * max_split * sizeof(__nv_bfloat16)));
  if (max_pad) {
    TRY(ctx->ws[4].ensure(ctx, (size_t)chunk * B * max_pad * sizeof(__nv_bfloat16)));
    TRY(ctx->ws[5].ensure(ctx, (size_t)chunk * B * max_pad * sizeof(__nv_bfloat16)));
  }
  size_t map_elems = 0;
  for (int j = 0; j <= layer; ++j) map_elems = std::max(map_elems, (size_t)ctx->net[j].H * ctx->net[j].W);
  map_elems = std::max(map_elems, (size_t)ctx->in_H * ctx->in_W);
  TRY(ctx->ws[2].ensure(ctx, (size_t)chunk * B * map_elems * sizeof(float)));
  TRY(ctx->ws[3].ensure(ctx, (size_t)chunk * B * map_elems * sizeof(float)));
  if (feat) {
    TRY(ctx->feat_dev.ensure(ctx, (size_t)n * sizeof(int)));
    CU(ctx, cudaMemcpyAsync(ctx->feat_dev.p, feat, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
  }
  if (pipelined) {
    for (int i = 0; i < 2; ++i) TRY(ctx->out_stage[i].ensure(ctx, (size_t)chunk * out_elems * esz));
  } else if (!out_on_device && fmt != DCV_OUT_F32) {
    TRY(ctx->out_stage[0].ensure(ctx, (size_t)chunk * out_elems * esz));
  }
  // Chunk schedule.  Host output: the copy of the LAST chunk is the only one nothing overlaps, so the last chunk is
  // halved repeatedly (128 -> 64, 32, 32): 0.34 ms of exposed copy instead of 1.4 ms for 512 maps of 224x224x3.
  std::vector<int> sizes;
  for (int f0 = 0; f0 < n; f0 += chunk) sizes.push_back(std::min(chunk, n - f0));
  static const int taper_env = getenv("DCV_PIPE_TAPER") ? atoi(getenv("DCV_PIPE_TAPER")) : -1;
  const int taper = taper_env >= 0 ? taper_env : (fmt == DCV_OUT_U8 ? 64 : 32);
  if (pipelined && taper > 0 && !sizes.empty()) {
    int last = sizes.back();
    sizes.pop_back();
    while (last > 2 * taper) { const int half = (last / 2 + 7) / 8 * 8; sizes.push_back(half); last -= half; }
    if (last > taper) { sizes.push_back(last - taper); last = taper; }
    sizes.push_back(last);
  }
  int ci = 0, f0 = 0;
  for (size_t k = 0; k < sizes.size(); f0 += sizes[k], ++k, ++ci) {
    const int nf = sizes[k];
    float* res = nullptr;
    char* dst = out + (size_t)f0 * out_elems * esz;
    // out_elems covers the B images of one feature map; every (feature, image) projection is its own item
    const size_t item_len = out_elems / B;
    if (pipelined) {
      const int s = ci & 1;
      void* stage = ctx->out_stage[s].p;
      if (ci >= 2) CU(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_copied[s], 0));    // stage buffer is free again
      if (fmt == DCV_OUT_F32) {
        TRY(project_chunk(ctx, layer, f0, nf, vis_mode, precision, (float*)stage, &res));
        if (res != stage)
          CU(ctx, cudaMemcpyAsync(stage, res, (size_t)nf * out_elems * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
      } else {
        TRY(project_chunk(ctx, layer, f0, nf, vis_mode, precision, nullptr, &res));
        TRY(convert_result(ctx, res, nf * B, item_len, fmt, stage));
      }
      CU(ctx, cudaEventRecord(ctx->ev_ready[s], ctx->stream));
      CU(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_ready[s], 0));
      CU(ctx, cudaMemcpyAsync(dst, stage, (size_t)nf * out_elems * esz, cudaMemcpyDeviceToHost, ctx->copy_stream));
      CU(ctx, cudaEventRecord(ctx->ev_copied[s], ctx->copy_stream));
      continue;
    }
    if (fmt != DCV_OUT_F32) {
      TRY(project_chunk(ctx, layer, f0, nf, vis_mode, precision, nullptr, &res));
      void* conv_dst = out_on_device ? (void*)dst : ctx->out_stage[0].p;
      TRY(convert_result(ctx, res, nf * B, item_len, fmt, conv_dst));
      if (!out_on_device)
        CU(ctx, cudaMemcpyAsync(dst, conv_dst, (size_t)nf * out_elems * esz, cudaMemcpyDeviceToHost, ctx->stream));
      continue;
    }
    TRY(project_chunk(ctx, layer, f0, nf, vis_mode, precision, out_on_device ? (float*)dst : nullptr, &res));
    if ((char*)res != dst)
      CU(ctx, cudaMemcpyAsync(dst, res, (size_t)nf * out_elems * sizeof(float),
                              out_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, ctx->stream));
  }
  if (pipelined) {      // the context's stream completes only after the last copies have landed
    for (int s = 0; s < 2 && s < ci; ++s) CU(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_copied[s], 0));
  }
  return 0;
}

extern "C" int dcv_project(dcv_ctx* ctx, int layer, const int* feat, int n, int vis_mode, int precision, float* out) {
  ENTER(ctx);
  if (!feat) return ctx->fail(DCV_ERR_ARG, "dcv_project: bad arguments");
  TRY(project_impl(ctx, layer, feat, n, vis_mode, precision, out, false));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_project_dev(dcv_ctx* ctx, int layer, const int* feat, int n, int vis_mode, int precision, float* out_dev) {
  ENTER(ctx);
  if (!feat) return ctx->fail(DCV_ERR_ARG, "dcv_project: bad arguments");
  return project_impl(ctx, layer, feat, n, vis_mode, precision, out_dev, true);
}

extern "C" int dcv_project_fmt(dcv_ctx* ctx, int layer, const int* feat, int n, int vis_mode, int precision, int out_format,
                               void* out) {
  ENTER(ctx);
  if (!feat) return ctx->fail(DCV_ERR_ARG, "dcv_project: bad arguments");
  TRY(project_impl(ctx, layer, feat, n, vis_mode, precision, out, false, out_format));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_project_fmt_dev(dcv_ctx* ctx, int layer, const int* feat, int n, int vis_mode, int precision, int out_format,
                                   void* out_dev) {
  ENTER(ctx);
  if (!feat) return ctx->fail(DCV_ERR_ARG, "dcv_project: bad arguments");
  return project_impl(ctx, layer, feat, n, vis_mode, precision, out_dev, true, out_format);
}

// ---- request tail (main.py:60-72) ---------------------------------------------------------------------
// uint8 BGR image -> caffe preprocess -> forward tape -> per-channel sums -> top-4 maps -> four projections ->
// 2x2 mosaic -> de-process, all on the device: 3*H*W bytes go up, 12*H*W bytes (+ 5 ints) come back.
extern "C" int dcv_request_mosaic(dcv_ctx* ctx, const uint8_t* img_bgr, int H, int W, int layer, int vis_mode, int precision,
                                  uint8_t* out_mosaic, int* feats_out) {
  ENTER(ctx);
  if (!img_bgr || !out_mosaic || H <= 0 || W <= 0) return ctx->fail(DCV_ERR_ARG, "dcv_request_mosaic: bad arguments");
  if (layer < 0 || layer >= (int)ctx->net.size()) return ctx->fail(DCV_ERR_ARG, "dcv_request_mosaic: layer %d out of range", layer);
  const size_t npix = (size_t)H * W;
  const size_t o_img = 0, o_pre = align256(npix * 3), o_mos = o_pre + align256(npix * 3 * sizeof(float));
  const size_t o_proj = o_mos + align256(npix * 12);
  TRY(ctx->stage.ensure(ctx, o_proj + align256(4 * npix * 3 * sizeof(float))));
  char* base = (char*)ctx->stage.p;
  CU(ctx, cudaMemcpyAsync(base + o_img, img_bgr, npix * 3, cudaMemcpyHostToDevice, ctx->stream));
  launch_preprocess_caffe((const uint8_t*)(base + o_img), npix, (float*)(base + o_pre), ctx->stream); ctx->launches++;
  TRY(forward_impl(ctx, (const float*)(base + o_pre), true, 1, H, W, layer));
  if (ctx->in_C != 3) return ctx->fail(DCV_ERR_UNSUPPORTED, "dcv_request_mosaic: the network takes %d input channels, the request image has 3", ctx->in_C);
  const Layer& l = ctx->net[layer];
  // find_top_filters (deepdream.py:369-380) on the device; the reference sums a float64 array at pool layers
  const size_t rows = (size_t)l.B * l.H * l.W;
  const int np = channel_sums_partials(rows, l.C);
  TRY(ctx->scratch.ensure(ctx, ((size_t)np + 1) * l.C * sizeof(double)));
  double* partial = (double*)ctx->scratch.p;
  double* sums = partial + (size_t)np * l.C;
  launch_channel_sums((const float*)l.act.p, rows, l.C, partial, np, sums, ctx->stream); ctx->launches += 2;
  TRY(ctx->feat_dev.ensure(ctx, 8 * sizeof(int)));
  launch_top_filters(sums, l.C, l.kind == LK_POOL ? 0 : 1, 4, (int*)ctx->feat_dev.p, ctx->stream); ctx->launches++;
  CHECK_LAUNCH(ctx);
  // main.py:67-68 indexes [0..3] of the top-8 list: with fewer than four positive maps the reference raises
  // IndexError.  The count is needed on the host before the projections are queued (indices of -1 must not run).
  int top4[5];
  CU(ctx, cudaMemcpyAsync(top4, ctx->feat_dev.p, sizeof top4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  if (feats_out) memcpy(feats_out, top4, 4 * sizeof(int));
  if (top4[4] < 4) return ctx->fail(DCV_ERR_STATE, "list index out of range: layer %d has only %d feature maps with a positive sum", layer, top4[4]);
  float* proj = (float*)(base + o_proj);
  TRY(project_impl(ctx, layer, nullptr, 4, vis_mode, precision, proj, true));
  const int npd = deprocess_partials(4 * npix * 3);
  TRY(ctx->scratch.ensure(ctx, (size_t)npd * 2 * sizeof(double)));
  {
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 9.0 * 4 * npix * 3);
    launch_deprocess_mosaic(proj, H, W, 3, (double*)ctx->scratch.p, npd, (uint8_t*)(base + o_mos), ctx->stream); ctx->launches += 2;
  }
  CHECK_LAUNCH(ctx);
  CU(ctx, cudaMemcpyAsync(out_mosaic, base + o_mos, npix * 12, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

// ---- de-process -------------------------------------------------------------------------------------
extern "C" int dcv_deprocess_dev(dcv_ctx* ctx, const float* x_dev, size_t n, uint8_t* out_dev) {
  ENTER(ctx);
  if (!x_dev || !out_dev) return ctx->fail(DCV_ERR_ARG, "dcv_deprocess: null pointer");
  if (n == 0) return 0;
  if (((uintptr_t)x_dev & 15) || ((uintptr_t)out_dev & 3)) return ctx->fail(DCV_ERR_ARG, "dcv_deprocess_dev: x must be 16-byte and out 4-byte aligned");
  const int np = deprocess_partials(n);
  TRY(ctx->scratch.ensure(ctx, (size_t)np * 2 * sizeof(double)));
  {
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 9.0 * n);      // two reads of x + one byte written
    launch_deprocess(x_dev, n, (double*)ctx->scratch.p, np, out_dev, nullptr, ctx->stream); ctx->launches += 2;
  }
  CHECK_LAUNCH(ctx);
  return 0;
}

extern "C" int dcv_deprocess(dcv_ctx* ctx, float* x, size_t n, uint8_t* out, int mutate) {
  ENTER(ctx);
  if (!x || !out) return ctx->fail(DCV_ERR_ARG, "dcv_deprocess: null pointer");
  if (n == 0) return 0;
  const size_t in_bytes = (n * sizeof(float) + 255) & ~(size_t)255;
  const size_t out_bytes = (n + 255) & ~(size_t)255;
  TRY(ctx->stage.ensure(ctx, in_bytes + out_bytes + (mutate ? in_bytes : 0)));
  float* xd = (float*)ctx->stage.p;
  uint8_t* od = (uint8_t*)ctx->stage.p + in_bytes;
  float* nd = mutate ? (float*)((char*)ctx->stage.p + in_bytes + out_bytes) : nullptr;
  CU(ctx, cudaMemcpyAsync(xd, x, n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
  const int np = deprocess_partials(n);
  TRY(ctx->scratch.ensure(ctx, (size_t)np * 2 * sizeof(double)));
  {
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 9.0 * n);      // two reads of x + one byte written
    launch_deprocess(xd, n, (double*)ctx->scratch.p, np, od, nd, ctx->stream); ctx->launches += 2;
  }
  CHECK_LAUNCH(ctx);
  CU(ctx, cudaMemcpyAsync(out, od, n, cudaMemcpyDeviceToHost, ctx->stream));
  if (mutate) CU(ctx, cudaMemcpyAsync(x, nd, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

// ---- standalone operators on host buffers -----------------------------------------------------------

extern "C" int dcv_maxpool_switch(dcv_ctx* ctx, const float* x, int B, int H, int W, int C, float* pooled, uint8_t* sw) {
  ENTER(ctx);
  if (!x || !pooled || !sw || B < 0 || H < 0 || W < 0 || C < 0) return ctx->fail(DCV_ERR_ARG, "dcv_maxpool_switch: bad arguments");
  const size_t nin = (size_t)B * H * W * C, nout = (size_t)B * (H / 2) * (W / 2) * C;
  if (nout == 0) return 0;
  const size_t o1 = align256(nin * 4), o2 = o1 + align256(nout * 4);
  TRY(ctx->stage.ensure(ctx, o2 + nout));
  char* base = (char*)ctx->stage.p;
  CU(ctx, cudaMemcpyAsync(base, x, nin * 4, cudaMemcpyHostToDevice, ctx->stream));
  {
    ProfScope ps(ctx, DCV_K_POOL, 0.0, (double)B * H * W * C * (4.0 + 5.0 / 4.0));
    launch_maxpool_switch((const float*)base, B, H, W, C, (float*)(base + o1), (uint8_t*)(base + o2), ctx->stream); ctx->launches++;
  }
  CHECK_LAUNCH(ctx);
  CU(ctx, cudaMemcpyAsync(pooled, base + o1, nout * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaMemcpyAsync(sw, base + o2, nout, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_unpool_switch(dcv_ctx* ctx, const float* x, const uint8_t* sw, int B, int H, int W, int C, float* out) {
  ENTER(ctx);
  if (!x || !sw || !out || B < 0 || H < 0 || W < 0 || C < 0) return ctx->fail(DCV_ERR_ARG, "dcv_unpool_switch: bad arguments");
  if ((H & 1) || (W & 1)) return ctx->fail(DCV_ERR_ARG, "dcv_unpool_switch: output size %dx%d must be even (the reference raises on odd sizes)", H, W);
  const size_t nout = (size_t)B * H * W * C, nin = nout / 4;
  if (nout == 0) return 0;
  const size_t o1 = align256(nin * 4), o2 = o1 + align256(nin);
  TRY(ctx->stage.ensure(ctx, o2 + nout * 4));
  char* base = (char*)ctx->stage.p;
  CU(ctx, cudaMemcpyAsync(base, x, nin * 4, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(base + o1, sw, nin, cudaMemcpyHostToDevice, ctx->stream));
  {
    ProfScope ps(ctx, DCV_K_UNPOOL, 0.0, (double)B * H * W * C * (4.0 + 5.0 / 4.0));
    launch_unpool_switch((const float*)base, (const uint8_t*)(base + o1), B, H, W, C, 0, (float*)(base + o2), ctx->stream); ctx->launches++;
  }
  CHECK_LAUNCH(ctx);
  CU(ctx, cudaMemcpyAsync(out, base + o2, nout * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_conv3x3(dcv_ctx* ctx, const float* x, const float* w_hwio, const float* bias, int B, int H, int W,
                           int cin, int cout, int relu, float* out) {
  ENTER(ctx);
  if (!x || !w_hwio || !out || B <= 0 || H <= 0 || W <= 0 || cin <= 0 || cout <= 0) return ctx->fail(DCV_ERR_ARG, "dcv_conv3x3: bad arguments");
  const size_t nin = (size_t)B * H * W * cin, nout = (size_t)B * H * W * cout, nw = (size_t)9 * cin * cout;
  const size_t o1 = align256(nin * 4), o2 = o1 + align256(nw * 4), o3 = o2 + align256((size_t)cout * 4);
  TRY(ctx->stage.ensure(ctx, o3 + nout * 4));
  char* base = (char*)ctx->stage.p;
  CU(ctx, cudaMemcpyAsync(base, x, nin * 4, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(base + o1, w_hwio, nw * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (bias) CU(ctx, cudaMemcpyAsync(base + o2, bias, (size_t)cout * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (ctx->fwd_mode == 1 && cin >= 16) {
    // the kernel the forward tape uses for this layer shape (forward_impl): split-operand tcgen05
    Layer tmp;
    tmp.kind = LK_CONV; tmp.cin = cin; tmp.cout = cout; tmp.relu = relu ? 1 : 0;
    tmp.w = (float*)(base + o1);
    tmp.bias = bias ? (float*)(base + o2) : nullptr;
    int r = tc_prepare_forward_exact(ctx, tmp);
    if (!r) r = tc_conv_fwd_exact(ctx, tmp, (const float*)base, B, H, W, (float*)(base + o3));
    if (!r && cudaStreamSynchronize(ctx->stream) != cudaSuccess) r = ctx->fail(DCV_ERR_CUDA, "dcv_conv3x3: kernel failed");
    tc_free_layer(tmp);
    if (r) return r;
  } else {
    ConvEpilogue e; e.bias = bias ? (const float*)(base + o2) : nullptr; e.relu = relu;
    launch_conv3x3_fp32((const float*)base, (const float*)(base + o1), B, H, W, cin, cout, e, (float*)(base + o3), ctx->stream);
    ctx->launches++;
  }
  CHECK_LAUNCH(ctx);
  CU(ctx, cudaMemcpyAsync(out, base + o3, nout * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_deconv3x3(dcv_ctx* ctx, const float* x, const float* w_hwio, int B, int H, int W, int cin, int cout,
                             int relu, int precision, float* out) {
  ENTER(ctx);
  if (!x || !w_hwio || !out || B <= 0 || H <= 0 || W <= 0 || cin <= 0 || cout <= 0) return ctx->fail(DCV_ERR_ARG, "dcv_deconv3x3: bad arguments");
  if (precision != DCV_FP32 && precision != DCV_BF16) return ctx->fail(DCV_ERR_ARG, "dcv_deconv3x3: unknown precision");
  const size_t nin = (size_t)B * H * W * cout, nout = (size_t)B * H * W * cin, nw = (size_t)9 * cin * cout;
  const size_t o1 = align256(nin * 4), o2 = o1 + align256(nw * 4), o3 = o2 + align256(nw * 4);
  TRY(ctx->stage.ensure(ctx, o3 + nout * 4));
  char* base = (char*)ctx->stage.p;
  CU(ctx, cudaMemcpyAsync(base, x, nin * 4, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(base + o1, w_hwio, nw * 4, cudaMemcpyHostToDevice, ctx->stream));
  launch_flip_transpose((const float*)(base + o1), cin, cout, (float*)(base + o2), ctx->stream); ctx->launches++;
  if (precision == DCV_BF16) {
    Layer tmp;
    tmp.kind = LK_CONV; tmp.cin = cin; tmp.cout = cout; tmp.relu = relu ? 1 : 0;
    tmp.wd = (float*)(base + o2);
    int r = tc_prepare_layer(ctx, tmp);
    if (!r) r = tc_deconv(ctx, tmp, (const float*)base, B, H, W, (float*)(base + o3));
    if (!r && cudaStreamSynchronize(ctx->stream) != cudaSuccess) r = ctx->fail(DCV_ERR_CUDA, "dcv_deconv3x3: kernel failed");
    tc_free_layer(tmp);
    if (r) return r;
  } else if (ctx->fp32_tc) {
    // the kernels the DCV_FP32 projection chain uses: split-operand tcgen05 (GEMM + col2im form for <= 3 channels)
    Layer tmp;
    tmp.kind = LK_CONV; tmp.cin = cin; tmp.cout = cout; tmp.relu = relu ? 1 : 0;
    tmp.wd = (float*)(base + o2);
    int r = tc_prepare_backward_exact(ctx, tmp);
    if (!r) r = tc_conv_bwd_exact(ctx, tmp, (const float*)base, B, H, W, (float*)(base + o3), tmp.relu);
    if (!r && cudaStreamSynchronize(ctx->stream) != cudaSuccess) r = ctx->fail(DCV_ERR_CUDA, "dcv_deconv3x3: kernel failed");
    tc_free_layer(tmp);
    if (r) return r;
  } else {
    ConvEpilogue e; e.relu = relu;
    launch_conv3x3_fp32((const float*)base, (const float*)(base + o2), B, H, W, cout, cin, e, (float*)(base + o3), ctx->stream);
    ctx->launches++;
  }
  CHECK_LAUNCH(ctx);
  CU(ctx, cudaMemcpyAsync(out, base + o3, nout * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

// Fused tail of the bf16 projection chain as one operator (unit parity of tc_strip.cuh): [DPooling.down] ->
// DConvolution2D.down of a 64 -> 64 conv -> [DConvolution2D.down of the conv below it].
extern "C" int dcv_deconv_tail(dcv_ctx* ctx, const float* x, const uint8_t* sw, int B, int H, int W, const float* w2_hwio, int cmid,
                               int c2, int relu2, const float* w1_hwio, int c0, int relu1, float* out) {
  ENTER(ctx);
  if (!x || !w2_hwio || !out || B <= 0 || H <= 0 || W <= 0 || cmid <= 0 || c2 <= 0 || (w1_hwio && c0 <= 0))
    return ctx->fail(DCV_ERR_ARG, "dcv_deconv_tail: bad arguments");
  if (sw && ((H & 1) || (W & 1))) return ctx->fail(DCV_ERR_ARG, "dcv_deconv_tail: unpool needs an even output size, got %dx%d", H, W);
  const int ih = sw ? H / 2 : H, iw = sw ? W / 2 : W;
  const size_t nin = (size_t)B * ih * iw * c2, nsw = sw ? nin : 0;
  const size_t nw2 = (size_t)9 * cmid * c2, nw1 = w1_hwio ? (size_t)9 * c0 * cmid : 0;
  const int cl = w1_hwio ? c0 : cmid;
  const size_t nout = (size_t)B * H * W * cl;
  const size_t o_sw = align256(nin * 4), o_w2 = o_sw + align256(nsw), o_w2d = o_w2 + align256(nw2 * 4), o_w1 = o_w2d + align256(nw2 * 4);
  const size_t o_w1d = o_w1 + align256(nw1 * 4), o_out = o_w1d + align256(nw1 * 4);
  const size_t o_mid = o_out + align256(nout * 4);
  const size_t pad_in = (size_t)B * tc_padded_elems(ih, iw, c2) * 2, pad_out = (size_t)B * tc_padded_elems(H, W, cmid) * 2;
  TRY(ctx->stage.ensure(ctx, o_mid + align256(pad_in) + align256(pad_out)));
  char* base = (char*)ctx->stage.p;
  CU(ctx, cudaMemcpyAsync(base, x, nin * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (sw) CU(ctx, cudaMemcpyAsync(base + o_sw, sw, nsw, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(base + o_w2, w2_hwio, nw2 * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (w1_hwio) CU(ctx, cudaMemcpyAsync(base + o_w1, w1_hwio, nw1 * 4, cudaMemcpyHostToDevice, ctx->stream));
  Layer l2, l1;
  l2.kind = LK_CONV; l2.cin = cmid; l2.cout = c2; l2.relu = relu2 ? 1 : 0; l2.wd = (float*)(base + o_w2d);
  launch_flip_transpose((const float*)(base + o_w2), cmid, c2, l2.wd, ctx->stream); ctx->launches++;
  int r = tc_prepare_layer(ctx, l2);
  if (!r && w1_hwio) {
    l1.kind = LK_CONV; l1.cin = c0; l1.cout = cmid; l1.relu = relu1 ? 1 : 0; l1.wd = (float*)(base + o_w1d);
    launch_flip_transpose((const float*)(base + o_w1), c0, cmid, l1.wd, ctx->stream); ctx->launches++;
    r = tc_prepare_layer(ctx, l1);
  }
  __nv_bfloat16* pin = (__nv_bfloat16*)(base + o_mid);
  __nv_bfloat16* pmid = (__nv_bfloat16*)(base + o_mid + align256(pad_in));
  if (!r) r = tc_pad_convert(ctx, (const float*)base, B, ih, iw, c2, pin);
  if (!r && !(tc_strip_conv_ok(l2) && (!w1_hwio || tc_strip_last_ok(l2, l1))))
    r = ctx->fail(DCV_ERR_UNSUPPORTED, "dcv_deconv_tail: the fused tail needs a 64 -> 64 conv (and <= 3 channels below it)");
  if (!r) r = tc_strip(ctx, l2, w1_hwio ? &l1 : nullptr, pin, sw ? (const uint8_t*)(base + o_sw) : nullptr, B, B, H, W,
                       w1_hwio ? nullptr : pmid, w1_hwio ? (float*)(base + o_out) : nullptr, -1);
  if (!r && !w1_hwio) {
    // padded-flat bf16 -> compact fp32 on the host side of the test operator: copy the padded tensor back as it is
    r = 0;
  }
  if (!r && cudaStreamSynchronize(ctx->stream) != cudaSuccess) r = ctx->fail(DCV_ERR_CUDA, "dcv_deconv_tail: kernel failed: %s", cudaGetErrorString(cudaGetLastError()));
  tc_free_layer(l2);
  if (w1_hwio) tc_free_layer(l1);
  if (r) return r;
  if (w1_hwio) {
    CU(ctx, cudaMemcpyAsync(out, base + o_out, nout * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
  } else {
    // un-pad on the host: rows (b, y, x) of the padded-flat bf16 tensor, 64 channels each
    std::vector<uint16_t> tmp((size_t)B * (H + 1) * (W + 1) * 64);
    CU(ctx, cudaMemcpyAsync(tmp.data(), pmid, tmp.size() * 2, cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    for (int b = 0; b < B; ++b)
      for (int y = 0; y < H; ++y)
        for (int xx = 0; xx < W; ++xx) {
          const uint16_t* src = tmp.data() + (((size_t)b * (H + 1) + y) * (W + 1) + xx) * 64;
          float* dst = out + (((size_t)b * H + y) * W + xx) * cmid;
          for (int c = 0; c < cmid; ++c) { uint32_t u = (uint32_t)src[c] << 16; memcpy(dst + c, &u, 4); }
        }
  }
  return 0;
}

extern "C" int dcv_activation(dcv_ctx* ctx, const float* x, int rows, int n, int act, float* out) {
  ENTER(ctx);
  if (!x || !out || rows < 0 || n < 0 || act < 0 || act > 2) return ctx->fail(DCV_ERR_ARG, "dcv_activation: bad arguments");
  const size_t total = (size_t)rows * n;
  if (total == 0) return 0;
  TRY(ctx->stage.ensure(ctx, total * 4));
  float* d = (float*)ctx->stage.p;
  CU(ctx, cudaMemcpyAsync(d, x, total * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (act == DCV_ACT_RELU) { launch_relu(d, total, d, ctx->stream); ctx->launches++; }
  if (act == DCV_ACT_SOFTMAX) { launch_softmax_rows(d, rows, n, d, ctx->stream); ctx->launches++; }
  CHECK_LAUNCH(ctx);
  CU(ctx, cudaMemcpyAsync(out, d, total * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_dense(dcv_ctx* ctx, const float* x, const float* w, const float* bias, int B, int nin, int nout,
                         int act, float* out) {
  ENTER(ctx);
  if (!x || !w || !out || B <= 0 || nin <= 0 || nout <= 0 || act < 0 || act > 2) return ctx->fail(DCV_ERR_ARG, "dcv_dense: bad arguments");
  const size_t o1 = align256((size_t)B * nin * 4), o2 = o1 + align256((size_t)nin * nout * 4);
  const size_t o3 = o2 + align256((size_t)nout * 4), o4 = o3 + align256((size_t)dense_splits(nin) * B * nout * 4);
  TRY(ctx->stage.ensure(ctx, o4 + (size_t)B * nout * 4));
  char* base = (char*)ctx->stage.p;
  CU(ctx, cudaMemcpyAsync(base, x, (size_t)B * nin * 4, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(base + o1, w, (size_t)nin * nout * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (bias) CU(ctx, cudaMemcpyAsync(base + o2, bias, (size_t)nout * 4, cudaMemcpyHostToDevice, ctx->stream));
  float* od = (float*)(base + o4);
  launch_dense_fwd((const float*)base, (const float*)(base + o1), bias ? (const float*)(base + o2) : nullptr, B, nin, nout,
                   act, (float*)(base + o3), od, ctx->stream);
  ctx->launches += 2;
  if (act == DCV_ACT_SOFTMAX) { launch_softmax_rows(od, B, nout, od, ctx->stream); ctx->launches++; }
  CHECK_LAUNCH(ctx);
  CU(ctx, cudaMemcpyAsync(out, od, (size_t)B * nout * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_dense_transpose(dcv_ctx* ctx, const float* x, const float* w, int B, int nin, int nout, float* out) {
  ENTER(ctx);
  if (!x || !w || !out || B <= 0 || nin <= 0 || nout <= 0) return ctx->fail(DCV_ERR_ARG, "dcv_dense_transpose: bad arguments");
  const size_t o1 = align256((size_t)B * nout * 4), o2 = o1 + align256((size_t)nin * nout * 4);
  TRY(ctx->stage.ensure(ctx, o2 + (size_t)B * nin * 4));
  char* base = (char*)ctx->stage.p;
  CU(ctx, cudaMemcpyAsync(base, x, (size_t)B * nout * 4, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(base + o1, w, (size_t)nin * nout * 4, cudaMemcpyHostToDevice, ctx->stream));
  launch_dense_bwd((const float*)base, (const float*)(base + o1), B, nin, nout, (float*)(base + o2), ctx->stream);
  ctx->launches++;
  CHECK_LAUNCH(ctx);
  CU(ctx, cudaMemcpyAsync(out, base + o2, (size_t)B * nin * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}
