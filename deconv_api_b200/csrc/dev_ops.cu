// Device-resident forms of the standalone HBM-streaming operators (pool, unpool, resize, roll): the caller owns
// the device buffers, nothing is copied or allocated, the call is asynchronous on the context's stream.  They are
// what a pipeline that keeps its images on the GPU calls, and what profiles/stream_kernels.py times back to back
// (a kernel timed right after a long pageable host copy runs its first ~100 us on ramping clocks).
#include "common.cuh"
#include "internal.cuh"

using namespace dcv;

#define ENTERV(ctx)                                                                                        \
  if (!(ctx)) return DCV_ERR_ARG;                                                                          \
  {                                                                                                        \
    cudaError_t e_ = cudaSetDevice((ctx)->device);                                                         \
    if (e_ != cudaSuccess) return (ctx)->fail(DCV_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e_));  \
  }

static int check_launch(dcv_ctx* ctx, const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
  return 0;
}

static bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

extern "C" int dcv_maxpool_switch_dev(dcv_ctx* ctx, const float* x_dev, int B, int H, int W, int C, float* pooled_dev,
                                      uint8_t* sw_dev) {
  ENTERV(ctx);
  if (!x_dev || !pooled_dev || !sw_dev || B < 0 || H < 0 || W < 0 || C < 0)
    return ctx->fail(DCV_ERR_ARG, "dcv_maxpool_switch_dev: bad arguments");
  if (!aligned16(x_dev) || !aligned16(pooled_dev) || !aligned16(sw_dev))
    return ctx->fail(DCV_ERR_ARG, "dcv_maxpool_switch_dev: device pointers must be 16-byte aligned");
  if ((size_t)B * (H / 2) * (W / 2) * C == 0) return 0;
  {
    ProfScope ps(ctx, DCV_K_POOL, 0.0, (double)B * H * W * C * (4.0 + 5.0 / 4.0));
    launch_maxpool_switch(x_dev, B, H, W, C, pooled_dev, sw_dev, ctx->stream);
    ctx->launches++;
  }
  return check_launch(ctx, "dcv_maxpool_switch_dev");
}

extern "C" int dcv_unpool_switch_dev(dcv_ctx* ctx, const float* x_dev, const uint8_t* sw_dev, int B, int H, int W, int C,
                                     float* out_dev) {
  ENTERV(ctx);
  if (!x_dev || !sw_dev || !out_dev || B < 0 || H < 0 || W < 0 || C < 0)
    return ctx->fail(DCV_ERR_ARG, "dcv_unpool_switch_dev: bad arguments");
  if ((H & 1) || (W & 1))
    return ctx->fail(DCV_ERR_ARG, "dcv_unpool_switch_dev: output size %dx%d must be even (the reference raises on odd sizes)", H, W);
  if (!aligned16(x_dev) || !aligned16(out_dev) || !aligned16(sw_dev))
    return ctx->fail(DCV_ERR_ARG, "dcv_unpool_switch_dev: device pointers must be 16-byte aligned");
  if ((size_t)B * H * W * C == 0) return 0;
  {
    ProfScope ps(ctx, DCV_K_UNPOOL, 0.0, (double)B * H * W * C * (4.0 + 5.0 / 4.0));
    launch_unpool_switch(x_dev, sw_dev, B, H, W, C, 0, out_dev, ctx->stream);
    ctx->launches++;
  }
  return check_launch(ctx, "dcv_unpool_switch_dev");
}

extern "C" int dcv_resize_bilinear_dev(dcv_ctx* ctx, const float* x_dev, int B, int H, int W, int C, int OH, int OW,
                                       float* out_dev) {
  ENTERV(ctx);
  if (!x_dev || !out_dev || B <= 0 || H <= 0 || W <= 0 || C <= 0 || OH <= 0 || OW <= 0)
    return ctx->fail(DCV_ERR_ARG, "dcv_resize_bilinear_dev: bad arguments");
  if (x_dev == out_dev) return ctx->fail(DCV_ERR_ARG, "dcv_resize_bilinear_dev: in-place resize is not supported");
  {
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 4.0 * ((double)B * H * W * C + (double)B * OH * OW * C));
    launch_resize_bilinear(x_dev, B, H, W, C, OH, OW, out_dev, ctx->stream);
    ctx->launches++;
  }
  return check_launch(ctx, "dcv_resize_bilinear_dev");
}

extern "C" int dcv_roll_dev(dcv_ctx* ctx, const float* x_dev, int B, int H, int W, int C, int oy, int ox, float* out_dev) {
  ENTERV(ctx);
  if (!x_dev || !out_dev || B <= 0 || H <= 0 || W <= 0 || C <= 0) return ctx->fail(DCV_ERR_ARG, "dcv_roll_dev: bad arguments");
  if (x_dev == out_dev) return ctx->fail(DCV_ERR_ARG, "dcv_roll_dev: in-place roll is not supported");
  {
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 8.0 * (double)B * H * W * C);
    launch_roll(x_dev, B, H, W, C, oy, ox, out_dev, ctx->stream);
    ctx->launches++;
  }
  return check_launch(ctx, "dcv_roll_dev");
}
