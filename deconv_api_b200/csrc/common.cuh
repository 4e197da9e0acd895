// Shared declarations of the dcv library (internal; the public surface is include/dcv.h).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stddef.h>
#include <string>
#include <vector>

#include "../../include/dcv.h"

namespace dcv {

constexpr int kNumSMs = 148;   // B200: 2 dies x 74 SMs; grids are sized in multiples of this

#ifdef __CUDACC__
// Sum of the fp64 (sum, sum of squares) block partials written by deprocess_stats_kernel, by all 256 threads of the
// calling block in a fixed order (strided per-thread sums, then a shuffle / shared-memory tree): every block and
// every run gets the same bits.  Result valid in all threads after the call.
__device__ __forceinline__ void block_sum_partials(const double* __restrict__ partial, int np, double& ts, double& tq) {
  __shared__ double ps_s[16], ps_q[16];   // up to 16 warps per block
  double s = 0.0, q = 0.0;
  for (int k = threadIdx.x; k < np; k += blockDim.x) { s += partial[2 * k]; q += partial[2 * k + 1]; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if ((threadIdx.x & 31) == 0) { ps_s[threadIdx.x >> 5] = s; ps_q[threadIdx.x >> 5] = q; }
  __syncthreads();
  ts = 0.0; tq = 0.0;
  const int nw = (blockDim.x + 31) >> 5;
  for (int k = 0; k < nw; ++k) { ts += ps_s[k]; tq += ps_q[k]; }
}
#endif

// ---- epilogue of the conv kernels -------------------------------------------------------------
struct ConvEpilogue {
  const float* bias = nullptr;      // per output channel, may be null (backward pass: zero bias)
  int relu = 0;                     // clamp at 0 (deconvnet rule, deepdream.py:69,86)
  const float* sign_mask = nullptr; // DeepDream backward: multiply by (mask > 0), same shape as out
  const float* add = nullptr;       // DeepDream backward: loss seed added after masking
  float add_scale = 0.f;
};

// ---- fp32 streaming / FFMA kernels (kernels_stream.cu, conv_fp32.cu) -----------------------------
void launch_maxpool_switch(const float* x, int B, int H, int W, int C, float* pooled, uint8_t* sw, cudaStream_t s);
// (H,W) = OUTPUT size; P items share the switches of Bsw images (item % Bsw)
void launch_unpool_switch_items(const float* x, const uint8_t* sw, int P, int Bsw, int H, int W, int C, int relu,
                                float* out, cudaStream_t s);
void launch_unpool_switch(const float* x, const uint8_t* sw, int B, int H, int W, int C, int relu, float* out, cudaStream_t s);
// per-channel fp64 sums over `rows` rows of a (rows, C) matrix; out[C] is overwritten
void launch_channel_sums(const float* x, size_t rows, int C, double* partial, int npartial, double* out, cudaStream_t s);
int channel_sums_partials(size_t rows, int C);
// de-process: block partials of (sum, sum of squares) in fp64, then the elementwise map
// norm (optional): the normalised floats the reference leaves in its argument (deepdream.py:485-490)
void launch_deprocess(const float* x, size_t n, double* partial, int npartial, uint8_t* out, float* norm, cudaStream_t s);
int deprocess_partials(size_t n);
void launch_sum_stats(const float* x, size_t n, double* partial, int npartial, cudaStream_t s);
// batched de-process: `items` arrays of `len` floats, each with its own mean / std; partial: items*kItemPartials*2 doubles
constexpr int kItemPartials = 8;
void launch_deprocess_items(const float* x, int items, size_t len, double* partial, uint8_t* out, cudaStream_t s);
void launch_convert_bf16(const float* x, size_t n, __nv_bfloat16* out, cudaStream_t s);
// request tail (main.py:60-72): caffe preprocess of a uint8 image, top-k feature selection, 2x2 mosaic + de-process
void launch_preprocess_caffe(const uint8_t* img, size_t npix, float* out, cudaStream_t s);
void launch_top_filters(const double* sums, int C, int as_f32, int top, int* feat_out, cudaStream_t s);
void launch_deprocess_mosaic(const float* x, int H, int W, int C, double* partial, int npartial, uint8_t* out, cudaStream_t s);
// seed maps: maps[(k*B+b), y, x] = act[b, y, x, feat[k]]
void launch_gather_map(const float* act, int B, int H, int W, int C, const int* feat_dev, int n, float* maps, cudaStream_t s);
// DCV_VIS_MAX: maps (n, len) -> keep only the entries equal to the map maximum
void launch_keep_max(float* maps, int n, size_t len, float* unused, cudaStream_t s);
// unpool of single-channel maps using channel feat[k] of the switch tensor: (n*B,h,w) -> (n*B,2h,2w)
void launch_unpool_map(const float* maps, const uint8_t* sw, int B, int h, int w, int C, const int* feat_dev, int n,
                       float* out, cudaStream_t s);
void launch_scatter_map(const float* maps, int n, int B, int H, int W, int C, const int* feat_dev, float* out, cudaStream_t s);
// rank-1 first backward conv: maps (n*B,H,W) x wd[tap][feat[k]][cin] -> out (n*B,H,W,cin), optional relu
void launch_rank1_deconv(const float* maps, const float* wd /*[9][cout][cin]*/, int cout, int cin, const int* feat_dev,
                         int n, int B, int H, int W, int relu, float* out, cudaStream_t s);
// direct 3x3 'same' convolution, NHWC fp32, weights [9][Cin][Cout]
void launch_conv3x3_fp32(const float* x, const float* w9, int B, int H, int W, int cin, int cout,
                         const ConvEpilogue& epi, float* out, cudaStream_t s);
// weight preparation (deepdream.py:80-81): HWIO (3,3,cin,cout) -> flipped+transposed [9][cout][cin]
void launch_flip_transpose(const float* w_hwio, int cin, int cout, float* wd, cudaStream_t s);
// dense layers
int dense_splits(int nin);
void launch_dense_fwd(const float* x, const float* w, const float* bias, int B, int nin, int nout, int act,
                      float* partial, float* out, cudaStream_t s);
void launch_dense_bwd(const float* x, const float* w, int B, int nin, int nout, float* out, cudaStream_t s);
void launch_softmax_rows(const float* x, int B, int n, float* out, cudaStream_t s);
void launch_relu(const float* x, size_t n, float* out, cudaStream_t s);
}  // namespace dcv
// dream.cu
void launch_resize_bilinear(const float* x, int B, int H, int W, int C, int OH, int OW, float* out, cudaStream_t s);
void launch_roll(const float* x, int B, int H, int W, int C, int oy, int ox, float* out, cudaStream_t s);
namespace dcv {

}  // namespace dcv
