// Tensor-core side of the library: padded-flat bf16 layout, weight packing, streaming helpers and the host
// dispatch of the tcgen05 kernels (tc_kernel.cuh single CTA, tc_kernel2.cuh CTA pair, tc_last.cuh GEMM+col2im for the
// 3-channel last layer, tc_dxg.cuh opt-in experiment) — hand-written for sm_100a.
//
// Layout ("padded-flat"): an activation of P items x (h,w) x C is stored as a 2-D bf16 matrix
//   rows = P*(h+1)*(w+1), cols = C_pad (multiple of 64); pixel (p,y,x) is row p*(h+1)*(w+1)+y*(w+1)+x,
// with one zero column (x == w) per image row and one zero row (y == h) per image.  In that layout the
// input a 3x3 tap (dy,dx) needs for output row r is simply row r + dy*(w+1) + dx: the halo of a 'same'
// convolution is the shared zero row/column, or an out-of-range row that TMA zero-fills.  So the A tile
// of every (tap, 64-channel block) is ONE plain 2-D TMA box of consecutive rows at a shifted row
// coordinate, landing in shared memory already in the 128-byte-swizzled K-major form UMMA reads.
//   GEMM view:  M = padded rows, N = output channels, K = 9 taps x C_pad;
//   A tile 128 x 64 (16 KB), B tile BLOCK_N x 64 from weights packed [tap][N_pad][K_pad] bf16.
// Output rows at pad positions are garbage by construction and are written as zeros (bf16 output, so
// the next layer's halo stays zero) or skipped (fp32 compact output).
// Used for: the backward projection chain (DCV_BF16), the fp32-accurate forward tape (three bf16 operand planes,
// six products) and the DeepDream forward/backward (bf16 tape).
#include "common.cuh"
#include "internal.cuh"
#include "tc_ptx.cuh"
#include "tc_kernel.cuh"
#include "tc_last.cuh"
#include "tc_kernel2.cuh"
#include "tc_dxg.cuh"
#include "tc_strip.cuh"
#include "tc_exact.cuh"

#include <cuda_fp16.h>

#include <algorithm>
#include <cstdlib>

using namespace dcv;
using namespace dcv::tc;

// =====================================================================================================
// Supporting kernels of the bf16 chain (HBM-streaming, 128-bit accesses along the channel axis).
namespace {

static inline int grid_for(size_t work, int block, int waves = 16) {
  size_t b = (work + block - 1) / block, cap = (size_t)kNumSMs * waves;
  if (b > cap) b = cap;
  return (int)(b < 1 ? 1 : b);
}

// Weights of the backward conv for UMMA: wd fp32 [9][K=cout][N=cin]  ->  bf16 [9][n_pad][k_pad] (K-major rows).
__global__ void pack_weights_tc_kernel(const float* __restrict__ wd, int kdim, int ndim, int k_pad, int n_pad,
                                       __nv_bfloat16* __restrict__ out) {
  const size_t total = (size_t)9 * n_pad * k_pad;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int k = (int)(i % k_pad);
    const int n = (int)((i / k_pad) % n_pad);
    const int t = (int)(i / ((size_t)k_pad * n_pad));
    float v = 0.f;
    if (k < kdim && n < ndim) v = __ldg(wd + ((size_t)t * kdim + k) * ndim + n);
    out[i] = __float2bfloat16_rn(v);
  }
}

// All-tap weights of a narrow last layer for deconv_last_kernel: wd fp32 [9][K][N] -> bf16 [32][k_pad], row = tap*N + c.
__global__ void pack_weights_alltap_kernel(const float* __restrict__ wd, int kdim, int ndim, int k_pad, __nv_bfloat16* __restrict__ out) {
  const int total = 32 * k_pad;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int k = i % k_pad, row = i / k_pad;
    float v = 0.f;
    if (row < 9 * ndim && k < kdim) {
      const int t = row / ndim, c = row % ndim;
      v = __ldg(wd + ((size_t)t * kdim + k) * ndim + c);
    }
    out[i] = __float2bfloat16_rn(v);
  }
}

// Exact-forward operands: x = hi + mid + lo with three bf16 planes (8 mantissa bits each, 24 together).
__device__ __forceinline__ void split3(float x, __nv_bfloat16& h, __nv_bfloat16& m, __nv_bfloat16& l) {
  h = __float2bfloat16_rn(x);
  const float r1 = x - __bfloat162float(h);
  m = __float2bfloat16_rn(r1);
  l = __float2bfloat16_rn(r1 - __bfloat162float(m));
}

// the all-tap weights of pack_weights_alltap_kernel as three planes: out [3][32][k_pad]
__global__ void pack_weights_alltap3_kernel(const float* __restrict__ wd, int kdim, int ndim, int k_pad, __nv_bfloat16* __restrict__ out) {
  const int total = 32 * k_pad;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int k = i % k_pad, row = i / k_pad;
    float v = 0.f;
    if (row < 9 * ndim && k < kdim) {
      const int t = row / ndim, c = row % ndim;
      v = __ldg(wd + ((size_t)t * kdim + k) * ndim + c);
    }
    __nv_bfloat16 h, m, l;
    split3(v, h, m, l);
    out[i] = h; out[total + i] = m; out[2 * total + i] = l;
  }
}

// fp32 compact (P,h,w,C) -> padded-flat bf16 with THREE planes per row: columns [hi | mid | lo], c_pad each.
__global__ void split3_pad_kernel(const float* __restrict__ x, int P, int h, int w, int C, int c_pad,
                                  __nv_bfloat16* __restrict__ out) {
  const int wp = w + 1, rp = (h + 1) * wp, cg = c_pad / 8;
  const size_t total = (size_t)P * rp * cg;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int g = (int)(i % cg);
    const size_t r = i / cg;
    const int q = (int)(r % rp);
    const int item = (int)(r / rp);
    const int y = q / wp, xx = q % wp;
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = 0.f;
    if (y < h && xx < w) {
      const float* src = x + (((size_t)item * h + y) * w + xx) * C + g * 8;
      if ((C & 3) == 0 && g * 8 + 8 <= C) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src + 4));
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (g * 8 + k < C) v[k] = __ldg(src + k);
      }
    }
    __nv_bfloat16 hh[8], mm[8], ll[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) split3(v[k], hh[k], mm[k], ll[k]);
    __nv_bfloat16* o = out + r * (3 * (size_t)c_pad) + g * 8;
    *reinterpret_cast<uint4*>(o) = *reinterpret_cast<const uint4*>(hh);
    *reinterpret_cast<uint4*>(o + c_pad) = *reinterpret_cast<const uint4*>(mm);
    *reinterpret_cast<uint4*>(o + 2 * c_pad) = *reinterpret_cast<const uint4*>(ll);
  }
}

// ---- fp16 two-plane scheme (tc_exact.cuh, SCHEME 1) ------------------------------------------------------------------
// Largest |x| of a tensor as a float bit pattern (non-negative floats order like unsigned integers): one atomicMax
// per block; *out must be zero before the launch.
__global__ void __launch_bounds__(256) absmax_kernel(const float* __restrict__ x, size_t n, uint32_t* __restrict__ out) {
  uint32_t m = 0;
  const size_t n4 = n / 4, stride = (size_t)gridDim.x * blockDim.x;
  const uint4* x4 = reinterpret_cast<const uint4*>(x);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const uint4 a = __ldg(x4 + i);
    m = max(max(m, a.x & 0x7FFFFFFFu), max(max(a.y & 0x7FFFFFFFu, a.z & 0x7FFFFFFFu), a.w & 0x7FFFFFFFu));
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) m = max(m, __float_as_uint(x[n4 * 4 + threadIdx.x]) & 0x7FFFFFFFu);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  __shared__ uint32_t red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 1; k < 8; ++k) m = max(m, red[k]);
    if (m) atomicMax(out, m);
  }
}

// The same per item: `items` consecutive arrays of `len` floats, out[item]; grid (blocks per item, items).  Every
// projection is scaled by ITS OWN maximum: the items of a batch are independent and can differ by many orders of
// magnitude (the ten softmax outputs of a `predictions` target span 1e-1 .. 1e-6), and a result that depends only on
// the item does not change with how the batch is chunked.
__global__ void __launch_bounds__(256) absmax_items_kernel(const float* __restrict__ x, size_t len, uint32_t* __restrict__ out) {
  const float* xi = x + (size_t)blockIdx.y * len;
  uint32_t m = 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x, i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if ((len & 3) == 0) {                               // every item starts 16-byte aligned
    const uint4* x4 = reinterpret_cast<const uint4*>(xi);
    const size_t n4 = len / 4;
    size_t i = i0;
    for (; i + stride < n4; i += 2 * stride) {
      const uint4 a = __ldg(x4 + i), b = __ldg(x4 + i + stride);
      m = max(m, max(max(a.x & 0x7FFFFFFFu, a.y & 0x7FFFFFFFu), max(a.z & 0x7FFFFFFFu, a.w & 0x7FFFFFFFu)));
      m = max(m, max(max(b.x & 0x7FFFFFFFu, b.y & 0x7FFFFFFFu), max(b.z & 0x7FFFFFFFu, b.w & 0x7FFFFFFFu)));
    }
    for (; i < n4; i += stride) {
      const uint4 a = __ldg(x4 + i);
      m = max(m, max(max(a.x & 0x7FFFFFFFu, a.y & 0x7FFFFFFFu), max(a.z & 0x7FFFFFFFu, a.w & 0x7FFFFFFFu)));
    }
  } else {
    for (size_t i = i0; i < len; i += stride) m = max(m, __float_as_uint(__ldg(xi + i)) & 0x7FFFFFFFu);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  __shared__ uint32_t red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 1; k < 8; ++k) m = max(m, red[k]);
    if (m) atomicMax(out + blockIdx.y, m);
  }
}

// x*s = hi + lo with two fp16 planes (11 significand bits each; s = power of two that puts the tensor's maximum into
// [2^13, 2^14)): exact to ~2^-22 relative, or 2^-39 of the maximum for the smallest values (fp16 subnormals).
__device__ __forceinline__ void split2h(float xs, __half& h, __half& l) {
  h = __float2half_rn(xs);
  l = __float2half_rn(xs - __half2float(h));
}

// Largest L1 norm of an output channel's backward weights: max over n of sum over (tap, k) of |wd[tap][k][n]|, as float bits
// (atomicMax on a zeroed word).  |conv output| <= that x the largest |input|: the scale of a fused split (tc_exact.cuh).
__global__ void weight_l1max_kernel(const float* __restrict__ wd, int kdim, int ndim, uint32_t* __restrict__ out) {
  const int n = blockIdx.x;
  float s = 0.f;
  for (int i = threadIdx.x; i < 9 * kdim; i += blockDim.x) s += fabsf(wd[(size_t)i * ndim + n]);
  __shared__ float part[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += part[k];
    atomicMax(out, __float_as_uint(t));
  }
}
// scale source of a fused split: bound[item] = largest |input| of the item x the layer's L1 maximum (a little above, for
// the rounding of both factors; pow2_scale_from_amax leaves a factor of four to spare besides)
__global__ void split_bound_kernel(const uint32_t* __restrict__ amax_in, const uint32_t* __restrict__ l1max, int P,
                                   uint32_t* __restrict__ bound) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < P) bound[i] = __float_as_uint(__uint_as_float(amax_in[i]) * __uint_as_float(*l1max) * 1.001f);
}

// fp32 compact (P,h,w,C) -> padded-flat fp16 with TWO planes per row: columns [hi | lo], c_pad each.
__global__ void split2_pad_kernel(const float* __restrict__ x, int P, int h, int w, int C, int c_pad,
                                  const uint32_t* __restrict__ amax, __half* __restrict__ out, int item0) {
  // one grid row (blockIdx.y) per item, 32-bit index arithmetic inside it
  const int wp = w + 1, rp = (h + 1) * wp, cg = c_pad / 8;
  const int item = item0 + blockIdx.y;
  (void)P;
  const unsigned per_item = (unsigned)rp * (unsigned)cg;
  const float s = pow2_scale_from_amax(__ldg(amax + item));        // this item's own scale
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < per_item; i += gridDim.x * blockDim.x) {
    const int g = (int)(i % (unsigned)cg);
    const int q = (int)(i / (unsigned)cg);
    const size_t r = (size_t)item * rp + q;
    const int y = q / wp, xx = q - y * wp;
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = 0.f;
    if (y < h && xx < w) {
      const float* src = x + (((size_t)item * h + y) * w + xx) * C + g * 8;
      if ((C & 3) == 0 && g * 8 + 8 <= C) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src + 4));
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (g * 8 + k < C) v[k] = __ldg(src + k);
      }
    }
    __half hh[8], ll[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) split2h(v[k] * s, hh[k], ll[k]);
    __half* o = out + r * (2 * (size_t)c_pad) + g * 8;
    *reinterpret_cast<uint4*>(o) = *reinterpret_cast<const uint4*>(hh);
    *reinterpret_cast<uint4*>(o + c_pad) = *reinterpret_cast<const uint4*>(ll);
  }
}

// The same with the switch-driven unpool (DPooling.down) in front: x is the POOLED fp32 tensor (P,h/2,w/2,C), sw its
// switches (Bsw,h/2,w/2,C); the unpooled fp32 tensor (4x the elements, written once and read once before) never exists.
__global__ void split2_unpool_pad_kernel(const float* __restrict__ x, const uint8_t* __restrict__ sw, int Bsw, int P, int h, int w,
                                         int C, int c_pad, const uint32_t* __restrict__ amax, __half* __restrict__ out, int item0) {
  // one grid row (blockIdx.y) per item, 32-bit index arithmetic inside it (the 64-bit divisions of a flat index over all
  // items cost as much as the split itself: 4.0 -> 4.5 TB/s of algorithmic bytes here, 5.5 -> 6.1 for the plain split)
  const int wp = w + 1, rp = (h + 1) * wp, cg = c_pad / 8;
  const int ph = h / 2, pw = w / 2;
  const int item = item0 + blockIdx.y;
  (void)P;
  const unsigned per_item = (unsigned)rp * (unsigned)cg;
  const float s = pow2_scale_from_amax(__ldg(amax + item));
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < per_item; i += gridDim.x * blockDim.x) {
    const int g = (int)(i % (unsigned)cg);
    const int q = (int)(i / (unsigned)cg);
    const size_t r = (size_t)item * rp + q;
    const int y = q / wp, xx = q - y * wp;
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = 0.f;
    if (y < h && xx < w) {
      const size_t pix = ((size_t)(y >> 1) * pw + (xx >> 1)) * C + g * 8;
      const float* src = x + (size_t)item * ph * pw * C + pix;
      const uint8_t* sp = sw + (size_t)(item % Bsw) * ph * pw * C + pix;
      const uint32_t corner = (uint32_t)((y & 1) * 2 + (xx & 1));
      if ((C & 7) == 0 && g * 8 + 8 <= C) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src + 4));
        const uint2 s2 = __ldg(reinterpret_cast<const uint2*>(sp));
        const float t[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t sb = ((k < 4 ? s2.x : s2.y) >> (8 * (k & 3))) & 0xFFu;
          v[k] = sb == corner ? t[k] : 0.f;
        }
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (g * 8 + k < C && __ldg(sp + k) == corner) v[k] = __ldg(src + k);
      }
    }
    __half hh[8], ll[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) split2h(v[k] * s, hh[k], ll[k]);
    __half* o = out + r * (2 * (size_t)c_pad) + g * 8;
    *reinterpret_cast<uint4*>(o) = *reinterpret_cast<const uint4*>(hh);
    *reinterpret_cast<uint4*>(o + c_pad) = *reinterpret_cast<const uint4*>(ll);
  }
}

// Weights for the two-plane scheme: w fp32 [9][K][N] -> fp16 [tap][plane (hi,lo)][n_pad][k_pad], scaled like the activations.
__global__ void pack_weights_2h_kernel(const float* __restrict__ w, int kdim, int ndim, int k_pad, int n_pad,
                                       const uint32_t* __restrict__ amax, __half* __restrict__ out) {
  const float s = pow2_scale_from_amax(__ldg(amax));
  const size_t plane = (size_t)n_pad * k_pad;
  const size_t total = (size_t)9 * plane;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int k = (int)(i % k_pad);
    const int n = (int)((i / k_pad) % n_pad);
    const int t = (int)(i / plane);
    float v = 0.f;
    if (k < kdim && n < ndim) v = __ldg(w + ((size_t)t * kdim + k) * ndim + n);
    __half h, l;
    split2h(v * s, h, l);
    const size_t o = ((size_t)t * 2) * plane + (size_t)n * k_pad + k;
    out[o] = h; out[o + plane] = l;
  }
}

// the all-tap weights of a narrow last layer in the same form: out [2][32][k_pad], row = tap*N + c
__global__ void pack_weights_alltap2h_kernel(const float* __restrict__ wd, int kdim, int ndim, int k_pad,
                                             const uint32_t* __restrict__ amax, __half* __restrict__ out) {
  const float s = pow2_scale_from_amax(__ldg(amax));
  const int total = 32 * k_pad;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int k = i % k_pad, row = i / k_pad;
    float v = 0.f;
    if (row < 9 * ndim && k < kdim) {
      const int t = row / ndim, c = row % ndim;
      v = __ldg(wd + ((size_t)t * kdim + k) * ndim + c);
    }
    __half h, l;
    split2h(v * s, h, l);
    out[i] = h; out[total + i] = l;
  }
}

// Forward weights for the exact path: w fp32 [9][K=cin][N=cout] -> bf16 [tap][plane (hi,mid,lo)][n_pad][k_pad].
__global__ void pack_weights_fwd3_kernel(const float* __restrict__ w, int kdim, int ndim, int k_pad, int n_pad,
                                         __nv_bfloat16* __restrict__ out) {
  const size_t plane = (size_t)n_pad * k_pad;
  const size_t total = (size_t)9 * plane;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int k = (int)(i % k_pad);
    const int n = (int)((i / k_pad) % n_pad);
    const int t = (int)(i / plane);
    float v = 0.f;
    if (k < kdim && n < ndim) v = __ldg(w + ((size_t)t * kdim + k) * ndim + n);
    __nv_bfloat16 h, m, l;
    split3(v, h, m, l);
    const size_t o = ((size_t)t * 3) * plane + (size_t)n * k_pad + k;
    out[o] = h; out[o + plane] = m; out[o + 2 * plane] = l;
  }
}

// First-layer im2col: fp32 image (P,h,w,C), C <= 4 -> rows of 64 bf16 in the padded-flat row order:
// columns [0, 9C) = bf16(x) of the 3x3 neighbourhood (tap-major, channel-minor), [9C, 18C) = bf16(x - hi), rest zero.
// With weights [n][64] holding bf16(w) under both halves, one K=64 GEMM gives sum (x_hi + x_lo) * w: the image keeps
// 16 mantissa bits, the weights are bf16 like everywhere else in the bf16 chain.  One thread per padded row.
template <int C>
__global__ void __launch_bounds__(128)
im2col_first_kernel(const float* __restrict__ x, int P, int h, int w, const int* __restrict__ jit,
                    __nv_bfloat16* __restrict__ out) {
  // one thread per padded row (measured: a 16-byte-chunk-per-thread form with coalesced stores is 3x slower,
  // the index arithmetic dominates).  jit (optional): {counter, oy0, ox0, ...} - the rows are taken from the image
  // rolled by (oy, ox) of iteration `counter` (numpy.roll), i.e. DeepDream's jitter is folded into this gather.
  griddep_wait();
  griddep_launch_dependents();
  int roll_y = 0, roll_x = 0;
  if (jit) { const int c = jit[0]; roll_y = jit[1 + 2 * c] % h; roll_x = jit[2 + 2 * c] % w; if (roll_y < 0) roll_y += h; if (roll_x < 0) roll_x += w; }
  const int wp = w + 1, rp = (h + 1) * wp;
  const size_t total = (size_t)P * rp;
  for (size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x; r < total; r += (size_t)gridDim.x * blockDim.x) {
    const int q = (int)(r % rp);
    const int item = (int)(r / rp);
    const int y = q / wp, xx = q % wp;
    __nv_bfloat16 row[64];
#pragma unroll
    for (int k = 0; k < 64; ++k) row[k] = __float2bfloat16(0.f);
    if (y < h && xx < w) {
      const float* img = x + (size_t)item * h * w * C;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const int yy = y + t / 3 - 1, xs = xx + t % 3 - 1;
        if (yy >= 0 && yy < h && xs >= 0 && xs < w) {
          int sy = yy - roll_y; if (sy < 0) sy += h;          // rolled[yy][xs] = img[(yy - oy) mod h][(xs - ox) mod w]
          int sx = xs - roll_x; if (sx < 0) sx += w;
#pragma unroll
          for (int c = 0; c < C; ++c) {
            const float v = __ldg(img + ((size_t)sy * w + sx) * C + c);
            const __nv_bfloat16 hi = __float2bfloat16(v);
            row[t * C + c] = hi;
            row[9 * C + t * C + c] = __float2bfloat16(v - __bfloat162float(hi));
          }
        }
      }
    }
    uint4* o = reinterpret_cast<uint4*>(out + r * 64);
    const uint4* src = reinterpret_cast<const uint4*>(row);
#pragma unroll
    for (int k = 0; k < 8; ++k) o[k] = src[k];
  }
}

// first-layer weights for the GEMM above: w fp32 [9][C][cout] -> bf16 [n_pad][64]
__global__ void pack_first_kernel(const float* __restrict__ w, int C, int cout, int n_pad, __nv_bfloat16* __restrict__ out) {
  const int total = n_pad * 64;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int k = i % 64, n = i / 64;
    float v = 0.f;
    if (n < cout && k < 18 * C) v = __ldg(w + (size_t)(k % (9 * C)) * cout + n);     // [t][c][n] flattened = (t*C + c)*cout + n
    out[i] = __float2bfloat16(v);
  }
}

// fp32 compact (P,h,w,C) -> bf16 padded-flat (P*(h+1)*(w+1), c_pad); one thread = 8 channels of one padded row.
__global__ void pad_convert_kernel(const float* __restrict__ x, int P, int h, int w, int C, int c_pad,
                                   __nv_bfloat16* __restrict__ out) {
  const int wp = w + 1, rp = (h + 1) * wp, cg = c_pad / 8;
  const size_t total = (size_t)P * rp * cg;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int g = (int)(i % cg);
    const size_t r = i / cg;
    const int q = (int)(r % rp);
    const int item = (int)(r / rp);
    const int y = q / wp, xx = q % wp;
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = 0.f;
    if (y < h && xx < w) {
      const float* src = x + (((size_t)item * h + y) * w + xx) * C + g * 8;
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (g * 8 + k < C) v[k] = __ldg(src + k);
    }
    uint32_t pk[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      __nv_bfloat162 h2 = __floats2bfloat162_rn(v[2 * k], v[2 * k + 1]);
      pk[k] = *reinterpret_cast<uint32_t*>(&h2);
    }
    *reinterpret_cast<uint4*>(out + r * c_pad + g * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
}

// Switch-driven unpool in the padded-flat bf16 layout (DPooling.down / max-pool backward): in
// (P,(ph+1),(pw+1),c_pad) at the pooled size, switches (Bsw, ph, pw, C) bytes, out (P,(h+1),(w+1),c_pad) with
// h in {2ph, 2ph+1}, w likewise (odd sizes only occur in DeepDream: the trailing row/column gets zeros).
// One thread = 8 channels of one 2x2 output quad; quads beyond the pooled grid are zero rows / columns.
__global__ void __launch_bounds__(256)
unpool_pad_kernel(const __nv_bfloat16* __restrict__ x, const uint8_t* __restrict__ sw, int P, int Bsw, int ph, int pw,
                  int h, int w, int C, int c_pad, __nv_bfloat16* __restrict__ out,
                  const __nv_bfloat16* __restrict__ prep_act, float prep_scale, int prep_relu) {
  griddep_wait();                    // no-ops unless launched as a programmatic dependent (launch_k below)
  griddep_launch_dependents();
  const int iwp = pw + 1, irp = (ph + 1) * iwp;
  const int owp = w + 1, orp = (h + 1) * owp;
  const int qh = (h + 2) / 2, qw = (w + 2) / 2;          // quads covering y in [0,h], x in [0,w]
  const int cg = c_pad / 8;
  // grid rows (blockIdx.y) walk the items, 32-bit index arithmetic inside an item (no 64-bit divisions)
  const unsigned per_item = (unsigned)qh * (unsigned)qw * (unsigned)cg;
  for (int item = blockIdx.y; item < P; item += gridDim.y)
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < per_item; i += gridDim.x * blockDim.x) {
    const int g = (int)(i % (unsigned)cg);
    unsigned t = i / (unsigned)cg;
    const int oy = (int)(t / (unsigned)qw);
    const int ox = (int)(t - (unsigned)oy * (unsigned)qw);
    uint4 val = make_uint4(0u, 0u, 0u, 0u);
    uint32_t sidx[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) sidx[k] = 0xFFu;
    if (oy < ph && ox < pw) {
      val = __ldg(reinterpret_cast<const uint4*>(x + ((size_t)item * irp + oy * iwp + ox) * c_pad + g * 8));
      const uint8_t* sp = sw + (((size_t)(item % Bsw) * ph + oy) * pw + ox) * C + g * 8;
      if (g * 8 + 8 <= C && (C & 7) == 0) {
        const uint2 s2 = __ldg(reinterpret_cast<const uint2*>(sp));
#pragma unroll
        for (int k = 0; k < 4; ++k) { sidx[k] = (s2.x >> (8 * k)) & 0xFFu; sidx[4 + k] = (s2.y >> (8 * k)) & 0xFFu; }
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (g * 8 + k < C) sidx[k] = __ldg(sp + k);
      }
    }
    const uint32_t vv[4] = {val.x, val.y, val.z, val.w};
#pragma unroll
    for (int qd = 0; qd < 4; ++qd) {
      const int y = 2 * oy + (qd >> 1), xx = 2 * ox + (qd & 1);
      if (y > h || xx > w) continue;
      uint32_t o[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t lo = (sidx[2 * k] == (uint32_t)qd) ? (vv[k] & 0xFFFFu) : 0u;
        const uint32_t hi = (sidx[2 * k + 1] == (uint32_t)qd) ? (vv[k] & 0xFFFF0000u) : 0u;
        o[k] = lo | hi;
      }
      const size_t oo = ((size_t)item * orp + y * owp + xx) * c_pad + g * 8;
      if (prep_act && y < h && xx < w) {                  // gradient prep of the layer below (pad rows / columns stay zero)
        const uint4 a4 = __ldg(reinterpret_cast<const uint4*>(prep_act + oo));
        const uint32_t aw[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float a0 = __uint_as_float(aw[k] << 16), a1 = __uint_as_float(aw[k] & 0xFFFF0000u);
          float v0 = fmaf(prep_scale, a0, __uint_as_float(o[k] << 16));
          float v1 = fmaf(prep_scale, a1, __uint_as_float(o[k] & 0xFFFF0000u));
          if (prep_relu) {
            if (!(a0 > 0.f)) v0 = 0.f;
            if (!(a1 > 0.f)) v1 = 0.f;
          }
          __nv_bfloat162 h2 = __floats2bfloat162_rn(v0, v1);
          o[k] = *reinterpret_cast<uint32_t*>(&h2);
        }
      }
      *reinterpret_cast<uint4*>(out + oo) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// 2x2/2 max-pool + switch in the padded-flat bf16 layout (DeepDream forward in bf16): in (P,(h+1),(w+1),c_pad),
// out (P,(ph+1),(pw+1),c_pad) with ph = h/2, pw = w/2 (floor), switches (P,ph,pw,C).  Same tie rule as the fp32
// kernel (strict '>' in row-major order).  One thread = 8 channels of one pooled position (incl. the zero pads).
__global__ void __launch_bounds__(256)
pool_pad_kernel(const __nv_bfloat16* __restrict__ x, int P, int h, int w, int C, int c_pad,
                __nv_bfloat16* __restrict__ out, uint8_t* __restrict__ sw) {
  griddep_wait();
  griddep_launch_dependents();
  const int ph = h / 2, pw = w / 2;
  const int iwp = w + 1, irp = (h + 1) * iwp;
  const int owp = pw + 1, orp = (ph + 1) * owp;
  const int cg = c_pad / 8;
  // grid rows (blockIdx.y) walk the items, 32-bit index arithmetic inside an item (no 64-bit divisions)
  const unsigned per_item = (unsigned)orp * (unsigned)cg;
  for (int item = blockIdx.y; item < P; item += gridDim.y)
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < per_item; i += gridDim.x * blockDim.x) {
    const int g = (int)(i % (unsigned)cg);
    const int q = (int)(i / (unsigned)cg);
    const size_t r = (size_t)item * orp + q;
    const int oy = q / owp, ox = q - oy * owp;
    uint4 res = make_uint4(0u, 0u, 0u, 0u);
    if (oy < ph && ox < pw) {
      const __nv_bfloat16* base = x + ((size_t)item * irp + (2 * oy) * iwp + 2 * ox) * c_pad + g * 8;
      uint4 q4[4];
      q4[0] = __ldg(reinterpret_cast<const uint4*>(base));
      q4[1] = __ldg(reinterpret_cast<const uint4*>(base + c_pad));
      q4[2] = __ldg(reinterpret_cast<const uint4*>(base + (size_t)iwp * c_pad));
      q4[3] = __ldg(reinterpret_cast<const uint4*>(base + (size_t)iwp * c_pad + c_pad));
      uint16_t best[8];
      uint8_t idx[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float bv = 0.f;
#pragma unroll
        for (int qd = 0; qd < 4; ++qd) {
          const uint32_t word = (&q4[qd].x)[k >> 1];
          const uint16_t bits = (k & 1) ? (uint16_t)(word >> 16) : (uint16_t)(word & 0xFFFFu);
          const float v = __uint_as_float((uint32_t)bits << 16);
          if (qd == 0 || v > bv) { bv = v; best[k] = bits; idx[k] = (uint8_t)qd; }
        }
      }
      res = make_uint4((uint32_t)best[0] | ((uint32_t)best[1] << 16), (uint32_t)best[2] | ((uint32_t)best[3] << 16),
                       (uint32_t)best[4] | ((uint32_t)best[5] << 16), (uint32_t)best[6] | ((uint32_t)best[7] << 16));
      uint8_t* sp = sw + (((size_t)item * ph + oy) * pw + ox) * C + g * 8;
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (g * 8 + k < C) sp[k] = idx[k];
    }
    *reinterpret_cast<uint4*>(out + r * c_pad + g * 8) = res;
  }
}

// DeepDream backward, padded-flat bf16: g = mask(act > 0) * (g_above + scale * act); g_above may be absent.
__global__ void dream_prep_pad_kernel(__nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ act, size_t n2,
                                      float scale, int has_g, int relu) {
  griddep_wait();
  griddep_launch_dependents();
  __nv_bfloat162* g2 = reinterpret_cast<__nv_bfloat162*>(g);
  const __nv_bfloat162* a2 = reinterpret_cast<const __nv_bfloat162*>(act);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (size_t)gridDim.x * blockDim.x) {
    const float2 a = __bfloat1622float2(a2[i]);
    float2 v = has_g ? __bfloat1622float2(g2[i]) : make_float2(0.f, 0.f);
    v.x += scale * a.x; v.y += scale * a.y;
    if (relu) { if (!(a.x > 0.f)) v.x = 0.f; if (!(a.y > 0.f)) v.y = 0.f; }
    g2[i] = __floats2bfloat162_rn(v.x, v.y);
  }
}

// fp64 block partials of sum(x^2) over a bf16 array (pads are zero, so padded-flat tensors sum correctly)
__global__ void __launch_bounds__(256) sumsq_bf16_kernel(const __nv_bfloat16* __restrict__ x, size_t n2, double* __restrict__ partial) {
  const __nv_bfloat162* x2 = reinterpret_cast<const __nv_bfloat162*>(x);
  double q = 0.0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (size_t)gridDim.x * blockDim.x) {
    const float2 a = __bfloat1622float2(x2[i]);
    q += (double)a.x * a.x + (double)a.y * a.y;
  }
  __shared__ double rq[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  if ((threadIdx.x & 31) == 0) rq[threadIdx.x >> 5] = q;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += rq[k];
    partial[2 * blockIdx.x] = 0.0;
    partial[2 * blockIdx.x + 1] = t;
  }
}

// Rank-1 first backward conv writing the padded-flat bf16 layout: out[r][ci] = relu(sum_tap map[p+tap] * wd[tap][f][ci]).
// One block = RANK1_PX consecutive padded rows of one item.  A thread owns four channels for the whole block
// (its 36 weights live in registers) and walks over the rows; the 9 map values of a row are shared-memory
// broadcasts.  Stores are 8 bytes, contiguous across the threads of a row.
constexpr int RANK1_PX = 128;
__global__ void __launch_bounds__(256)
rank1_pad_kernel(const float* __restrict__ maps, const float* __restrict__ wd, int cout, int cin, int c_pad,
                 const int* __restrict__ feat, int B, int h, int w, int relu, __nv_bfloat16* __restrict__ out) {
  __shared__ __align__(16) float tap[RANK1_PX * 12];          // 9 taps per row, padded to 12 floats: three 16-byte reads
  const int item = blockIdx.y;
  const int f = feat[item / B];
  const int wp = w + 1, rp = (h + 1) * wp;
  const int q0 = blockIdx.x * RANK1_PX;
  const float* m = maps + (size_t)item * h * w;
  for (int i = threadIdx.x; i < RANK1_PX * 9; i += blockDim.x) {
    const int pp = i / 9, t = i - pp * 9;
    const int q = q0 + pp;
    float v = 0.f;
    if (q < rp) {
      const int yc = q / wp, xc = q - yc * wp;
      const int y = yc + t / 3 - 1, x = xc + t % 3 - 1;
      if (yc < h && xc < w && y >= 0 && y < h && x >= 0 && x < w) v = __ldg(m + y * w + x);
    }
    tap[pp * 12 + t] = v;           // all-zero taps at pad positions -> relu(0) = 0 is written there
  }
  __syncthreads();
  // A thread owns FOUR channels for the whole block (36 weights in registers) and walks over the rows: three 16-byte
  // shared-memory broadcasts, 36 FMAs and one 8-byte store per row (two channels per thread: 62 instructions per 8
  // bytes against 46 now - the kernel is bound by instruction issue, not by its HBM writes).  Same FMA order per output.
  const int cq = c_pad / 4;                         // channel quads (c_pad is a multiple of 64)
  const int nrows = min(RANK1_PX, rp - q0);
  for (int quad = threadIdx.x % cq; quad < cq; quad += (cq > (int)blockDim.x ? blockDim.x : cq)) {
    const int c4 = quad * 4;
    float wv[9][4];
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float* wr = wd + ((size_t)t * cout + f) * cin;
#pragma unroll
      for (int e = 0; e < 4; ++e) wv[t][e] = c4 + e < cin ? __ldg(wr + c4 + e) : 0.f;
    }
    const int lanes = cq >= (int)blockDim.x ? 1 : (int)blockDim.x / cq;
    for (int pp = (cq >= (int)blockDim.x ? 0 : threadIdx.x / cq); pp < nrows; pp += lanes) {
      const float4 t0 = *reinterpret_cast<const float4*>(tap + pp * 12);
      const float4 t1 = *reinterpret_cast<const float4*>(tap + pp * 12 + 4);
      const float t8 = tap[pp * 12 + 8];
      const float tv[9] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w, t8};
      float a[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int e = 0; e < 4; ++e) a[e] = fmaf(tv[t], wv[t][e], a[e]);
      if (relu) {
#pragma unroll
        for (int e = 0; e < 4; ++e) a[e] = fmaxf(a[e], 0.f);
      }
      const __nv_bfloat162 p0 = __floats2bfloat162_rn(a[0], a[1]), p1 = __floats2bfloat162_rn(a[2], a[3]);
      *reinterpret_cast<uint2*>(out + ((size_t)item * rp + q0 + pp) * c_pad + c4) =
          make_uint2(*reinterpret_cast<const uint32_t*>(&p0), *reinterpret_cast<const uint32_t*>(&p1));
    }
    if (cq <= (int)blockDim.x) break;
  }
}

}  // namespace

// =====================================================================================================
// Host side.
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D bf16 row-major matrix (rows x cols), box = (box_rows x 64 cols), 128-byte swizzle, zero OOB fill.
int make_tmap(dcv_ctx* ctx, CUtensorMap* m, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return ctx->fail(DCV_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * sizeof(__nv_bfloat16)};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return ctx->fail(DCV_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

// Padded-flat activation viewed as (cols = C_pad, x = w+1, y = P*(h+1)); box = 64 channels x 16 x 16.
int make_tmap_patch(dcv_ctx* ctx, CUtensorMap* m, const void* base, uint64_t c_pad, uint64_t wp, uint64_t ytot) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return ctx->fail(DCV_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[3] = {c_pad, wp, ytot};
  cuuint64_t strides[2] = {c_pad * sizeof(__nv_bfloat16), wp * c_pad * sizeof(__nv_bfloat16)};
  cuuint32_t box[3] = {64, (cuuint32_t)LAST_H, (cuuint32_t)LAST_H};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return ctx->fail(DCV_ERR_CUDA, "cuTensorMapEncodeTiled (3-D) failed with CUresult %d", (int)r);
  return 0;
}

// Launch on the context's stream; while DeepDream runs (ctx->latency_tiles) as a programmatic dependent of the
// previous kernel: its CTAs may start (prologue, constant loads) while the predecessor drains, and every kernel
// launched through here calls griddep_wait() before it touches anything the predecessor wrote.  DCV_PDL=0: plain.
// `cluster` > 1: the grid is launched as thread-block clusters of that many CTAs (along x).
template <typename... KArgs, typename... Args>
cudaError_t launch_kc(dcv_ctx* ctx, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, int cluster, Args... args) {
  static const int use_pdl = getenv("DCV_PDL") ? atoi(getenv("DCV_PDL")) : 1;     // measured: +2-4 % on DeepDream 512x512
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = ctx->stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
  // not right after a kernel that wrote weights: deconv_tc_kernel fetches its first weight tiles BEFORE griddepcontrol.wait
  if (use_pdl && ctx->latency_tiles && !ctx->pdl_hold) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = (unsigned)cluster; attr[n].val.clusterDim.y = 1; attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  ctx->pdl_hold = 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
template <typename... KArgs, typename... Args>
cudaError_t launch_k(dcv_ctx* ctx, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, Args... args) {
  return launch_kc(ctx, kernel, grid, block, smem, 1, args...);
}

template <int BLOCK_N, int MODE, int SPLIT = 1>
int launch_tc(dcv_ctx* ctx, const CUtensorMap& ta, const CUtensorMap& tb, TcParams p) {
  const int stage_bytes = tc_stage_bytes(MODE, BLOCK_N);
  const int res_bytes = tc_resident_bytes(MODE, BLOCK_N, p.k_blocks);
  const int red_bytes = (SPLIT - 1) * BLOCK_M * BLOCK_N * 4;        // the peers' partial sums (K split over a cluster)
  int stages = (SMEM_BUDGET - res_bytes - red_bytes) / stage_bytes;
  if (stages > 8) stages = 8;
  if (stages < 2) return ctx->fail(DCV_ERR_UNSUPPORTED, "tc kernel: %d resident bytes leave no room for the pipeline", res_bytes);
  p.stages = stages;
  const int smem = res_bytes + stages * stage_bytes + 1024 /*align slack*/ + 256 /*barriers*/ + red_bytes;
  // the attribute is per (kernel instantiation, device); remember the largest value already set
  static int smem_set[64] = {};
  cudaError_t e = cudaSuccess;
  if (smem_set[ctx->device & 63] < smem) {
    e = cudaFuncSetAttribute(deconv_tc_kernel<BLOCK_N, MODE, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaFuncSetAttribute(smem %d): %s", smem, cudaGetErrorString(e));
    smem_set[ctx->device & 63] = smem;
  }
  const int tiles = p.m_tiles * p.n_tiles;
  const int grid = SPLIT * std::min(tiles, ctx->num_sms / SPLIT);
  e = launch_kc(ctx, deconv_tc_kernel<BLOCK_N, MODE, SPLIT>, dim3(grid), dim3(TC1_THREADS), (size_t)smem, SPLIT, ta, tb, p);
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "launch(deconv_tc_kernel<%d,%d>): %s", BLOCK_N, MODE, cudaGetErrorString(e));
  ctx->launches++;
  e = cudaGetLastError();
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "deconv_tc_kernel<%d,%d> launch: %s", BLOCK_N, MODE, cudaGetErrorString(e));
  return 0;
}

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

}  // namespace

struct TcConvArgs {
  const __nv_bfloat16* w;   // [9][n_pad][k_pad]
  int k_pad, n_pad;
  int k_real, n_real;
  int relu;
  const float* bias;
  int nprod = 1;            // 6: exact fp32 emulation (three operand planes, six products)
  const TcUnpool* up = nullptr;
  const TcPrep* prep = nullptr;
  int ntaps = 9;            // 1: un-shifted GEMM (first layer over im2col rows)
  int prof_class = -1;      // profiling class override (default: conv_fwd for nprod 6, conv_tc otherwise)
  int scheme = 0;           // split-operand mode (nprod 6): 0 = three bf16 planes / six products, 1 = two fp16 planes / three
  const uint32_t* amax_a = nullptr;
  const uint32_t* amax_b = nullptr;
  uint32_t* amax_out = nullptr;
  bool allow_ksplit = false;                // forward pass: few-tile layers may split K over a cluster (bits then depend on the batch's tile count, never on how feature maps are chunked)
  uint16_t* split_out = nullptr;            // scheme 1: write the result as the next conv's fp16 planes (TcParams::split_out)
  const uint32_t* split_scale = nullptr;
};

static int tc_conv_run(dcv_ctx* ctx, const TcConvArgs& a, const __nv_bfloat16* in, int P, int h, int w,
                       __nv_bfloat16* out_bf16, float* out_f32);
static int tc_strip_impl(dcv_ctx* ctx, const Layer& conv, const Layer* last, const __nv_bfloat16* in, const uint8_t* sw, int Bsw, int P,
                         int h, int w, __nv_bfloat16* out_bf16, float* out_f32, int relu, bool prof);


int tc_cpad(int c) { return round_up(c, 64); }

size_t tc_padded_elems(int h, int w, int c) { return (size_t)(h + 1) * (w + 1) * tc_cpad(c); }

int tc_prepare_layer(dcv_ctx* ctx, Layer& l) {
  // backward conv: K = l.cout (input channels of the backward pass), N = l.cin
  l.tc.k_pad = tc_cpad(l.cout);
  l.tc.n_pad = tc_cpad(l.cin);
  const size_t n = (size_t)9 * l.tc.n_pad * l.tc.k_pad;
  cudaError_t e = cudaMalloc(&l.tc.wd_bf16, n * sizeof(__nv_bfloat16));
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_NOMEM, "cudaMalloc(bf16 weights): %s", cudaGetErrorString(e));
  pack_weights_tc_kernel<<<grid_for(n, 256), 256, 0, ctx->stream>>>(l.wd, l.cout, l.cin, l.tc.k_pad, l.tc.n_pad, l.tc.wd_bf16);
  ctx->pdl_hold = 1;       // weights written by a kernel: the next conv must not start early (launch_kc)
  ctx->launches++;
  if (9 * l.cin <= LAST_N) {     // narrow last layer: all-tap weights for the GEMM + col2im kernel
    e = cudaMalloc(&l.tc.wall_bf16, (size_t)LAST_N * l.tc.k_pad * sizeof(__nv_bfloat16));
    if (e != cudaSuccess) return ctx->fail(DCV_ERR_NOMEM, "cudaMalloc(all-tap weights): %s", cudaGetErrorString(e));
    pack_weights_alltap_kernel<<<grid_for((size_t)LAST_N * l.tc.k_pad, 256), 256, 0, ctx->stream>>>(l.wd, l.cout, l.cin, l.tc.k_pad, l.tc.wall_bf16);
    ctx->pdl_hold = 1;       // weights written by a kernel: the next conv must not start early (launch_kc)
    ctx->launches++;
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "pack_weights_tc_kernel: %s", cudaGetErrorString(e));
  return 0;
}

void tc_free_layer(Layer& l) {
  if (l.tc.wd_bf16) cudaFree(l.tc.wd_bf16);
  if (l.tc.wf_bf16) cudaFree(l.tc.wf_bf16);
  if (l.tc.wall_bf16) cudaFree(l.tc.wall_bf16);
  if (l.tc.wf3_bf16) cudaFree(l.tc.wf3_bf16);
  if (l.tc.wd3_bf16) cudaFree(l.tc.wd3_bf16);
  if (l.tc.wall3_bf16) cudaFree(l.tc.wall3_bf16);
  l.tc.wall3_bf16 = nullptr;
  if (l.tc.wf0_bf16) cudaFree(l.tc.wf0_bf16);
  l.tc.wf0_bf16 = nullptr;
  if (l.tc.wd2_f16) cudaFree(l.tc.wd2_f16);
  if (l.tc.wall2_f16) cudaFree(l.tc.wall2_f16);
  if (l.tc.wd_amax) cudaFree(l.tc.wd_amax);
  l.tc.wd2_f16 = l.tc.wall2_f16 = nullptr;
  l.tc.wd_amax = nullptr;
  l.tc.wd_bf16 = l.tc.wf_bf16 = l.tc.wall_bf16 = l.tc.wf3_bf16 = l.tc.wd3_bf16 = nullptr;
}

int tc_pad_convert(dcv_ctx* ctx, const float* x, int P, int h, int w, int C, __nv_bfloat16* out) {
  const int c_pad = tc_cpad(C);
  const size_t work = (size_t)P * (h + 1) * (w + 1) * (c_pad / 8);
  ProfScope ps(ctx, DCV_K_OTHER, 0.0, 4.0 * P * h * w * C + 2.0 * P * (h + 1) * (w + 1) * c_pad);
  pad_convert_kernel<<<grid_for(work, 256), 256, 0, ctx->stream>>>(x, P, h, w, C, c_pad, out);
  ctx->launches++;
  return 0;
}

int tc_rank1(dcv_ctx* ctx, const float* maps, const Layer& l, const int* feat, int n, int B, int h, int w,
             __nv_bfloat16* out) {
  const int c_pad = l.tc.n_pad;
  const int rp = (h + 1) * (w + 1);
  dim3 grid((rp + RANK1_PX - 1) / RANK1_PX, n * B);
  ProfScope ps(ctx, DCV_K_RANK1, 2.0 * 9 * n * B * h * w * (double)l.cin, (double)n * B * h * w * (4.0 + 2.0 * l.cin));
  rank1_pad_kernel<<<grid, 256, 0, ctx->stream>>>(maps, l.wd, l.cout, l.cin, c_pad, feat, B, h, w, l.relu, out);
  ctx->launches++;
  return 0;
}

int tc_unpool(dcv_ctx* ctx, const __nv_bfloat16* x, const uint8_t* sw, int P, int Bsw, int ph, int pw, int h, int w,
              int C, __nv_bfloat16* out, const TcPrep* prep) {
  const int c_pad = tc_cpad(C);
  const size_t work = (size_t)P * ((h + 2) / 2) * ((w + 2) / 2) * (c_pad / 8);
  // algorithmic bytes in this layout: 2 B/elem read at pooled size + 1 B switch, 2 B/elem written at full size
  ProfScope ps(ctx, DCV_K_UNPOOL, 0.0, (double)P * ph * pw * C * (2.0 + 1.0 + 8.0));
  const size_t per_item = work / P;
  const unsigned gx = (unsigned)std::min<size_t>((per_item + 255) / 256, (size_t)16 * kNumSMs);
  const unsigned gy = (unsigned)std::min<size_t>((size_t)P, std::max<size_t>(1, (size_t)16 * kNumSMs / gx));
  launch_k(ctx, unpool_pad_kernel, dim3(gx, gy), dim3(256), (size_t)0, x, sw, P, Bsw, ph, pw, h, w, C, c_pad, out,
           prep ? prep->act : (const __nv_bfloat16*)nullptr, prep ? prep->scale : 0.f, prep ? prep->relu : 0);
  ctx->launches++;
  return 0;
}

int tc_pool(dcv_ctx* ctx, const __nv_bfloat16* x, int P, int h, int w, int C, __nv_bfloat16* out, uint8_t* sw) {
  const int c_pad = tc_cpad(C);
  const size_t work = (size_t)P * (h / 2 + 1) * (w / 2 + 1) * (c_pad / 8);
  ProfScope ps(ctx, DCV_K_POOL, 0.0, (double)P * h * w * C * (2.0 + 3.0 / 4.0));
  const size_t per_item = work / P;
  const unsigned gx = (unsigned)std::min<size_t>((per_item + 255) / 256, (size_t)16 * kNumSMs);
  const unsigned gy = (unsigned)std::min<size_t>((size_t)P, std::max<size_t>(1, (size_t)16 * kNumSMs / gx));
  launch_k(ctx, pool_pad_kernel, dim3(gx, gy), dim3(256), (size_t)0, x, P, h, w, C, c_pad, out, sw);
  ctx->launches++;
  return 0;
}

int tc_dream_prep(dcv_ctx* ctx, __nv_bfloat16* g, const __nv_bfloat16* act, size_t elems, float scale, int has_g, int relu) {
  ProfScope ps(ctx, DCV_K_OTHER, 0.0, 6.0 * elems);
  launch_k(ctx, dream_prep_pad_kernel, dim3(grid_for(elems / 2, 256)), dim3(256), (size_t)0, g, act, elems / 2, scale, has_g, relu);
  ctx->launches++;
  return 0;
}

int tc_sumsq_partials(size_t elems) { return grid_for(elems / 2, 256, 4); }

int tc_sumsq(dcv_ctx* ctx, const __nv_bfloat16* x, size_t elems, double* partial, int np) {
  sumsq_bf16_kernel<<<np, 256, 0, ctx->stream>>>(x, elems / 2, partial);
  ctx->launches++;
  return 0;
}

// Exact (fp32-accurate) forward conv on the tensor cores: both operands are split into three bf16 planes and the
// six significant cross products are accumulated in fp32 (x*w = hh + hm + mh + hl + lh + mm, dropped terms ~2^-24).
int tc_prepare_forward_exact(dcv_ctx* ctx, Layer& l) {
  if (l.tc.wf3_bf16) return 0;
  l.tc.fk_pad = tc_cpad(l.cin);
  l.tc.fn_pad = tc_cpad(l.cout);
  const size_t n = (size_t)9 * 3 * l.tc.fn_pad * l.tc.fk_pad;
  cudaError_t e = cudaMalloc(&l.tc.wf3_bf16, n * sizeof(__nv_bfloat16));
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_NOMEM, "cudaMalloc(split forward weights): %s", cudaGetErrorString(e));
  pack_weights_fwd3_kernel<<<grid_for(n / 3, 256), 256, 0, ctx->stream>>>(l.w, l.cin, l.cout, l.tc.fk_pad, l.tc.fn_pad, l.tc.wf3_bf16);
  ctx->pdl_hold = 1;       // weights written by a kernel: the next conv must not start early (launch_kc)
  ctx->launches++;
  return 0;
}

// in: fp32 compact (P,h,w,cin) -> out: fp32 compact (P,h,w,cout) = act(conv + bias); scratch: ctx->ws[4]
// planes_in: -1 = split `in` here; 0 / 1 = the previous conv's epilogue already wrote this conv's three operand planes
// into ws[4 + planes_in].  planes_out (optional): set to the workspace index this conv wrote the NEXT conv's planes to
// (fuse_next: the conv that consumes this result directly, with nothing in between).
int tc_conv_fwd_exact(dcv_ctx* ctx, const Layer& l, const float* in, int P, int h, int w, float* out, int planes_in,
                      const Layer* fuse_next, int* planes_out) {
  if (!l.tc.wf3_bf16) return ctx->fail(DCV_ERR_STATE, "split forward weights were not prepared for this layer");
  static const int fuse = getenv("DCV_FWD_FUSE_SPLIT") ? atoi(getenv("DCV_FWD_FUSE_SPLIT")) : 1;
  if (planes_out) *planes_out = -1;
  const int sel = planes_in >= 0 ? planes_in : 0;
  int r = 0;
  if (planes_in < 0) {
    const size_t elems = (size_t)P * (h + 1) * (w + 1) * 3 * l.tc.fk_pad;
    r = ctx->ws[4].ensure(ctx, elems * sizeof(__nv_bfloat16));
    if (r) return r;
    const size_t work = (size_t)P * (h + 1) * (w + 1) * (l.tc.fk_pad / 8);
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 4.0 * P * h * w * l.cin + 6.0 * P * (h + 1) * (w + 1) * l.tc.fk_pad);
    split3_pad_kernel<<<grid_for(work, 256), 256, 0, ctx->stream>>>(in, P, h, w, l.cin, l.tc.fk_pad, (__nv_bfloat16*)ctx->ws[4].p);
    ctx->launches++;
  }
  TcConvArgs a{l.tc.wf3_bf16, l.tc.fk_pad, l.tc.fn_pad, l.cin, l.cout, l.relu, l.bias, 6};
  a.allow_ksplit = true;
  if (fuse && fuse_next && planes_out && tc_cpad(fuse_next->cin) == l.tc.fn_pad && fuse_next->cin == l.cout) {
    r = ctx->ws[4 + (sel ^ 1)].ensure(ctx, (size_t)P * (h + 1) * (w + 1) * 3 * l.tc.fn_pad * sizeof(__nv_bfloat16));
    if (r) return r;
    a.split_out = (uint16_t*)ctx->ws[4 + (sel ^ 1)].p;
    *planes_out = sel ^ 1;
  }
  return tc_conv_run(ctx, a, (const __nv_bfloat16*)ctx->ws[4 + sel].p, P, h, w, nullptr, out);
}

size_t tc_split_elems(int h, int w, int c) { return (size_t)(h + 1) * (w + 1) * 3 * tc_cpad(c); }

// Backward conv of the fp32 projection mode on the tensor cores, same six-product scheme.  l.wd is the flipped,
// transposed fp32 kernel [9][cout][cin] = [tap][K][N], i.e. the layout pack_weights_fwd3_kernel expects.
// DCV_FP32_SCHEME: 1 (default) = two fp16 planes, three products; 0 = three bf16 planes, six products (the forward's scheme)
static int fp32_scheme() {
  static const int s = getenv("DCV_FP32_SCHEME") ? atoi(getenv("DCV_FP32_SCHEME")) : 1;
  return s;
}

static int tc_prepare_backward_2h(dcv_ctx* ctx, Layer& l) {
  if (l.tc.wd2_f16) return 0;
  if (!l.wd) return ctx->fail(DCV_ERR_STATE, "backward weights were not prepared for this layer");
  const int k_pad = tc_cpad(l.cout), n_pad = tc_cpad(l.cin);
  const size_t nw = (size_t)9 * l.cin * l.cout;
  cudaError_t e = cudaMalloc(&l.tc.wd_amax, 256);
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_NOMEM, "cudaMalloc(weight maximum): %s", cudaGetErrorString(e));
  cudaMemsetAsync(l.tc.wd_amax, 0, 256, ctx->stream);
  absmax_kernel<<<grid_for(nw / 4 + 1, 256, 4), 256, 0, ctx->stream>>>(l.wd, nw, l.tc.wd_amax);
  weight_l1max_kernel<<<l.cin, 256, 0, ctx->stream>>>(l.wd, l.cout, l.cin, l.tc.wd_amax + 1);       // second word of the same block
  ctx->launches += 2;
  const size_t n = (size_t)9 * 2 * n_pad * k_pad;
  e = cudaMalloc(&l.tc.wd2_f16, n * sizeof(__half));
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_NOMEM, "cudaMalloc(two-plane backward weights): %s", cudaGetErrorString(e));
  pack_weights_2h_kernel<<<grid_for(n / 2, 256), 256, 0, ctx->stream>>>(l.wd, l.cout, l.cin, k_pad, n_pad, l.tc.wd_amax, (__half*)l.tc.wd2_f16);
  ctx->pdl_hold = 1;       // weights written by a kernel: the next conv must not start early (launch_kc)
  ctx->launches++;
  if (9 * l.cin <= LAST_N) {
    e = cudaMalloc(&l.tc.wall2_f16, (size_t)2 * LAST_N * k_pad * sizeof(__half));
    if (e != cudaSuccess) return ctx->fail(DCV_ERR_NOMEM, "cudaMalloc(two-plane all-tap weights): %s", cudaGetErrorString(e));
    pack_weights_alltap2h_kernel<<<grid_for((size_t)LAST_N * k_pad, 256), 256, 0, ctx->stream>>>(l.wd, l.cout, l.cin, k_pad, l.tc.wd_amax, (__half*)l.tc.wall2_f16);
    ctx->pdl_hold = 1;       // weights written by a kernel: the next conv must not start early (launch_kc)
    ctx->launches++;
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "two-plane weight packing: %s", cudaGetErrorString(e));
  return 0;
}

int tc_prepare_backward_exact(dcv_ctx* ctx, Layer& l) {
  if (fp32_scheme() == 1) return tc_prepare_backward_2h(ctx, l);
  if (l.tc.wd3_bf16) return 0;
  if (!l.wd) return ctx->fail(DCV_ERR_STATE, "backward weights were not prepared for this layer");
  const int k_pad = tc_cpad(l.cout), n_pad = tc_cpad(l.cin);
  const size_t n = (size_t)9 * 3 * n_pad * k_pad;
  cudaError_t e = cudaMalloc(&l.tc.wd3_bf16, n * sizeof(__nv_bfloat16));
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_NOMEM, "cudaMalloc(split backward weights): %s", cudaGetErrorString(e));
  pack_weights_fwd3_kernel<<<grid_for(n / 3, 256), 256, 0, ctx->stream>>>(l.wd, l.cout, l.cin, k_pad, n_pad, l.tc.wd3_bf16);
  ctx->pdl_hold = 1;       // weights written by a kernel: the next conv must not start early (launch_kc)
  ctx->launches++;
  if (9 * l.cin <= LAST_N) {          // narrow last layer: all-tap weights for the exact GEMM + col2im kernel
    e = cudaMalloc(&l.tc.wall3_bf16, (size_t)3 * LAST_N * k_pad * sizeof(__nv_bfloat16));
    if (e != cudaSuccess) return ctx->fail(DCV_ERR_NOMEM, "cudaMalloc(split all-tap weights): %s", cudaGetErrorString(e));
    pack_weights_alltap3_kernel<<<grid_for((size_t)LAST_N * k_pad, 256), 256, 0, ctx->stream>>>(l.wd, l.cout, l.cin, k_pad, l.tc.wall3_bf16);
    ctx->pdl_hold = 1;       // weights written by a kernel: the next conv must not start early (launch_kc)
    ctx->launches++;
  }
  return 0;
}

// fp16 two-plane form of the backward conv: absmax -> scaled split -> conv_exact_kernel<.., 1> (or the GEMM + col2im last layer)
// chain: the per-item maxima of `in` were left by the conv that produced it (TcParams::amax_out; an unpool or a ReLU in
// between does not raise them) - the extra pass over the input is skipped
// up_sw != null: `in` is the pooled tensor (P,h/2,w/2,cout) and the unpool by these switches is done by the split pass
static int tc_conv_bwd_2h(dcv_ctx* ctx, const Layer& l, const float* in, int P, int h, int w, float* out, int relu, bool chain,
                          const uint8_t* up_sw = nullptr, int up_Bsw = 1, const Layer* fuse_next = nullptr) {
  if (!l.tc.wd2_f16) return ctx->fail(DCV_ERR_STATE, "two-plane backward weights were not prepared for this layer");
  const int k_pad = tc_cpad(l.cout), n_pad = tc_cpad(l.cin);
  // Per-item arrays, four of them: E[0], E[1] = exact maxima of |tensor| (read for this conv's input / filled by its
  // epilogue for its output), S[0], S[1] = scale sources of fused splits (the bound this conv's input planes were scaled
  // by when the previous conv wrote them / the bound this conv scales its output planes by).
  const size_t slot = ((size_t)P * sizeof(uint32_t) + 255) & ~(size_t)255;
  const bool grew = ctx->amax.cap < 4 * slot;
  int r = ctx->amax.ensure(ctx, 4 * slot);
  if (r) return r;
  // input planes already written by the previous conv of the chain (fused split)?
  const int in_planes = (chain && !grew && !up_sw && ctx->amax_cur >= 0) ? ctx->split_ready : -1;
  ctx->split_ready = -1;
  if (grew || !chain || ctx->amax_cur < 0) { chain = false; ctx->amax_cur = 0; }
  auto arr = [&](int i) { return (uint32_t*)((char*)ctx->amax.p + (size_t)i * slot); };
  uint32_t* amax = arr(ctx->amax_cur);                 // E of the input
  uint32_t* amax_next = arr(ctx->amax_cur ^ 1);        // E of the output
  const uint32_t* scale_in = in_planes >= 0 ? arr(2 + ctx->scale_cur) : amax;     // what the input planes were scaled by
  uint32_t* scale_out = arr(2 + (in_planes >= 0 ? ctx->scale_cur ^ 1 : 0));
  // plane buffers: ws[4] / ws[5]; the input sits in ws[4 + in_sel], a fused output goes to the other one
  const int in_sel = in_planes >= 0 ? in_planes : 0;
  const size_t elems = (size_t)P * (h + 1) * (w + 1) * 2 * k_pad;
  if (in_planes < 0) { r = ctx->ws[4].ensure(ctx, elems * sizeof(__half)); if (r) return r; }
  const bool fuse = fuse_next && fuse_next->tc.wd2_f16 && tc_cpad(fuse_next->cout) == n_pad && !(l.tc.wall2_f16);
  if (fuse) {
    r = ctx->ws[4 + (in_sel ^ 1)].ensure(ctx, (size_t)P * (h + 1) * (w + 1) * 2 * n_pad * sizeof(__half));
    if (r) return r;
  }
  const __half* planes = (const __half*)ctx->ws[4 + in_sel].p;
  const size_t len = (size_t)(up_sw ? (h / 2) * (w / 2) : h * w) * l.cout, n_in = (size_t)P * len;   // elements actually read
  if (in_planes < 0) {
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, (chain ? 4.0 : 8.0) * n_in + (up_sw ? 1.0 * n_in : 0.0) + 4.0 * P * (h + 1) * (w + 1) * k_pad);
    if (!chain) {
      cudaMemsetAsync(amax, 0, (size_t)P * sizeof(uint32_t), ctx->stream);
      const int bpi = (int)std::max<size_t>(1, std::min<size_t>((len + 8191) / 8192, std::max<size_t>(1, (size_t)16 * kNumSMs / P)));
      absmax_items_kernel<<<dim3(bpi, P), 256, 0, ctx->stream>>>(in, len, amax);
      ctx->launches++;
    }
    const size_t per_item = (size_t)(h + 1) * (w + 1) * (k_pad / 8);
    const unsigned gx = (unsigned)std::min<size_t>((per_item + 255) / 256, 4096);
    for (int i0 = 0; i0 < P; i0 += 65535) {              // gridDim.y <= 65535
      const int np = std::min(P - i0, 65535);
      if (up_sw)
        split2_unpool_pad_kernel<<<dim3(gx, np), 256, 0, ctx->stream>>>(in, up_sw, up_Bsw, P, h, w, l.cout, k_pad, amax, (__half*)ctx->ws[4].p, i0);
      else
        split2_pad_kernel<<<dim3(gx, np), 256, 0, ctx->stream>>>(in, P, h, w, l.cout, k_pad, amax, (__half*)ctx->ws[4].p, i0);
      ctx->launches++;
    }
  }
  cudaMemsetAsync(amax_next, 0, (size_t)P * sizeof(uint32_t), ctx->stream);
  if (fuse) {
    split_bound_kernel<<<(P + 255) / 256, 256, 0, ctx->stream>>>(amax, l.tc.wd_amax + 1, P, scale_out);
    ctx->launches++;
  }
  ctx->amax_cur = -1;          // set again below once a kernel that fills amax_next has been queued
  static const int use_last = getenv("DCV_TC_LAST") ? atoi(getenv("DCV_TC_LAST")) : 1;
  if (l.tc.wall2_f16 && use_last) {
    LastParams p = {};
    p.P = P; p.h = h; p.w = w; p.cout = l.cin; p.k_blocks = k_pad / 64;
    p.tiles_x = (w + LAST_T - 1) / LAST_T; p.tiles_y = (h + LAST_T - 1) / LAST_T;
    const long long tt = (long long)P * p.tiles_x * p.tiles_y;
    if (tt > 0x7FFFFFFFLL || (long long)P * (h + 1) > 0x7FFFFFFFLL) return ctx->fail(DCV_ERR_ARG, "tc_conv: too many tiles");
    p.total_tiles = (int)tt;
    p.relu = relu;
    p.out = out;
    p.amax_a = scale_in; p.amax_b = l.tc.wd_amax;
    CUtensorMap ta, tb;
    r = make_tmap_patch(ctx, &ta, (void*)planes, (uint64_t)2 * k_pad, (uint64_t)(w + 1), (uint64_t)P * (h + 1));
    if (r) return r;
    r = make_tmap(ctx, &tb, l.tc.wall2_f16, (uint64_t)2 * LAST_N, (uint64_t)k_pad, LAST_N);
    if (r) return r;
    const int smem = last_smem_bytes(p.k_blocks, 2);
    if (smem <= 227 * 1024) {
      static int smem_set[64] = {};
      cudaError_t e = cudaSuccess;
      if (smem_set[ctx->device & 63] < smem) {
        e = cudaFuncSetAttribute(deconv_last_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        smem_set[ctx->device & 63] = smem;
      }
      ProfScope ps(ctx, DCV_K_CONV_TC, 2.0 * 9 * P * h * w * (double)l.cin * l.cout, 2.0 * P * h * w * (2.0 * l.cout + 2.0 * l.cin));
      deconv_last_kernel<2><<<std::min(p.total_tiles, ctx->num_sms), LAST_THREADS, smem, ctx->stream>>>(ta, tb, p);
      ctx->launches++;
      e = cudaGetLastError();
      if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "deconv_last_kernel<fp16> launch: %s", cudaGetErrorString(e));
      return 0;
    }
  }
  TcConvArgs a{l.tc.wd2_f16, k_pad, n_pad, l.cout, l.cin, relu, nullptr, 6};
  a.prof_class = DCV_K_CONV_TC;
  a.scheme = 1;
  a.amax_a = scale_in; a.amax_b = l.tc.wd_amax;
  a.amax_out = amax_next;
  if (fuse) { a.split_out = (uint16_t*)ctx->ws[4 + (in_sel ^ 1)].p; a.split_scale = scale_out; }
  r = tc_conv_run(ctx, a, (const __nv_bfloat16*)planes, P, h, w, nullptr, fuse ? nullptr : out);
  if (!r) {
    ctx->amax_cur = (int)(amax_next != (uint32_t*)ctx->amax.p);
    if (fuse) { ctx->split_ready = in_sel ^ 1; ctx->scale_cur = (int)(scale_out != arr(2)); }
  }
  return r;
}

bool tc_bwd_exact_fuses_unpool() { return fp32_scheme() == 1; }
bool tc_bwd_exact_can_fuse_split() { return fp32_scheme() == 1; }

int tc_conv_bwd_exact_unpool(dcv_ctx* ctx, const Layer& l, const float* pooled, const uint8_t* sw, int Bsw, int P, int h, int w,
                             float* out, int relu, bool chain, const Layer* fuse_next) {
  if (fp32_scheme() != 1) return ctx->fail(DCV_ERR_UNSUPPORTED, "fused unpool needs the two-plane scheme");
  if ((h & 1) || (w & 1)) return ctx->fail(DCV_ERR_ARG, "fused unpool needs an even size");
  static const int fuse_split = getenv("DCV_FP32_FUSE_SPLIT") ? atoi(getenv("DCV_FP32_FUSE_SPLIT")) : 1;
  return tc_conv_bwd_2h(ctx, l, pooled, P, h, w, out, relu, chain, sw, Bsw, fuse_split ? fuse_next : nullptr);
}

int tc_conv_bwd_exact(dcv_ctx* ctx, const Layer& l, const float* in, int P, int h, int w, float* out, int relu, bool chain,
                      const Layer* fuse_next) {
  static const int fuse_split = getenv("DCV_FP32_FUSE_SPLIT") ? atoi(getenv("DCV_FP32_FUSE_SPLIT")) : 1;
  if (fp32_scheme() == 1) return tc_conv_bwd_2h(ctx, l, in, P, h, w, out, relu, chain, nullptr, 1, fuse_split ? fuse_next : nullptr);
  if (!l.tc.wd3_bf16) return ctx->fail(DCV_ERR_STATE, "split backward weights were not prepared for this layer");
  const int k_pad = tc_cpad(l.cout), n_pad = tc_cpad(l.cin);
  const size_t elems = (size_t)P * tc_split_elems(h, w, l.cout);
  int r = ctx->ws[4].ensure(ctx, elems * sizeof(__nv_bfloat16));
  if (r) return r;
  {
    const size_t work = (size_t)P * (h + 1) * (w + 1) * (k_pad / 8);
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 4.0 * P * h * w * l.cout + 6.0 * P * (h + 1) * (w + 1) * k_pad);
    split3_pad_kernel<<<grid_for(work, 256), 256, 0, ctx->stream>>>(in, P, h, w, l.cout, k_pad, (__nv_bfloat16*)ctx->ws[4].p);
    ctx->launches++;
  }
  static const int use_last = getenv("DCV_TC_LAST") ? atoi(getenv("DCV_TC_LAST")) : 1;
  if (l.tc.wall3_bf16 && use_last) {
    // narrow last layer: GEMM + col2im with split operands (tc_last.cuh, EXACT)
    LastParams p;
    p.P = P; p.h = h; p.w = w; p.cout = l.cin; p.k_blocks = k_pad / 64;
    p.tiles_x = (w + LAST_T - 1) / LAST_T; p.tiles_y = (h + LAST_T - 1) / LAST_T;
    const long long tt = (long long)P * p.tiles_x * p.tiles_y;
    if (tt > 0x7FFFFFFFLL || (long long)P * (h + 1) > 0x7FFFFFFFLL) return ctx->fail(DCV_ERR_ARG, "tc_conv: too many tiles");
    p.total_tiles = (int)tt;
    p.relu = relu;
    p.out = out;
    CUtensorMap ta, tb;
    r = make_tmap_patch(ctx, &ta, ctx->ws[4].p, (uint64_t)3 * k_pad, (uint64_t)(w + 1), (uint64_t)P * (h + 1));
    if (r) return r;
    r = make_tmap(ctx, &tb, l.tc.wall3_bf16, (uint64_t)3 * LAST_N, (uint64_t)k_pad, LAST_N);
    if (r) return r;
    const int smem = last_smem_bytes(p.k_blocks, 1);
    if (smem <= 227 * 1024) {
      static int smem_set[64] = {};
      cudaError_t e = cudaSuccess;
      if (smem_set[ctx->device & 63] < smem) {
        e = cudaFuncSetAttribute(deconv_last_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        smem_set[ctx->device & 63] = smem;
      }
      ProfScope ps(ctx, DCV_K_CONV_TC, 2.0 * 9 * P * h * w * (double)l.cin * l.cout, 2.0 * P * h * w * (3.0 * l.cout + 2.0 * l.cin));
      deconv_last_kernel<1><<<std::min(p.total_tiles, ctx->num_sms), LAST_THREADS, smem, ctx->stream>>>(ta, tb, p);
      ctx->launches++;
      e = cudaGetLastError();
      if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "deconv_last_kernel<exact> launch: %s", cudaGetErrorString(e));
      return 0;
    }
  }
  TcConvArgs a{l.tc.wd3_bf16, k_pad, n_pad, l.cout, l.cin, relu, nullptr, 6};
  a.prof_class = DCV_K_CONV_TC;
  return tc_conv_run(ctx, a, (const __nv_bfloat16*)ctx->ws[4].p, P, h, w, nullptr, out);
}

bool tc_first_layer_ok(const Layer& l) { return l.kind == LK_CONV && l.cin >= 1 && l.cin <= 3; }

int tc_conv_first(dcv_ctx* ctx, Layer& l, const float* x, int P, int h, int w, __nv_bfloat16* out_bf16, const int* jit) {
  if (!tc_first_layer_ok(l)) return ctx->fail(DCV_ERR_UNSUPPORTED, "tc_conv_first: %d input channels", l.cin);
  const int n_pad = tc_cpad(l.cout);
  if (!l.tc.wf0_bf16) {
    cudaError_t e = cudaMalloc(&l.tc.wf0_bf16, (size_t)n_pad * 64 * sizeof(__nv_bfloat16));
    if (e != cudaSuccess) return ctx->fail(DCV_ERR_NOMEM, "cudaMalloc(first-layer weights): %s", cudaGetErrorString(e));
    pack_first_kernel<<<grid_for((size_t)n_pad * 64, 256), 256, 0, ctx->stream>>>(l.w, l.cin, l.cout, n_pad, l.tc.wf0_bf16);
    ctx->pdl_hold = 1;       // weights written by a kernel: the next conv must not start early (launch_kc)
    ctx->launches++;
  }
  const size_t rows = (size_t)P * (h + 1) * (w + 1);
  int r = ctx->ws[4].ensure(ctx, rows * 64 * sizeof(__nv_bfloat16));
  if (r) return r;
  __nv_bfloat16* cols = (__nv_bfloat16*)ctx->ws[4].p;
  {
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 4.0 * P * h * w * l.cin + 128.0 * rows);
    const int grid = grid_for(rows, 128);
    switch (l.cin) {
      case 1: launch_k(ctx, im2col_first_kernel<1>, dim3(grid), dim3(128), (size_t)0, x, P, h, w, jit, cols); break;
      case 2: launch_k(ctx, im2col_first_kernel<2>, dim3(grid), dim3(128), (size_t)0, x, P, h, w, jit, cols); break;
      default: launch_k(ctx, im2col_first_kernel<3>, dim3(grid), dim3(128), (size_t)0, x, P, h, w, jit, cols); break;
    }
    ctx->launches++;
  }
  TcConvArgs a{l.tc.wf0_bf16, 64, n_pad, l.cin, l.cout, l.relu, l.bias};
  a.ntaps = 1;
  return tc_conv_run(ctx, a, cols, P, h, w, out_bf16, nullptr);
}

int tc_prepare_forward(dcv_ctx* ctx, Layer& l) {
  if (l.tc.wf_bf16) return 0;
  l.tc.fk_pad = tc_cpad(l.cin);
  l.tc.fn_pad = tc_cpad(l.cout);
  const size_t n = (size_t)9 * l.tc.fn_pad * l.tc.fk_pad;
  cudaError_t e = cudaMalloc(&l.tc.wf_bf16, n * sizeof(__nv_bfloat16));
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_NOMEM, "cudaMalloc(bf16 forward weights): %s", cudaGetErrorString(e));
  // forward conv: K = cin, N = cout; l.w is [9][cin][cout] = [tap][k][n]
  pack_weights_tc_kernel<<<grid_for(n, 256), 256, 0, ctx->stream>>>(l.w, l.cin, l.cout, l.tc.fk_pad, l.tc.fn_pad, l.tc.wf_bf16);
  ctx->pdl_hold = 1;       // weights written by a kernel: the next conv must not start early (launch_kc)
  ctx->launches++;
  return 0;
}

int tc_conv(dcv_ctx* ctx, const Layer& l, const __nv_bfloat16* in, int P, int h, int w, __nv_bfloat16* out_bf16,
            float* out_f32, int relu, const TcUnpool* up, const TcPrep* prep) {
  if (up && !out_bf16) return ctx->fail(DCV_ERR_ARG, "tc_conv: fused unpool needs the bf16 output form");
  if (prep && (!out_bf16 || up)) return ctx->fail(DCV_ERR_ARG, "tc_conv: fused gradient prep needs the plain bf16 output form");
  if (!l.tc.wd_bf16) return ctx->fail(DCV_ERR_STATE, "tensor-core weights were not prepared for this layer");
  static const int use_last = getenv("DCV_TC_LAST") ? atoi(getenv("DCV_TC_LAST")) : 1;
  if (out_f32 && l.tc.wall_bf16 && use_last) {
    // narrow last layer: GEMM + col2im (tc_last.cuh)
    LastParams p;
    p.P = P; p.h = h; p.w = w; p.cout = l.cin; p.k_blocks = l.tc.k_pad / 64;
    p.tiles_x = (w + LAST_T - 1) / LAST_T; p.tiles_y = (h + LAST_T - 1) / LAST_T;
    const long long tt = (long long)P * p.tiles_x * p.tiles_y;
    if (tt > 0x7FFFFFFFLL || (long long)P * (h + 1) > 0x7FFFFFFFLL) return ctx->fail(DCV_ERR_ARG, "tc_conv: too many tiles");
    p.total_tiles = (int)tt;
    p.relu = relu < 0 ? l.relu : relu;
    p.out = out_f32;
    CUtensorMap ta, tb;
    int r = make_tmap_patch(ctx, &ta, in, (uint64_t)l.tc.k_pad, (uint64_t)(w + 1), (uint64_t)P * (h + 1));
    if (r) return r;
    r = make_tmap(ctx, &tb, l.tc.wall_bf16, LAST_N, (uint64_t)l.tc.k_pad, LAST_N);
    if (r) return r;
    const int smem = last_smem_bytes(p.k_blocks);
    if (smem > 227 * 1024) return ctx->fail(DCV_ERR_UNSUPPORTED, "tc_conv: last-layer kernel needs %d bytes of shared memory", smem);
    static int smem_set[64] = {};
    cudaError_t e = cudaSuccess;
    if (smem_set[ctx->device & 63] < smem) {
      e = cudaFuncSetAttribute(deconv_last_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      smem_set[ctx->device & 63] = smem;
    }
    ProfScope ps(ctx, DCV_K_CONV_TC, 2.0 * 9 * P * h * w * (double)l.cin * l.cout, 2.0 * P * h * w * ((double)l.cout + 2.0 * l.cin));
    launch_k(ctx, deconv_last_kernel<0>, dim3(std::min(p.total_tiles, ctx->num_sms)), dim3(LAST_THREADS), (size_t)smem, ta, tb, p);
    ctx->launches++;
    e = cudaGetLastError();
    if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "deconv_last_kernel launch: %s", cudaGetErrorString(e));
    return 0;
  }
  TcConvArgs a{l.tc.wd_bf16, l.tc.k_pad, l.tc.n_pad, l.cout, l.cin, relu < 0 ? l.relu : relu, nullptr};
  a.up = up;
  a.prep = prep;
  return tc_conv_run(ctx, a, in, P, h, w, out_bf16, out_f32);
}

int tc_conv_fwd(dcv_ctx* ctx, const Layer& l, const __nv_bfloat16* in, int P, int h, int w, __nv_bfloat16* out_bf16) {
  if (!l.tc.wf_bf16) return ctx->fail(DCV_ERR_STATE, "tensor-core forward weights were not prepared for this layer");
  TcConvArgs a{l.tc.wf_bf16, l.tc.fk_pad, l.tc.fn_pad, l.cin, l.cout, l.relu, l.bias};
  return tc_conv_run(ctx, a, in, P, h, w, out_bf16, nullptr);
}

static int tc_conv_run(dcv_ctx* ctx, const TcConvArgs& l, const __nv_bfloat16* in, int P, int h, int w,
                       __nv_bfloat16* out_bf16, float* out_f32) {
  const long long rows_ll = (long long)P * (h + 1) * (w + 1);
  if (rows_ll > 0x7FFFFF00LL) return ctx->fail(DCV_ERR_ARG, "tc_conv: %lld padded rows exceed the TMA coordinate range", rows_ll);
  TcParams p = {};
  p.rows = (int)rows_ll; p.h = h; p.w = w;
  p.k_blocks = l.k_pad / 64;
  p.n_pad = l.n_pad; p.n_real = l.n_real;
  p.relu = l.relu;
  p.bias = l.bias;
  p.nprod = l.nprod;
  p.ntaps = l.ntaps;
  p.nprod_planes = l.nprod == 6 ? 3 : 1;
  p.prep_act = l.prep ? l.prep->act : nullptr;
  p.prep_scale = l.prep ? l.prep->scale : 0.f;
  p.prep_relu = l.prep ? l.prep->relu : 0;
  p.up_sw = l.up ? l.up->sw : nullptr;
  p.up_B = l.up ? l.up->B : 0; p.up_C = l.up ? l.up->C : 0; p.up_h = l.up ? l.up->h : 0; p.up_w = l.up ? l.up->w : 0;
  p.out_bf16 = out_bf16; p.out_f32 = out_f32;
  p.m_tiles = (p.rows + BLOCK_M - 1) / BLOCK_M;
  // fp32-accurate mode: accumulation in registers (tc_exact.cuh); DCV_TC_EXACT=2 forces N tile 64
  static const int exact_regs = getenv("DCV_TC_EXACT") ? atoi(getenv("DCV_TC_EXACT")) : 1;
  if (l.nprod == 6) {
    if (!out_f32 && !(l.split_out && l.scheme == 1))   // (scheme 0 writes the planes besides the fp32 result)
      return ctx->fail(DCV_ERR_UNSUPPORTED, "tc_conv: the split-operand mode writes fp32 (or, two-plane scheme, the next conv's planes)");
    int bn = l.n_real <= 16 ? 16 : (l.n_pad % 128 == 0 ? 128 : 64);
    if (bn == 128 && (long long)p.m_tiles * (l.n_pad / 128) < ctx->num_sms) bn = 64;      // few rows: more, smaller tiles
    if (exact_regs == 2 && bn == 128) bn = 64;
    // Forward pass: the layers from 112x112 down have 16 - 200 tiles per image.  Pick the N tile and the K split (clusters
    // of 2 / 4 CTAs per tile) that minimise waves x K steps x cycles per step (85 at N = 64, 92 at N = 128; a split costs
    // about half a 64-channel block for its reduction): 112x112x128 one wave of 100 N = 128 tiles instead of two of N = 64,
    // 56x56x256 52 N = 128 tiles split in two, 28x28x512 56 N = 64 tiles split in two, 14x14x512 16 in four.  The SPLIT is
    // decided on the geometry of ONE image and then used for any batch: it changes the order of the K sum, and an image's
    // tape (its switches) must not depend on how many images share the call.  The N tile does not touch the sums.
    static const int exact_split = getenv("DCV_EXACT_SPLITK") ? atoi(getenv("DCV_EXACT_SPLITK")) : 1;
    int split = 1;
    if (exact_split && l.allow_ksplit && l.scheme == 0 && bn >= 64) {
      auto choose = [&](long long m_tiles, bool with_split, int& obn, int& osplit) {
        double best = 1e30;
        for (int b = 64; b <= 128; b += 64) {
          if (l.n_pad % b) continue;
          const long long tiles = m_tiles * (l.n_pad / b);
          for (int sp = 1; sp <= (with_split ? (b == 64 ? 4 : 2) : 1); sp *= 2) {
            if (p.k_blocks % sp) continue;
            if (sp > 1 && tiles * sp > ctx->num_sms) continue;
            const long long waves = sp > 1 ? 1 : (tiles + ctx->num_sms - 1) / ctx->num_sms;
            const double c = (double)waves * ((double)p.k_blocks / sp + (sp > 1 ? 0.5 : 0.0)) * (b == 64 ? 85.0 : 92.0);
            if (c < best - 1e-9) { best = c; obn = b; osplit = sp; }
          }
        }
      };
      int bn1 = bn, split1 = 1;
      choose(((long long)(h + 1) * (w + 1) + BLOCK_M - 1) / BLOCK_M, true, bn1, split1);
      if (split1 > 1) { bn = bn1; split = split1; }
      else { int s1 = 1; choose(p.m_tiles, false, bn, s1); }
    }
    p.n_tiles = bn == 16 ? 1 : l.n_pad / bn;
    const int red_bytes = (split - 1) * BLOCK_M * bn * 4;
    int stages = (SMEM_BUDGET + (split > 1 ? 24 * 1024 : 0) - red_bytes) / exact_stage_bytes(bn);
    if (stages > 8) stages = 8;
    p.stages = stages;
    p.base_off = 0;
    const int smem = stages * exact_stage_bytes(bn) + 1024 + 256 + red_bytes;
    const int npl = l.scheme == 1 ? 2 : 3;
    p.amax_a = l.amax_a; p.amax_b = l.amax_b; p.amax_out = l.amax_out;
    p.split_out = l.split_out; p.split_scale = l.split_scale;
    CUtensorMap ta, tb;
    int r = make_tmap(ctx, &ta, in, (uint64_t)p.rows, (uint64_t)npl * l.k_pad, BAND_ROWS);
    if (r) return r;
    r = make_tmap(ctx, &tb, l.w, (uint64_t)9 * npl * l.n_pad, (uint64_t)l.k_pad, (uint32_t)bn);
    if (r) return r;
    ProfScope ps(ctx, l.prof_class >= 0 ? l.prof_class : DCV_K_CONV_FWD, 2.0 * 9 * P * h * w * (double)l.k_real * l.n_real,
                 2.0 * P * h * w * ((double)l.k_real + 2.0 * l.n_real));
    const int grid = std::min(p.m_tiles * p.n_tiles, ctx->num_sms);
    cudaError_t e = cudaSuccess;
    if (split > 1) {
      const int tiles = p.m_tiles * p.n_tiles;
      const dim3 g2(split * std::min(tiles, ctx->num_sms / split));
      const int which = bn == 128 ? 2 : (split == 4 ? 1 : 0);
      static int smem_set2[3][64] = {};
      if (smem_set2[which][ctx->device & 63] < smem) {
        e = which == 2 ? cudaFuncSetAttribute(conv_exact_kernel<128, 0, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
          : which == 1 ? cudaFuncSetAttribute(conv_exact_kernel<64, 0, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
                       : cudaFuncSetAttribute(conv_exact_kernel<64, 0, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaFuncSetAttribute(exact kernel, K split): %s", cudaGetErrorString(e));
        smem_set2[which][ctx->device & 63] = smem;
      }
      e = which == 2 ? launch_kc(ctx, conv_exact_kernel<128, 0, 2>, g2, dim3(EXACT_THREADS), (size_t)smem, 2, ta, tb, p)
        : which == 1 ? launch_kc(ctx, conv_exact_kernel<64, 0, 4>, g2, dim3(EXACT_THREADS), (size_t)smem, 4, ta, tb, p)
                     : launch_kc(ctx, conv_exact_kernel<64, 0, 2>, g2, dim3(EXACT_THREADS), (size_t)smem, 2, ta, tb, p);
      if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "conv_exact_kernel (K split %d) launch: %s", split, cudaGetErrorString(e));
      ctx->launches++;
      e = cudaGetLastError();
      if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "conv_exact_kernel (K split %d): %s", split, cudaGetErrorString(e));
      return 0;
    }
#define EXACT_LAUNCH(BN)                                                                                         \
    {                                                                                                              \
      static int smem_set[64] = {};                                                                                \
      if (smem_set[ctx->device & 63] < smem) {                                                                     \
        e = cudaFuncSetAttribute(conv_exact_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);        \
        if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaFuncSetAttribute(exact kernel): %s", cudaGetErrorString(e)); \
        smem_set[ctx->device & 63] = smem;                                                                         \
      }                                                                                                            \
      static int smem_set1[64] = {};                                                                               \
      if (l.scheme == 1 && smem_set1[ctx->device & 63] < smem) {                                                   \
        e = cudaFuncSetAttribute(conv_exact_kernel<BN, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);     \
        if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaFuncSetAttribute(exact kernel): %s", cudaGetErrorString(e)); \
        smem_set1[ctx->device & 63] = smem;                                                                        \
      }                                                                                                            \
      if (l.scheme == 1) conv_exact_kernel<BN, 1><<<grid, EXACT_THREADS, smem, ctx->stream>>>(ta, tb, p);           \
      else conv_exact_kernel<BN, 0><<<grid, EXACT_THREADS, smem, ctx->stream>>>(ta, tb, p);                         \
    }
    switch (bn) {
      case 16: EXACT_LAUNCH(16) break;
      case 64: EXACT_LAUNCH(64) break;
      default: EXACT_LAUNCH(128) break;
    }
#undef EXACT_LAUNCH
    ctx->launches++;
    e = cudaGetLastError();
    if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "conv_exact_kernel<%d> launch: %s", bn, cudaGetErrorString(e));
    return 0;
  }
  int bn;
  if (out_f32 && l.n_real <= 16) bn = 16;
  else if (l.n_pad % 256 == 0) bn = 256;
  else if (l.n_pad % 128 == 0) bn = 128;
  else bn = 64;
  // Few rows (single images at 14x14 .. 129x129; DeepDream sets ctx->latency_tiles, the projection path keeps the
  // fixed rule below so that its bits do not depend on how the feature maps are chunked): pick the N tile that
  // minimises waves x cycles per K step.
  // Measured cycles per 128-row MMA step: ~85 at N=64, ~92 at N=128 (A-operand fetch bound), ~140 at N=256 (single
  // CTA).  A narrower tile gives more CTAs, but pays only if it does not add a wave: 65x65x512 runs one wave of 136
  // N=128 tiles instead of two waves of N=64 tiles, 129x129x256 one wave of 133 N=256 tiles instead of two of N=128.
  static const int bn_model = getenv("DCV_TC_BN_MODEL") ? atoi(getenv("DCV_TC_BN_MODEL")) : 1;
  bool latency_pair = false;      // DeepDream: the rule below chose the CTA-pair kernel, whatever the tile count
  if (bn_model && ctx->latency_tiles) {
    auto cost = [&](int b) {
      const long long tiles = (long long)p.m_tiles * (l.n_pad / b);
      const long long waves = (tiles + ctx->num_sms - 1) / ctx->num_sms;
      return waves * (b == 256 ? 200 : (b == 128 ? 92 : 85));      // N=256 single-CTA: measured slower than two N=128 waves at 129x129
    };
    int best = bn;
    for (int b = bn >> 1; b >= 64 && bn > 64; b >>= 1)
      if (cost(b) < cost(best)) best = b;
    // The CTA-pair kernel (256 x 256 tiles on two SMs, tc_kernel2.cuh; ~146 on this scale per K step of a pair tile) is a
    // candidate for layers of at least two waves of pairs.  Measured with eight epilogue warps per CTA (with four the
    // bias / gradient-prep epilogues of this path set the kernel's pace): one 2048x2048 iteration 6.78 -> 6.07 ms of kernels;
    // a single wave (129x129x256 at 512x512: 67 pairs) is slower than two waves of N = 128 tiles (0.752 -> 0.788 ms per
    // iteration), hence the two-wave rule.  DCV_TC_LATENCY_PAIR=0 switches the candidate off.
    static const int lat_pair = getenv("DCV_TC_LATENCY_PAIR") ? atoi(getenv("DCV_TC_LATENCY_PAIR")) : 1;
    if (lat_pair && bn == 256 && l.ntaps == 9 && l.nprod == 1 && out_bf16) {
      const long long pairs = (long long)((p.rows + 255) / 256) * (l.n_pad / 256);
      const long long pw = (pairs + ctx->num_sms / 2 - 1) / (ctx->num_sms / 2);
      if ((pw >= 2 || lat_pair == 2) && pw * 146 < cost(best)) { best = 256; latency_pair = true; }
    }
    bn = best;
  } else {
    while (bn > 64 && (long long)p.m_tiles * (l.n_pad / bn) < ctx->num_sms) bn >>= 1;
  }
  p.n_tiles = bn == 16 ? 1 : l.n_pad / bn;
  // operand staging mode (tc_kernel.cuh): bands for narrow N, resident weights when they fit beside >= 4 stages
  static const int force_mode = getenv("DCV_TC_MODE") ? atoi(getenv("DCV_TC_MODE")) : -1;
  // Measured on B200: the 128-byte swizzle is applied to the absolute shared-memory address, so an operand
  // that starts 1-2 rows into a TMA box needs base offset 0 (1 gives wrong results; tests/test_tc_gpu.py).
  static const int base_off = getenv("DCV_TC_BASEOFF") ? atoi(getenv("DCV_TC_BASEOFF")) : 0;
  int mode = MODE_TAP;
  if (bn <= 128 && l.ntaps == 9) {
    mode = MODE_BAND;
    if (l.nprod == 1 && p.n_tiles == 1 && tc_resident_bytes(MODE_BAND_RES, bn, p.k_blocks) + 4 * A_BAND_BYTES <= SMEM_BUDGET)
      mode = MODE_BAND_RES;
  }
  if (force_mode >= 0 && bn <= 128 && l.ntaps == 9) mode = force_mode;
  if (mode == MODE_BAND_RES && (p.n_tiles != 1 || l.nprod != 1)) mode = MODE_BAND;
  p.base_off = base_off;
  CUtensorMap ta, tb;
  const int planes = l.nprod == 6 ? 3 : 1;
  int r = make_tmap(ctx, &ta, in, (uint64_t)p.rows, (uint64_t)planes * l.k_pad, mode == MODE_TAP ? BLOCK_M : BAND_ROWS);
  if (r) return r;
  r = make_tmap(ctx, &tb, l.w, (uint64_t)l.ntaps * planes * l.n_pad, (uint64_t)l.k_pad, (uint32_t)bn);
  if (r) return r;
  ProfScope ps(ctx, l.prof_class >= 0 ? l.prof_class : (l.nprod == 6 ? DCV_K_CONV_FWD : DCV_K_CONV_TC),
               2.0 * 9 * P * h * w * (double)l.k_real * l.n_real,
               2.0 * P * h * w * ((double)l.k_real + (out_f32 ? 2.0 * l.n_real : (double)l.n_real)));
  // DeepDream's small layers: too few tiles to fill the GPU -> split K over a cluster of 2 or 4 CTAs per tile
  static const int use_split = getenv("DCV_TC_SPLITK") ? atoi(getenv("DCV_TC_SPLITK")) : 1;
  if (use_split && ctx->latency_tiles && mode == MODE_BAND && l.nprod == 1 && (bn == 64 || bn == 128)) {
    const int tiles = p.m_tiles * p.n_tiles;
    int split = 1;
    if (tiles * 4 <= ctx->num_sms && p.k_blocks % 4 == 0 && p.k_blocks >= 8) split = 4;
    else if (tiles * 2 <= ctx->num_sms && p.k_blocks % 2 == 0 && p.k_blocks >= 4) split = 2;
    if (use_split > 1 && p.k_blocks % use_split == 0 && (use_split == 2 || use_split == 4)) split = use_split;   // forced (tests)
    if (split == 4 && bn == 128) split = 2;      // three 64 KB partial accumulators would leave no room for the pipeline
    if (split == 4) return launch_tc<64, MODE_BAND, 4>(ctx, ta, tb, p);
    if (split == 2) return bn == 64 ? launch_tc<64, MODE_BAND, 2>(ctx, ta, tb, p) : launch_tc<128, MODE_BAND, 2>(ctx, ta, tb, p);
  }
#define TC_DISPATCH(BN)                                                    \
  switch (mode) {                                                          \
    case MODE_BAND: return launch_tc<BN, MODE_BAND>(ctx, ta, tb, p);       \
    case MODE_BAND_RES: return launch_tc<BN, MODE_BAND_RES>(ctx, ta, tb, p); \
    default: return launch_tc<BN, MODE_TAP>(ctx, ta, tb, p);               \
  }
  // Measured (round 1): block1_conv2 2.40 ms vs 2.55-2.9 ms, block2_conv1 0.90 vs 1.07 ms against the band + resident
  // kernels — once the epilogue got 16 warps and per-warp overlapping A segments (no cross-warp exchange).
  static const int use_dxg = getenv("DCV_TC_DXG") ? atoi(getenv("DCV_TC_DXG")) : 1;
  if (use_dxg && l.ntaps == 1 && l.nprod == 1 && !l.up && !l.prep && out_bf16 && l.n_pad == 64) {
    // plain GEMM over un-shifted rows (first layer through im2col): the 16-warp epilogue + TMA store of tc_dxg.cuh
    TcParams q = p;
    q.m_tiles = (p.rows + 127) / 128;
    q.n_tiles = 1;
    const int res_bytes = q.k_blocks * 64 * 128, out_bytes = 128 * 128;
    int stages = (SMEM_BUDGET - res_bytes - out_bytes) / A_TILE_BYTES;
    if (stages > 8) stages = 8;
    if (stages >= 2) {
      q.stages = stages;
      const int smem = res_bytes + stages * A_TILE_BYTES + out_bytes + 1024 + 256;
      CUtensorMap ta2, tb2, to2;
      r = make_tmap(ctx, &ta2, in, (uint64_t)p.rows, (uint64_t)l.k_pad, 32);
      if (r) return r;
      r = make_tmap(ctx, &tb2, l.w, (uint64_t)l.n_pad, (uint64_t)l.k_pad, 64);
      if (r) return r;
      r = make_tmap(ctx, &to2, out_bf16, (uint64_t)p.rows, (uint64_t)l.n_pad, 128);
      if (r) return r;
      const int grid = std::min(q.m_tiles, ctx->num_sms);
      cudaError_t e = cudaSuccess;
#define GEMM64_LAUNCH(EPI)                                                                                             \
      {                                                                                                                \
        static int smem_set[64] = {};                                                                                  \
        if (smem_set[ctx->device & 63] < smem) {                                                                       \
          e = cudaFuncSetAttribute(deconv_dxg_kernel<1, EPI, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);   \
          if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaFuncSetAttribute(gemm64 kernel): %s", cudaGetErrorString(e)); \
          smem_set[ctx->device & 63] = smem;                                                                           \
        }                                                                                                              \
        launch_k(ctx, deconv_dxg_kernel<1, EPI, 0>, dim3(grid), dim3(DXG_THREADS), (size_t)smem, ta2, tb2, to2, q);   \
      }
      if (q.bias) GEMM64_LAUNCH(1) else GEMM64_LAUNCH(0)
#undef GEMM64_LAUNCH
      ctx->launches++;
      e = cudaGetLastError();
      if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "gemm64 kernel launch: %s", cudaGetErrorString(e));
      return 0;
    }
  }
  if (use_dxg && l.ntaps == 9 && l.nprod == 1 && !l.up && out_bf16 && l.n_pad == 64 && (p.rows + DXG_ROWS_OUT - 1) / DXG_ROWS_OUT >= ctx->num_sms) {
    // K = 64 without epilogue extras: the strip form of the same GEMM (one A tile per tile instead of three)
    static const int use_strip = getenv("DCV_TC_STRIP") ? atoi(getenv("DCV_TC_STRIP")) : 1;
    if (use_strip && l.k_pad == 64 && !l.bias && !l.prep) {
      Layer tmp;
      tmp.kind = LK_CONV; tmp.cin = l.n_real; tmp.cout = l.k_real; tmp.relu = l.relu;
      tmp.tc.wd_bf16 = const_cast<__nv_bfloat16*>(l.w); tmp.tc.k_pad = 64; tmp.tc.n_pad = 64;
      return tc_strip_impl(ctx, tmp, nullptr, in, nullptr, 0, P, h, w, out_bf16, nullptr, l.relu, false);
    }
    // 64 output channels: dx-grouped GEMM (tc_dxg.cuh), tiles of 126 rows
    TcParams q = p;
    q.m_tiles = (p.rows + DXG_ROWS_OUT - 1) / DXG_ROWS_OUT;
    q.n_tiles = 1;
    const int xs_bytes = DXG_OUT_BYTES;
    // Resident weights whenever they fit the full 227 KB of shared memory next to >= 4 A stages: for K = 128
    // (block2_conv1, 144 KB of weights) re-loading the 24 KB weight tile with every stage made the L2 -> SM stream the
    // bottleneck (7.8 GB per launch).  DCV_TC_DXG_RES=0 restores the 200 KB rule.
    static const int big_res = getenv("DCV_TC_DXG_RES") ? atoi(getenv("DCV_TC_DXG_RES")) : 1;
    const int budget = big_res ? 232448 - 1024 - 256 : SMEM_BUDGET;
    const int resident = dxg_resident_bytes(1, q.k_blocks) + 4 * A_TILE_BYTES + xs_bytes <= budget ? 1 : 0;
    const int res_bytes = dxg_resident_bytes(resident, q.k_blocks), stage_bytes = dxg_stage_bytes(resident);
    int stages = ((resident ? budget : SMEM_BUDGET) - res_bytes - xs_bytes) / stage_bytes;
    if (stages > 8) stages = 8;
    if (stages >= 2) {
      q.stages = stages;
      const int smem = res_bytes + stages * stage_bytes + xs_bytes + 1024 + 256;
      CUtensorMap ta2, tb2;
      r = make_tmap(ctx, &ta2, in, (uint64_t)p.rows, (uint64_t)l.k_pad, 32);
      if (r) return r;
      r = make_tmap(ctx, &tb2, l.w, (uint64_t)9 * l.n_pad, (uint64_t)l.k_pad, DXG_N);
      if (r) return r;
      CUtensorMap to2;
      r = make_tmap(ctx, &to2, out_bf16, (uint64_t)p.rows, (uint64_t)l.n_pad, DXG_ROWS_OUT);
      if (r) return r;
      cudaError_t e = cudaSuccess;
      const int epi = (q.bias ? 1 : 0) | (q.prep_act ? 2 : 0);
      const int grid = std::min(q.m_tiles, ctx->num_sms);
#define DXG_LAUNCH(RES, EPI)                                                                                          \
      {                                                                                                                \
        static int smem_set[64] = {};                                                                                  \
        if (smem_set[ctx->device & 63] < smem) {                                                                       \
          e = cudaFuncSetAttribute(deconv_dxg_kernel<RES, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);    \
          if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaFuncSetAttribute(dxg kernel): %s", cudaGetErrorString(e)); \
          smem_set[ctx->device & 63] = smem;                                                                           \
        }                                                                                                              \
        launch_k(ctx, deconv_dxg_kernel<RES, EPI, 1>, dim3(grid), dim3(DXG_THREADS), (size_t)smem, ta2, tb2, to2, q);      \
      }
#define DXG_EPIS(RES)                                \
      switch (epi) {                                 \
        case 0: DXG_LAUNCH(RES, 0) break;            \
        case 1: DXG_LAUNCH(RES, 1) break;            \
        case 2: DXG_LAUNCH(RES, 2) break;            \
        default: DXG_LAUNCH(RES, 3) break;           \
      }
      if (resident) { DXG_EPIS(1) } else { DXG_EPIS(0) }
#undef DXG_EPIS
#undef DXG_LAUNCH
      ctx->launches++;
      e = cudaGetLastError();
      if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "deconv_dxg_kernel launch: %s", cudaGetErrorString(e));
      return 0;
    }
  }
  static const int use_pair = getenv("DCV_TC_2CTA") ? atoi(getenv("DCV_TC_2CTA")) : 1;
  const int m_tiles2 = (p.rows + 255) / 256;
  // Measured (profiles/r1_summary.md): the pair form gains 9-13 % at N tile 256 and nothing at N <= 128, where the
  // A-operand read of each SM (unchanged by pairing) sets the pace; DCV_TC_2CTA=2 forces it everywhere for tests.
  if (l.ntaps == 9 && l.nprod == 1 && bn >= 64 && (use_pair == 2 || (use_pair == 1 && bn == 256)) &&
      ((long long)m_tiles2 * p.n_tiles >= ctx->num_sms / 2 || latency_pair)) {
    // CTA-pair kernels (tc_kernel2.cuh): tiles of 256 rows x bn channels, each CTA stages half of the B tile
    TcParams q = p;
    q.m_tiles = m_tiles2;
    int mode2 = mode;
    if (mode2 == MODE_BAND && q.n_tiles == 1 && tc2_resident_bytes(MODE_BAND_RES, bn, q.k_blocks) + 4 * A_BAND_BYTES <= SMEM_BUDGET)
      mode2 = MODE_BAND_RES;       // half the weights per CTA: fits more layers than the single-CTA form
    const int stage_bytes = tc2_stage_bytes(mode2, bn), res_bytes = tc2_resident_bytes(mode2, bn, q.k_blocks);
    int stages = (SMEM_BUDGET - res_bytes) / stage_bytes;
    if (stages > 8) stages = 8;
    q.stages = stages;
    const int smem = res_bytes + stages * stage_bytes + 1024 + 256;
    CUtensorMap tb2;
    r = make_tmap(ctx, &tb2, l.w, (uint64_t)9 * l.n_pad, (uint64_t)l.k_pad, (uint32_t)(bn / 2));
    if (r) return r;
    if (mode2 != mode) {           // A map box differs between tap (128 rows) and band (136 rows) modes
      r = make_tmap(ctx, &ta, in, (uint64_t)p.rows, (uint64_t)l.k_pad, mode2 == MODE_TAP ? BLOCK_M : BAND_ROWS);
      if (r) return r;
    }
    const int tiles = q.m_tiles * q.n_tiles;
    const int pairs = std::min(tiles, ctx->num_sms / 2);
    cudaError_t e = cudaSuccess;
#define TC2_LAUNCH(BN, MD)                                                                                           \
    {                                                                                                                  \
      static int smem_set[64] = {};                                                                                    \
      if (smem_set[ctx->device & 63] < smem) {                                                                         \
        e = cudaFuncSetAttribute(deconv_tc2_kernel<BN, MD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);        \
        if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaFuncSetAttribute(pair kernel): %s", cudaGetErrorString(e)); \
        smem_set[ctx->device & 63] = smem;                                                                             \
      }                                                                                                                \
      deconv_tc2_kernel<BN, MD><<<2 * pairs, NUM_THREADS, smem, ctx->stream>>>(ta, tb2, q);                            \
    }
#define TC2_MODES(BN)                                                     \
    switch (mode2) {                                                      \
      case MODE_BAND: TC2_LAUNCH(BN, MODE_BAND) break;                    \
      case MODE_BAND_RES: TC2_LAUNCH(BN, MODE_BAND_RES) break;            \
      default: TC2_LAUNCH(BN, MODE_TAP) break;                            \
    }
    switch (bn) {
      case 64: TC2_MODES(64) break;
      case 128: TC2_MODES(128) break;
      default: TC2_LAUNCH(256, MODE_TAP) break;
    }
#undef TC2_MODES
#undef TC2_LAUNCH
    ctx->launches++;
    e = cudaGetLastError();
    if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "deconv_tc2_kernel<%d,%d> launch: %s", bn, mode2, cudaGetErrorString(e));
    return 0;
  }
  switch (bn) {
    case 16: TC_DISPATCH(16)
    case 64: TC_DISPATCH(64)
    case 128: TC_DISPATCH(128)
    default: return launch_tc<256, MODE_TAP>(ctx, ta, tb, p);
  }
#undef TC_DISPATCH
}

// ---- strip kernel (tc_strip.cuh): 64 -> 64 backward conv, optionally with the unpool in front of it and the last
// (64 -> <= 3) conv behind it fused in -----------------------------------------------------------------------------
bool tc_strip_conv_ok(const Layer& l) {
  return l.kind == LK_CONV && l.tc.wd_bf16 && l.tc.k_pad == 64 && l.tc.n_pad == 64;
}
bool tc_strip_last_ok(const Layer& conv, const Layer& last) {
  return last.kind == LK_CONV && last.tc.wall_bf16 && last.tc.k_pad == 64 && last.cout == conv.cin && 9 * last.cin <= 27;
}

template <bool UNPOOL, bool LAST>
static int launch_strip(dcv_ctx* ctx, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tbl, const CUtensorMap& to,
                        const StripParams& p) {
  const int smem = strip_smem_bytes(UNPOOL, LAST);
  static int smem_set[64] = {};
  cudaError_t e = cudaSuccess;
  if (smem_set[ctx->device & 63] < smem) {
    e = cudaFuncSetAttribute(deconv_strip_kernel<UNPOOL, LAST>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaFuncSetAttribute(strip kernel, %d bytes): %s", smem, cudaGetErrorString(e));
    smem_set[ctx->device & 63] = smem;
  }
  const int grid = std::min(p.total_strips, ctx->num_sms);
  e = launch_k(ctx, deconv_strip_kernel<UNPOOL, LAST>, dim3(grid), dim3(strip_threads(UNPOOL, LAST)), (size_t)smem, ta, tb, tbl, to, p);
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "launch(deconv_strip_kernel<%d,%d>): %s", (int)UNPOOL, (int)LAST, cudaGetErrorString(e));
  ctx->launches++;
  e = cudaGetLastError();
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "deconv_strip_kernel<%d,%d> launch: %s", (int)UNPOOL, (int)LAST, cudaGetErrorString(e));
  return 0;
}

// in: padded-flat bf16 at (h, w) — or, with sw != null, at the POOLED size (h/2, w/2), unpooled on the fly with the
// switches (Bsw, h/2, w/2, conv.cout).  last == null: out_bf16 = relu?(conv) padded-flat (P,(h+1),(w+1),64);
// last != null: out_f32 = act(last(act(conv))) compact (P, h, w, last->cin).
int tc_strip(dcv_ctx* ctx, const Layer& conv, const Layer* last, const __nv_bfloat16* in, const uint8_t* sw, int Bsw, int P, int h,
             int w, __nv_bfloat16* out_bf16, float* out_f32, int relu) {
  return tc_strip_impl(ctx, conv, last, in, sw, Bsw, P, h, w, out_bf16, out_f32, relu, true);
}

static int tc_strip_impl(dcv_ctx* ctx, const Layer& conv, const Layer* last, const __nv_bfloat16* in, const uint8_t* sw, int Bsw, int P,
                         int h, int w, __nv_bfloat16* out_bf16, float* out_f32, int relu, bool prof) {
  if (!tc_strip_conv_ok(conv)) return ctx->fail(DCV_ERR_UNSUPPORTED, "tc_strip: the layer is not a 64 -> 64 channel conv");
  if (last && !tc_strip_last_ok(conv, *last)) return ctx->fail(DCV_ERR_UNSUPPORTED, "tc_strip: the last layer cannot be fused");
  if (sw && ((h & 1) || (w & 1) || conv.cout != 64)) return ctx->fail(DCV_ERR_UNSUPPORTED, "tc_strip: fused unpool needs even sizes and 64 channels");
  if (last ? !out_f32 : !out_bf16) return ctx->fail(DCV_ERR_ARG, "tc_strip: missing output");
  const long long rows_ll = (long long)P * (h + 1) * (w + 1);
  if (rows_ll > 0x7FFFFF00LL) return ctx->fail(DCV_ERR_ARG, "tc_strip: %lld padded rows exceed the TMA coordinate range", rows_ll);
  StripParams p = {};
  p.rows = (int)rows_ll; p.h = h; p.w = w; p.T = P * (h + 1);
  p.n_cb = w + 1 <= STRIP_POS ? 1 : (w - (STRIP_POS - 1) + STRIP_CB - 1) / STRIP_CB + 1;
  // strip height: tall enough that the two halo rows (four with the fused last layer) cost a few percent, short enough
  // for several strips per SM
  int R = 64;
  while (R > 8 && (long long)p.n_cb * ((p.T + R - 1) / R) < 6LL * ctx->num_sms) R >>= 1;
  static const int force_r = getenv("DCV_TC_STRIP_ROWS") ? atoi(getenv("DCV_TC_STRIP_ROWS")) : 0;
  if (force_r > 0) R = force_r;
  p.R = R;
  p.total_strips = p.n_cb * ((p.T + R - 1) / R);
  p.relu = relu < 0 ? conv.relu : relu;
  p.pooled = sw ? in : nullptr; p.sw = sw; p.Bsw = Bsw; p.C = conv.cout; p.ph = h / 2; p.pw = w / 2;
  p.cout3 = last ? last->cin : 0; p.relu_last = last ? last->relu : 0; p.out = out_f32;
  CUtensorMap ta, tb, tbl, to;
  int r = make_tmap(ctx, &tb, conv.tc.wd_bf16, (uint64_t)9 * 64, 64, STRIP_N);
  if (r) return r;
  ta = tb; tbl = tb; to = tb;
  if (!sw) { r = make_tmap(ctx, &ta, in, (uint64_t)p.rows, 64, 32); if (r) return r; }
  if (last) { r = make_tmap(ctx, &tbl, last->tc.wall_bf16, LAST_N, 64, LAST_N); if (r) return r; }
  else { r = make_tmap(ctx, &to, out_bf16, (uint64_t)p.rows, 64, (uint32_t)std::min(STRIP_POS, w + 1)); if (r) return r; }
  const double px = (double)P * h * w;
  double flops = 2.0 * 9 * px * conv.cin * conv.cout, bytes = 2.0 * px * conv.cout * (sw ? 0.25 + 0.125 : 1.0);
  if (last) { flops += 2.0 * 9 * px * last->cin * last->cout; bytes += 4.0 * px * last->cin; } else bytes += 2.0 * px * conv.cin;
  ProfScope ps(ctx, DCV_K_CONV_TC, flops, bytes, prof);      // prof = false: the caller's scope is already open
  if (sw) return last ? launch_strip<true, true>(ctx, ta, tb, tbl, to, p) : launch_strip<true, false>(ctx, ta, tb, tbl, to, p);
  return last ? launch_strip<false, true>(ctx, ta, tb, tbl, to, p) : launch_strip<false, false>(ctx, ta, tb, tbl, to, p);
}

// Stand-alone operator (dcv_deconv3x3 with DCV_BF16): fp32 compact in -> fp32 compact out.
int tc_deconv(dcv_ctx* ctx, const Layer& l, const float* in, int P, int h, int w, float* out) {
  const size_t elems = (size_t)P * tc_padded_elems(h, w, l.cout);
  int r = ctx->ws[4].ensure(ctx, elems * sizeof(__nv_bfloat16));
  if (r) return r;
  r = tc_pad_convert(ctx, in, P, h, w, l.cout, (__nv_bfloat16*)ctx->ws[4].p);
  if (r) return r;
  return tc_conv(ctx, l, (const __nv_bfloat16*)ctx->ws[4].p, P, h, w, nullptr, out, -1);
}
