// HBM-streaming kernels of the deconvnet path: max-pool with switches, switch-driven unpool,
// per-channel sums, de-process, seed-map handling and the rank-1 first backward conv.
// All of them are bandwidth-bound: 128-bit accesses along the NHWC channel axis, grids sized in
// multiples of the SM count, no shared-memory staging (no reuse to exploit).
#include "common.cuh"
#include "tc_ptx.cuh"

namespace dcv {

static inline int grid_for(size_t work_items, int block, int max_waves = 16) {
  size_t blocks = (work_items + block - 1) / block;
  size_t cap = (size_t)kNumSMs * max_waves;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

// ------------------------------------------------------------------------------------------------
// 2x2/2 max-pool + switch (DPooling.up, deepdream.py:152-188).
// One thread = VEC channels of one pooled pixel.  Strict '>' scan in the order (0,0),(0,1),(1,0),(1,1)
// reproduces "first maximum in row-major order" (deepdream.py:180-183); an all-equal patch gives 0.
// Algorithmic bytes per pooled element: 16 read + 4 written + 1 switch byte.
template <int VEC>
__global__ void __launch_bounds__(256) maxpool_switch_kernel(const float* __restrict__ x, int B, int H, int W, int C,
                                                             float* __restrict__ pooled, uint8_t* __restrict__ sw) {
  const int OH = H >> 1, OW = W >> 1;
  const int CV = C / VEC;
  const size_t total = (size_t)B * OH * OW * CV;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % CV);
    size_t p = i / CV;
    const int ow = (int)(p % OW); p /= OW;
    const int oh = (int)(p % OH);
    const int b = (int)(p / OH);
    const float* base = x + (((size_t)b * H + 2 * oh) * W + 2 * ow) * C + (size_t)cv * VEC;
    float v[4][VEC];
    if constexpr (VEC == 4) {
      const float4 a0 = __ldg(reinterpret_cast<const float4*>(base));
      const float4 a1 = __ldg(reinterpret_cast<const float4*>(base + C));
      const float4 a2 = __ldg(reinterpret_cast<const float4*>(base + (size_t)W * C));
      const float4 a3 = __ldg(reinterpret_cast<const float4*>(base + (size_t)W * C + C));
      v[0][0] = a0.x; v[0][1] = a0.y; v[0][2] = a0.z; v[0][3] = a0.w;
      v[1][0] = a1.x; v[1][1] = a1.y; v[1][2] = a1.z; v[1][3] = a1.w;
      v[2][0] = a2.x; v[2][1] = a2.y; v[2][2] = a2.z; v[2][3] = a2.w;
      v[3][0] = a3.x; v[3][1] = a3.y; v[3][2] = a3.z; v[3][3] = a3.w;
    } else {
#pragma unroll
      for (int k = 0; k < VEC; ++k) {
        v[0][k] = __ldg(base + k);
        v[1][k] = __ldg(base + C + k);
        v[2][k] = __ldg(base + (size_t)W * C + k);
        v[3][k] = __ldg(base + (size_t)W * C + C + k);
      }
    }
    float m[VEC];
    uint8_t s[VEC];
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      float best = v[0][k];
      int idx = 0;
      if (v[1][k] > best) { best = v[1][k]; idx = 1; }
      if (v[2][k] > best) { best = v[2][k]; idx = 2; }
      if (v[3][k] > best) { best = v[3][k]; idx = 3; }
      m[k] = best;
      s[k] = (uint8_t)idx;
    }
    const size_t o = (((size_t)b * OH + oh) * OW + ow) * C + (size_t)cv * VEC;
    if constexpr (VEC == 4) {
      *reinterpret_cast<float4*>(pooled + o) = make_float4(m[0], m[1], m[2], m[3]);
      *reinterpret_cast<uchar4*>(sw + o) = make_uchar4(s[0], s[1], s[2], s[3]);
    } else {
#pragma unroll
      for (int k = 0; k < VEC; ++k) { pooled[o + k] = m[k]; sw[o + k] = s[k]; }
    }
  }
}

void launch_maxpool_switch(const float* x, int B, int H, int W, int C, float* pooled, uint8_t* sw, cudaStream_t s) {
  const size_t outs = (size_t)B * (H / 2) * (W / 2);
  if (outs == 0 || C == 0) return;
  if (C % 4 == 0) {
    maxpool_switch_kernel<4><<<grid_for(outs * (C / 4), 256), 256, 0, s>>>(x, B, H, W, C, pooled, sw);
  } else {
    maxpool_switch_kernel<1><<<grid_for(outs * C, 256), 256, 0, s>>>(x, B, H, W, C, pooled, sw);
  }
}

// ------------------------------------------------------------------------------------------------
// Switch-driven unpool (DPooling.down, deepdream.py:191-209) as a gather: one thread owns VEC
// channels of one pooled pixel and writes its whole 2x2 output quad (value at the switch position,
// zeros elsewhere) with coalesced 128-bit stores, so there are no atomics and no pre-zeroing.
// (H,W) is the OUTPUT size; rows/cols beyond 2*(H/2), 2*(W/2) (odd sizes, DeepDream only) are zeroed.
// `item % Bsw` selects the switch image, so P = n*B projection items can share B switch tensors.
template <int VEC>
__global__ void __launch_bounds__(256) unpool_switch_kernel(const float* __restrict__ x, const uint8_t* __restrict__ sw,
                                                            int P, int Bsw, int H, int W, int C, int relu,
                                                            float* __restrict__ out) {
  const int OH = (H + 1) >> 1, OW = (W + 1) >> 1;   // quads incl. the partial ones at odd borders
  const int PH = H >> 1, PW = W >> 1;               // pooled size
  const int CV = C / VEC;
  const size_t total = (size_t)P * OH * OW * CV;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % CV);
    size_t p = i / CV;
    const int ow = (int)(p % OW); p /= OW;
    const int oh = (int)(p % OH);
    const int item = (int)(p / OH);
    float v[VEC];
    int sidx[VEC];
    const bool inside = (oh < PH) && (ow < PW);
    if (inside) {
      const size_t pi = (((size_t)item * PH + oh) * PW + ow) * C + (size_t)cv * VEC;
      const size_t si = (((size_t)(item % Bsw) * PH + oh) * PW + ow) * C + (size_t)cv * VEC;
      if constexpr (VEC == 4) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(x + pi));
        const uchar4 q = __ldg(reinterpret_cast<const uchar4*>(sw + si));
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
        sidx[0] = q.x; sidx[1] = q.y; sidx[2] = q.z; sidx[3] = q.w;
      } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k) { v[k] = __ldg(x + pi + k); sidx[k] = __ldg(sw + si + k); }
      }
      if (relu) {
#pragma unroll
        for (int k = 0; k < VEC; ++k) v[k] = fmaxf(v[k], 0.f);
      }
    } else {
#pragma unroll
      for (int k = 0; k < VEC; ++k) { v[k] = 0.f; sidx[k] = -1; }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int y = 2 * oh + (q >> 1), xx = 2 * ow + (q & 1);
      if (y >= H || xx >= W) continue;
      const size_t o = (((size_t)item * H + y) * W + xx) * C + (size_t)cv * VEC;
      float r[VEC];
#pragma unroll
      for (int k = 0; k < VEC; ++k) r[k] = (sidx[k] == q) ? v[k] : 0.f;
      if constexpr (VEC == 4) {
        *reinterpret_cast<float4*>(out + o) = make_float4(r[0], r[1], r[2], r[3]);
      } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k) out[o + k] = r[k];
      }
    }
  }
}

// P items, switches of B images (item % B)
void launch_unpool_switch_items(const float* x, const uint8_t* sw, int P, int Bsw, int H, int W, int C, int relu,
                                float* out, cudaStream_t s) {
  const size_t quads = (size_t)P * ((H + 1) / 2) * ((W + 1) / 2);
  if (quads == 0 || C == 0) return;
  if (C % 4 == 0) {
    unpool_switch_kernel<4><<<grid_for(quads * (C / 4), 256), 256, 0, s>>>(x, sw, P, Bsw, H, W, C, relu, out);
  } else {
    unpool_switch_kernel<1><<<grid_for(quads * C, 256), 256, 0, s>>>(x, sw, P, Bsw, H, W, C, relu, out);
  }
}

void launch_unpool_switch(const float* x, const uint8_t* sw, int B, int H, int W, int C, int relu, float* out,
                          cudaStream_t s) {
  launch_unpool_switch_items(x, sw, B, B, H, W, C, relu, out, s);
}

// ------------------------------------------------------------------------------------------------
// Per-channel fp64 sums of a (rows, C) fp32 matrix (find_top_filters, deepdream.py:369-380).
// Stage 1: block (32 x 8) strides over rows; thread x = channel lane (coalesced), 8 row lanes;
// partial[blockIdx.y][c].  Stage 2: one thread per channel adds the partials in fixed order, so
// the result is deterministic.
__global__ void __launch_bounds__(256) channel_sums_stage1(const float* __restrict__ x, size_t rows, int C,
                                                           double* __restrict__ partial) {
  __shared__ double red[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  double acc = 0.0;
  if (c < C) {
    for (size_t r = (size_t)blockIdx.y * 8 + threadIdx.y; r < rows; r += (size_t)gridDim.y * 8)
      acc += (double)__ldg(x + r * C + c);
  }
  red[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k][threadIdx.x];
    partial[(size_t)blockIdx.y * C + c] = t;
  }
}

__global__ void channel_sums_stage2(const double* __restrict__ partial, int npartial, int C, double* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double t = 0.0;
  for (int k = 0; k < npartial; ++k) t += partial[(size_t)k * C + c];
  out[c] = t;
}

int channel_sums_partials(size_t rows, int C) {
  const int xb = (C + 31) / 32;
  size_t want = (size_t)(4 * kNumSMs + xb - 1) / xb;          // ~4 blocks per SM in total
  size_t maxy = (rows + 7) / 8;
  if (want > maxy) want = maxy;
  if (want < 1) want = 1;
  return (int)want;
}

void launch_channel_sums(const float* x, size_t rows, int C, double* partial, int npartial, double* out, cudaStream_t s) {
  dim3 grid((C + 31) / 32, npartial), block(32, 8);
  channel_sums_stage1<<<grid, block, 0, s>>>(x, rows, C, partial);
  channel_sums_stage2<<<(C + 127) / 128, 128, 0, s>>>(partial, npartial, C, out);
}

// ------------------------------------------------------------------------------------------------
// De-process (deprocess_image, deepdream.py:483-498).  Pass 1: fp64 sum and sum of squares (block
// partials in scratch, fixed-order final reduce inside pass 2's prologue).  Pass 2: the elementwise
// map in fp32, op by op in the reference's order.
__global__ void __launch_bounds__(256) deprocess_stats_kernel(const float* __restrict__ x, size_t n,
                                                              double* __restrict__ partial) {
  double s = 0.0, q = 0.0;
  const size_t n4 = n / 4;
  const float4* x4 = reinterpret_cast<const float4*>(x);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  // four independent 16-byte loads in flight per thread; per-load fp32 partials are folded into fp64 at once
  for (; i + 3 * stride < n4; i += 4 * stride) {
    const float4 a = __ldg(x4 + i), b = __ldg(x4 + i + stride), c = __ldg(x4 + i + 2 * stride), d = __ldg(x4 + i + 3 * stride);
    s += ((double)a.x + (double)a.y + (double)a.z + (double)a.w) + ((double)b.x + (double)b.y + (double)b.z + (double)b.w) +
         ((double)c.x + (double)c.y + (double)c.z + (double)c.w) + ((double)d.x + (double)d.y + (double)d.z + (double)d.w);
    q += ((double)a.x * a.x + (double)a.y * a.y + (double)a.z * a.z + (double)a.w * a.w) +
         ((double)b.x * b.x + (double)b.y * b.y + (double)b.z * b.z + (double)b.w * b.w) +
         ((double)c.x * c.x + (double)c.y * c.y + (double)c.z * c.z + (double)c.w * c.w) +
         ((double)d.x * d.x + (double)d.y * d.y + (double)d.z * d.z + (double)d.w * d.w);
  }
  for (; i < n4; i += stride) {
    const float4 a = __ldg(x4 + i);
    s += (double)a.x + (double)a.y + (double)a.z + (double)a.w;
    q += (double)a.x * a.x + (double)a.y * a.y + (double)a.z * a.z + (double)a.w * a.w;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const float a = x[n4 * 4 + threadIdx.x];
    s += a; q += (double)a * a;
  }
  __shared__ double rs[8], rq[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if ((threadIdx.x & 31) == 0) { rs[threadIdx.x >> 5] = s; rq[threadIdx.x >> 5] = q; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ts = 0.0, tq = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) { ts += rs[k]; tq += rq[k]; }
    partial[2 * blockIdx.x] = ts;
    partial[2 * blockIdx.x + 1] = tq;
  }
}

__device__ __forceinline__ float deprocess_norm(float v, float mean, float denom) {
  v = v - mean;            // x -= x.mean()
  v = v / denom;           // x /= (x.std() + epsilon)
  v = __fmul_rn(v, 0.1f);  // x *= 0.1   (separate roundings like numpy: no contraction into one FMA)
  v = __fadd_rn(v, 0.5f);  // x += 0.5
  return v;
}

__device__ __forceinline__ uint8_t deprocess_one(float v, float mean, float denom) {
  v = deprocess_norm(v, mean, denom);
  v = fminf(fmaxf(v, 0.f), 1.f);
  v = v * 255.f;
  v = fminf(fmaxf(v, 0.f), 255.f);
  return (uint8_t)v;       // astype('uint8') truncates
}

__global__ void __launch_bounds__(256) deprocess_map_kernel(const float* __restrict__ x, size_t n,
                                                            const double* __restrict__ partial, int npartial,
                                                            uint8_t* __restrict__ out, float* __restrict__ norm) {
  double ts, tq;
  block_sum_partials(partial, npartial, ts, tq);
  const double dmean = ts / (double)n;
  double var = tq / (double)n - dmean * dmean;      // population variance (ddof = 0)
  if (var < 0.0) var = 0.0;
  // std of the CENTRED data equals std of x; the reference adds K.epsilon() = 1e-7 to it
  const float mean = (float)dmean, denom = (float)(sqrt(var) + 1e-7);
  const size_t n4 = n / 4;
  const float4* x4 = reinterpret_cast<const float4*>(x);
  // Back to front: the statistics pass has just streamed x front to back, so its tail is what the 126 MB L2 still
  // holds - the blocks scheduled first start there (154 MB of input: about half of this pass's reads hit L2).
#pragma unroll 4
  for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n4; k += (size_t)gridDim.x * blockDim.x) {
    const size_t i = n4 - 1 - k;
    const float4 a = __ldg(x4 + i);
    uchar4 r;
    r.x = deprocess_one(a.x, mean, denom);
    r.y = deprocess_one(a.y, mean, denom);
    r.z = deprocess_one(a.z, mean, denom);
    r.w = deprocess_one(a.w, mean, denom);
    reinterpret_cast<uchar4*>(out)[i] = r;
    if (norm)
      reinterpret_cast<float4*>(norm)[i] = make_float4(deprocess_norm(a.x, mean, denom), deprocess_norm(a.y, mean, denom),
                                                       deprocess_norm(a.z, mean, denom), deprocess_norm(a.w, mean, denom));
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const size_t j = n4 * 4 + threadIdx.x;
    out[j] = deprocess_one(x[j], mean, denom);
    if (norm) norm[j] = deprocess_norm(x[j], mean, denom);
  }
}

// Four quotients x / d by one divisor, correctly rounded (the bits of IEEE division) without a division per element:
// with r = RN(1 / d): q = RN(x * r), remainder e = x - q * d (exact in one FMA), q' = RN(q + e * r) - the correction step
// of the compiler's own division sequence, which spends a reciprocal, a Newton step and a range check on every
// element.  The step is exact only while nothing under- or overflows: quotients outside [1e-20, 1e30] (zero, infinities
// and NaNs included - the sum carries them) send all four through the division instruction; the caller passes r = NaN
// when d itself is out of range (1 / d subnormal).  tests/test_ops_gpu.py compares with the dividing form bit for bit.
__device__ __forceinline__ void quot4(const float (&x)[4], float d, float r, float (&q)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float q0 = x[i] * r;
    q[i] = fmaf(fmaf(-q0, d, x[i]), r, q0);
  }
  const float sum = (fabsf(q[0]) + fabsf(q[1])) + (fabsf(q[2]) + fabsf(q[3]));
  const float lo = fminf(fminf(fabsf(q[0]), fabsf(q[1])), fminf(fabsf(q[2]), fabsf(q[3])));
  if (!(sum < 1e30f && lo > 1e-20f)) {
#pragma unroll
    for (int i = 0; i < 4; ++i) q[i] = x[i] / d;
  }
}

// The map pass as a pipeline: a producer warp streams x through a ring of shared-memory stages with bulk async copies
// (cp.async.bulk, completion on an mbarrier), eight consumer warps read 16 bytes per lane from the stage, apply the map
// and store 4 bytes per lane.  The grid-stride form above waits for its own loads (ncu: 61 registers -> 4 blocks per
// SM, 9 warps stalled on the long scoreboard per issued instruction, 3.3 TB/s); here the loads of the next DP_STAGES
// chunks are always in flight.  Same arithmetic (deprocess_one / deprocess_norm).
constexpr int DP_CONS_WARPS = 8;
constexpr int DP_THREADS = 32 * (DP_CONS_WARPS + 1);
constexpr int DP_CHUNK = 2048;                  // floats per stage (8 KB)
constexpr int DP_STAGES = 5;

__global__ void __launch_bounds__(DP_THREADS) deprocess_map_pipe_kernel(const float* __restrict__ x, size_t n,
                                                                        const double* __restrict__ partial, int npartial,
                                                                        uint8_t* __restrict__ out, float* __restrict__ norm) {
  __shared__ __align__(128) float stage[DP_STAGES][DP_CHUNK];
  __shared__ uint64_t full_bar[DP_STAGES], empty_bar[DP_STAGES];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const size_t n4 = n / 4;                                        // whole 16-byte units; the last n % 4 floats: block 0 below
  const size_t chunks = (n4 * 4 + DP_CHUNK - 1) / DP_CHUNK;
  if (tid == 0) {
    for (int s = 0; s < DP_STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], DP_CONS_WARPS); }
    tc::fence_barrier_init();
  }
  __syncthreads();
  auto issue = [&](size_t c, int s) {
    const size_t first = c * DP_CHUNK;
    const uint32_t bytes = (uint32_t)(min((size_t)DP_CHUNK, n4 * 4 - first) * 4);
    tc::mbar_arrive_expect_tx(&full_bar[s], bytes);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     tc::smem_u32(&stage[s][0])),
                 "l"(reinterpret_cast<unsigned long long>(x + first)), "r"(bytes), "r"(tc::smem_u32(&full_bar[s]))
                 : "memory");
  };
  // the first ring of copies goes out before the statistics are reduced
  if (warp == DP_CONS_WARPS && lane == 0) {
    size_t c = blockIdx.x;
    for (int s = 0; s < DP_STAGES && c < chunks; ++s, c += gridDim.x) issue(c, s);
  }
  double ts, tq;
  block_sum_partials(partial, npartial, ts, tq);
  const double dmean = ts / (double)n;
  double var = tq / (double)n - dmean * dmean;      // population variance (ddof = 0)
  if (var < 0.0) var = 0.0;
  const float mean = (float)dmean, denom = (float)(sqrt(var) + 1e-7);
  const float rd = denom < 1e30f ? 1.f / denom : __int_as_float(0x7fc00000);     // denom >= 1e-7 by construction
  if (warp == DP_CONS_WARPS) {
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      size_t k = 0;
      for (size_t c = blockIdx.x; c < chunks; c += gridDim.x, ++k) {
        if (k >= DP_STAGES) {                        // re-use of a stage: wait until the consumers have drained it
          tc::mbar_wait(&empty_bar[s], ph ^ 1u);
          issue(c, s);
        }
        if (++s == DP_STAGES) { s = 0; ph ^= 1u; }
      }
    }
  } else {
    int s = 0;
    uint32_t ph = 0;
    for (size_t c = blockIdx.x; c < chunks; c += gridDim.x) {
      const size_t first4 = c * (DP_CHUNK / 4);
      const int cnt4 = (int)min((size_t)(DP_CHUNK / 4), n4 - first4);
      tc::mbar_wait(&full_bar[s], ph);
      const float4* st = reinterpret_cast<const float4*>(&stage[s][0]);
#pragma unroll
      for (int j = 0; j < DP_CHUNK / 4 / (32 * DP_CONS_WARPS); ++j) {
        const int i = tid + j * 32 * DP_CONS_WARPS;
        if (i < cnt4) {
          const float4 a = st[i];
          const float c[4] = {a.x - mean, a.y - mean, a.z - mean, a.w - mean};      // x -= x.mean()
          float q[4];
          quot4(c, denom, rd, q);                                                   // x /= (x.std() + epsilon)
          uint32_t r = 0;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            q[e] = __fadd_rn(__fmul_rn(q[e], 0.1f), 0.5f);                          // x *= 0.1; x += 0.5 (two roundings)
            // clip to [0, 1] (NaN -> 0), * 255, truncate; the reference's second clip cannot change a value in [0, 255]
            r |= (uint32_t)(__saturatef(q[e]) * 255.f) << (8 * e);
          }
          reinterpret_cast<uint32_t*>(out)[first4 + i] = r;
          if (norm) reinterpret_cast<float4*>(norm)[first4 + i] = make_float4(q[0], q[1], q[2], q[3]);
        }
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&empty_bar[s]);
      if (++s == DP_STAGES) { s = 0; ph ^= 1u; }
    }
  }
  if (blockIdx.x == 0 && tid < (int)(n & 3)) {
    const size_t j = n4 * 4 + tid;
    out[j] = deprocess_one(x[j], mean, denom);
    if (norm) norm[j] = deprocess_norm(x[j], mean, denom);
  }
}

int deprocess_partials(size_t n) {
  return grid_for((n + 3) / 4, 256, 4);
}

void launch_deprocess(const float* x, size_t n, double* partial, int npartial, uint8_t* out, float* norm,
                      cudaStream_t s) {
  deprocess_stats_kernel<<<npartial, 256, 0, s>>>(x, n, partial);
  static const int pipe = getenv("DCV_DEPROCESS_PIPE") ? atoi(getenv("DCV_DEPROCESS_PIPE")) : 1;
  // bulk copies need 16-byte aligned sources (the grid-stride form reads float4, so x is aligned that far anyway)
  if (pipe && n >= 4 * DP_CHUNK && (reinterpret_cast<uintptr_t>(x) & 15) == 0)
    deprocess_map_pipe_kernel<<<grid_for((n + DP_CHUNK - 1) / DP_CHUNK, 1, 5), DP_THREADS, 0, s>>>(x, n, partial, npartial, out, norm);
  else
    deprocess_map_kernel<<<grid_for((n + 3) / 4, 256, 8), 256, 0, s>>>(x, n, partial, npartial, out, norm);
}

// ------------------------------------------------------------------------------------------------
// Batched de-process: `items` projections of `len` floats each, every one normalised by ITS OWN mean / std
// (deprocess_image applied per projection, deepdream.py:483-498), uint8 out.  Same arithmetic as the kernels above;
// kItemPartials blocks per item in the statistics pass, their partials summed in a fixed order by every map block.
// VEC: len % 4 == 0 (every item starts 16-byte aligned) -> float4 / uchar4 accesses; otherwise scalar.
template <bool VEC>
__global__ void __launch_bounds__(256) deprocess_items_stats_kernel(const float* __restrict__ x, size_t len,
                                                                    double* __restrict__ partial) {
  const float* xi = x + (size_t)blockIdx.y * len;
  double s = 0.0, q = 0.0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (VEC) {
    const float4* x4 = reinterpret_cast<const float4*>(xi);
    const size_t n4 = len / 4;
    size_t i = i0;
    for (; i + stride < n4; i += 2 * stride) {
      const float4 a = __ldg(x4 + i), b = __ldg(x4 + i + stride);
      s += ((double)a.x + (double)a.y + (double)a.z + (double)a.w) + ((double)b.x + (double)b.y + (double)b.z + (double)b.w);
      q += ((double)a.x * a.x + (double)a.y * a.y + (double)a.z * a.z + (double)a.w * a.w) +
           ((double)b.x * b.x + (double)b.y * b.y + (double)b.z * b.z + (double)b.w * b.w);
    }
    for (; i < n4; i += stride) {
      const float4 a = __ldg(x4 + i);
      s += (double)a.x + (double)a.y + (double)a.z + (double)a.w;
      q += (double)a.x * a.x + (double)a.y * a.y + (double)a.z * a.z + (double)a.w * a.w;
    }
  } else {
    for (size_t i = i0; i < len; i += stride) { const float a = __ldg(xi + i); s += a; q += (double)a * a; }
  }
  __shared__ double rs[8], rq[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if ((threadIdx.x & 31) == 0) { rs[threadIdx.x >> 5] = s; rq[threadIdx.x >> 5] = q; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ts = 0.0, tq = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) { ts += rs[k]; tq += rq[k]; }
    double* p = partial + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 2;
    p[0] = ts; p[1] = tq;
  }
}

template <bool VEC>
__global__ void __launch_bounds__(256) deprocess_items_map_kernel(const float* __restrict__ x, size_t len,
                                                                  const double* __restrict__ partial, int npartial,
                                                                  uint8_t* __restrict__ out) {
  const float* xi = x + (size_t)blockIdx.y * len;
  uint8_t* oi = out + (size_t)blockIdx.y * len;
  double ts, tq;
  block_sum_partials(partial + (size_t)blockIdx.y * npartial * 2, npartial, ts, tq);
  const double dmean = ts / (double)len;
  double var = tq / (double)len - dmean * dmean;
  if (var < 0.0) var = 0.0;
  const float mean = (float)dmean, denom = (float)(sqrt(var) + 1e-7);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (VEC) {
    const float4* x4 = reinterpret_cast<const float4*>(xi);
    const size_t n4 = len / 4;
#pragma unroll 4
    for (size_t i = i0; i < n4; i += stride) {
      const float4 a = __ldg(x4 + i);
      uchar4 r;
      r.x = deprocess_one(a.x, mean, denom);
      r.y = deprocess_one(a.y, mean, denom);
      r.z = deprocess_one(a.z, mean, denom);
      r.w = deprocess_one(a.w, mean, denom);
      reinterpret_cast<uchar4*>(oi)[i] = r;
    }
  } else {
    for (size_t i = i0; i < len; i += stride) oi[i] = deprocess_one(__ldg(xi + i), mean, denom);
  }
}

void launch_deprocess_items(const float* x, int items, size_t len, double* partial, uint8_t* out, cudaStream_t s) {
  if (items <= 0 || len == 0) return;
  const dim3 grid(kItemPartials, items);
  if ((len & 3) == 0) {
    deprocess_items_stats_kernel<true><<<grid, 256, 0, s>>>(x, len, partial);
    deprocess_items_map_kernel<true><<<grid, 256, 0, s>>>(x, len, partial, kItemPartials, out);
  } else {
    deprocess_items_stats_kernel<false><<<grid, 256, 0, s>>>(x, len, partial);
    deprocess_items_map_kernel<false><<<grid, 256, 0, s>>>(x, len, partial, kItemPartials, out);
  }
}

// fp32 -> bf16 (round to nearest even), the compact result format of the projection entry points
__global__ void __launch_bounds__(256) convert_bf16_kernel(const float* __restrict__ x, size_t n, __nv_bfloat16* __restrict__ out) {
  const size_t n4 = n / 4, stride = (size_t)gridDim.x * blockDim.x;
  const float4* x4 = reinterpret_cast<const float4*>(x);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 a = __ldg(x4 + i);
    __nv_bfloat162 lo = __floats2bfloat162_rn(a.x, a.y), hi = __floats2bfloat162_rn(a.z, a.w);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(out)[i] = pk;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) out[n4 * 4 + threadIdx.x] = __float2bfloat16_rn(x[n4 * 4 + threadIdx.x]);
}

void launch_convert_bf16(const float* x, size_t n, __nv_bfloat16* out, cudaStream_t s) {
  if (!n) return;
  convert_bf16_kernel<<<grid_for((n + 3) / 4, 256, 8), 256, 0, s>>>(x, n, out);
}

// ------------------------------------------------------------------------------------------------
// Request tail of the API (reference main.py:60-72) on the device.
// 'caffe' preprocess of Keras-Applications 1.0.8 as main.py:60-61 applies it: float32(pixel), channel axis
// reversed, minus [103.939, 116.779, 123.68] in the reversed order.  img: npix x 3 bytes.
__global__ void preprocess_caffe_kernel(const uint8_t* __restrict__ img, size_t npix, float* __restrict__ out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
    const uint8_t* p = img + 3 * i;
    out[3 * i + 0] = (float)p[2] - 103.939f;
    out[3 * i + 1] = (float)p[1] - 116.779f;
    out[3 * i + 2] = (float)p[0] - 123.68f;
  }
}

void launch_preprocess_caffe(const uint8_t* img, size_t npix, float* out, cudaStream_t s) {
  if (!npix) return;
  preprocess_caffe_kernel<<<grid_for(npix, 256), 256, 0, s>>>(img, npix, out);
}

// find_top_filters (deepdream.py:369-380) on the per-channel fp64 sums, on the device: keep sums > 0, order by the
// sum (as float32 where the reference sums a float32 array) descending, ties in index order (Python's sort is
// stable), first `top`.  One block; feat_out[0..top) = indices (-1 where fewer exist), feat_out[top] = how many.
__global__ void __launch_bounds__(256) top_filters_kernel(const double* __restrict__ sums, int C, int as_f32, int top,
                                                          int* __restrict__ feat_out) {
  __shared__ int chosen[16];
  __shared__ double bv[256];
  __shared__ int bi[256];
  int count = 0;
  for (int t = 0; t < top; ++t) {
    double best = 0.0;
    int besti = -1;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      double v = sums[c];
      if (as_f32) v = (double)(float)v;
      if (!(v > 0.0)) continue;
      bool taken = false;
      for (int k = 0; k < t; ++k) taken |= (chosen[k] == c);
      if (taken) continue;
      if (besti < 0 || v > best) { best = v; besti = c; }          // c ascends: the first of equal sums is kept
    }
    bv[threadIdx.x] = best; bi[threadIdx.x] = besti;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if ((int)threadIdx.x < o) {
        const double v2 = bv[threadIdx.x + o];
        const int i2 = bi[threadIdx.x + o];
        const int i1 = bi[threadIdx.x];
        if (i2 >= 0 && (i1 < 0 || v2 > bv[threadIdx.x] || (v2 == bv[threadIdx.x] && i2 < i1))) { bv[threadIdx.x] = v2; bi[threadIdx.x] = i2; }
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) { chosen[t] = bi[0]; feat_out[t] = bi[0]; }
    if (bi[0] >= 0) ++count;
    __syncthreads();
  }
  if (threadIdx.x == 0) feat_out[top] = count;
}

void launch_top_filters(const double* sums, int C, int as_f32, int top, int* feat_out, cudaStream_t s) {
  top_filters_kernel<<<1, 256, 0, s>>>(sums, C, as_f32, top, feat_out);
}

// 2x2 mosaic + de-process (main.py:67-72): x = four projections (4,H,W,C) fp32, out = (2H,2W,C) uint8 with
// projection 0 | 1 on top of 2 | 3; mean / std over the whole mosaic = over the four projections together
// (partials from deprocess_stats_kernel over the 4*H*W*C floats).
__global__ void __launch_bounds__(256) deprocess_mosaic_kernel(const float* __restrict__ x, int H, int W, int C,
                                                               const double* __restrict__ partial, int npartial,
                                                               uint8_t* __restrict__ out) {
  double ts, tq;
  block_sum_partials(partial, npartial, ts, tq);
  const size_t per = (size_t)H * W * C, n = 4 * per;
  const double dmean = ts / (double)n;
  double var = tq / (double)n - dmean * dmean;
  if (var < 0.0) var = 0.0;
  const float mean = (float)dmean, denom = (float)(sqrt(var) + 1e-7);
  const size_t rowlen = (size_t)W * C;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int item = (int)(i / per);
    const size_t r = i % per;
    const size_t y = r / rowlen, xc = r % rowlen;
    const size_t oy = y + (size_t)(item >> 1) * H, ox = xc + (size_t)(item & 1) * rowlen;
    out[oy * 2 * rowlen + ox] = deprocess_one(__ldg(x + i), mean, denom);
  }
}

void launch_deprocess_mosaic(const float* x, int H, int W, int C, double* partial, int npartial, uint8_t* out, cudaStream_t s) {
  const size_t n = (size_t)4 * H * W * C;
  deprocess_stats_kernel<<<npartial, 256, 0, s>>>(x, n, partial);
  deprocess_mosaic_kernel<<<grid_for(n, 256, 8), 256, 0, s>>>(x, H, W, C, partial, npartial, out);
}

// fp64 (sum, sum of squares) block partials of x[0..n): partial[2*b], partial[2*b+1]; x 16-byte aligned
void launch_sum_stats(const float* x, size_t n, double* partial, int npartial, cudaStream_t s) {
  deprocess_stats_kernel<<<npartial, 256, 0, s>>>(x, n, partial);
}

// ------------------------------------------------------------------------------------------------
// Seed maps of the projections (deepdream.py:450-465): map[(k*B+b), y, x] = act[b, y, x, feat[k]].
__global__ void gather_map_kernel(const float* __restrict__ act, int B, size_t HW, int C,
                                  const int* __restrict__ feat, int n, float* __restrict__ maps) {
  const size_t total = (size_t)n * B * HW;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t pix = i % HW;
    const size_t item = i / HW;
    const int b = (int)(item % B), k = (int)(item / B);
    maps[i] = __ldg(act + ((size_t)b * HW + pix) * C + feat[k]);
  }
}

void launch_gather_map(const float* act, int B, int H, int W, int C, const int* feat_dev, int n, float* maps,
                       cudaStream_t s) {
  const size_t total = (size_t)n * B * H * W;
  if (!total) return;
  gather_map_kernel<<<grid_for(total, 256), 256, 0, s>>>(act, B, (size_t)H * W, C, feat_dev, n, maps);
}

// 'max' visualize mode (deepdream.py:454-457): keep the entries equal to the map maximum.  One
// block per map; activations are >= 0 so the int-ordered max of the float bits would do, but a plain
// float reduction keeps it general.
__global__ void __launch_bounds__(256) keep_max_kernel(float* __restrict__ maps, size_t len) {
  float* m = maps + (size_t)blockIdx.x * len;
  float best = -INFINITY;
  for (size_t i = threadIdx.x; i < len; i += blockDim.x) best = fmaxf(best, m[i]);
  __shared__ float red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = fmaxf(best, __shfl_xor_sync(0xffffffffu, best, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = best;
  __syncthreads();
  best = red[0];
#pragma unroll
  for (int k = 1; k < 8; ++k) best = fmaxf(best, red[k]);
  for (size_t i = threadIdx.x; i < len; i += blockDim.x) {
    const float v = m[i];
    m[i] = (v == best) ? v : 0.f;
  }
}

void launch_keep_max(float* maps, int n, size_t len, float*, cudaStream_t s) {
  if (n <= 0 || len == 0) return;
  keep_max_kernel<<<n, 256, 0, s>>>(maps, len);
}

// Unpool of single-channel maps with channel feat[k] of the switches: (n*B, h, w) -> (n*B, H, W).
__global__ void unpool_map_kernel(const float* __restrict__ maps, const uint8_t* __restrict__ sw, int B, int h, int w,
                                  int H, int W, int C, const int* __restrict__ feat, int n, float* __restrict__ out) {
  const size_t total = (size_t)n * B * H * W;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % W);
    size_t p = i / W;
    const int y = (int)(p % H);
    const size_t item = p / H;
    const int b = (int)(item % B), k = (int)(item / B);
    const int py = y >> 1, px = x >> 1;
    float r = 0.f;
    if (py < h && px < w) {
      const int q = ((y & 1) << 1) | (x & 1);
      const uint8_t sidx = __ldg(sw + (((size_t)b * h + py) * w + px) * C + feat[k]);
      if (sidx == q) r = __ldg(maps + (item * h + py) * w + px);
    }
    out[i] = r;
  }
}

void launch_unpool_map(const float* maps, const uint8_t* sw, int B, int h, int w, int C, const int* feat_dev, int n,
                       float* out, cudaStream_t s) {
  // output size is exactly (2h, 2w): the deconvnet path only sees even sizes
  const size_t total = (size_t)n * B * (2 * h) * (2 * w);
  if (!total) return;
  unpool_map_kernel<<<grid_for(total, 256), 256, 0, s>>>(maps, sw, B, h, w, 2 * h, 2 * w, C, feat_dev, n, out);
}

// Expand maps into channel feat[k] of an otherwise zero (n*B, H, W, C) tensor (only needed when the
// projected layer has no conv below it).
__global__ void scatter_map_kernel(const float* __restrict__ maps, size_t HW, int C, int B, const int* __restrict__ feat,
                                   size_t total, float* __restrict__ out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const size_t pi = i / C;              // item*HW + pix
    const int k = (int)((pi / HW) / B);
    out[i] = (c == feat[k]) ? maps[pi] : 0.f;
  }
}

void launch_scatter_map(const float* maps, int n, int B, int H, int W, int C, const int* feat_dev, float* out,
                        cudaStream_t s) {
  const size_t total = (size_t)n * B * H * W * C;
  if (!total) return;
  scatter_map_kernel<<<grid_for(total, 256), 256, 0, s>>>(maps, (size_t)H * W, C, B, feat_dev, total, out);
}

// ------------------------------------------------------------------------------------------------
// Rank-1 first backward conv (a3 in SURVEY.md §8a): the projected tensor has ONE non-zero channel f,
// so relu(conv3x3_same(x, flipT(W)))[p, ci] = relu(sum_tap map[p + tap] * wd[tap][f][ci]).
// One block = 64 pixels of one item; the 9 x cin weight slice sits in shared memory; thread layout
// (cin lane, pixel lane) keeps the NHWC stores coalesced.  9*H*W*cin MACs instead of 9*H*W*cin*cout.
__global__ void __launch_bounds__(256) rank1_deconv_kernel(const float* __restrict__ maps, const float* __restrict__ wd,
                                                           int cout, int cin, const int* __restrict__ feat, int B,
                                                           int H, int W, int relu, float* __restrict__ out) {
  extern __shared__ float smem[];
  float* wsl = smem;                 // [9][cin]
  float* tap = smem + 9 * cin;       // [64][9] map values around each pixel
  const int item = blockIdx.y;
  const int f = feat[item / B];
  for (int i = threadIdx.x; i < 9 * cin; i += blockDim.x) {
    const int t = i / cin, ci = i % cin;
    wsl[i] = __ldg(wd + ((size_t)t * cout + f) * cin + ci);
  }
  const int HW = H * W;
  const int p0 = blockIdx.x * 64;
  const float* m = maps + (size_t)item * HW;
  for (int i = threadIdx.x; i < 64 * 9; i += blockDim.x) {
    const int pp = i / 9, t = i % 9;
    const int p = p0 + pp;
    float v = 0.f;
    if (p < HW) {
      const int y = p / W + t / 3 - 1, x = p % W + t % 3 - 1;
      if (y >= 0 && y < H && x >= 0 && x < W) v = __ldg(m + y * W + x);
    }
    tap[i] = v;
  }
  __syncthreads();
  const size_t obase = ((size_t)item * HW + p0) * cin;
  const int total = 64 * cin;
  if ((cin & 3) == 0) {
    // four channels per thread: one broadcast tap and one 16-byte weight read per tap and four outputs, one 16-byte store
    // (the scalar form below spends ~40 instructions per 4-byte store and ran at 1.2 TB/s); same FMA order per output
    const int c4 = cin >> 2;
    const int npix = min(64, HW - p0);
    for (int i = threadIdx.x; i < npix * c4; i += blockDim.x) {
      const int pp = i / c4, cq = i - pp * c4;
      const float* tp = tap + pp * 9;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const float tv = tp[t];
        const float4 wv = *reinterpret_cast<const float4*>(wsl + t * cin + 4 * cq);
        acc.x = fmaf(tv, wv.x, acc.x); acc.y = fmaf(tv, wv.y, acc.y); acc.z = fmaf(tv, wv.z, acc.z); acc.w = fmaf(tv, wv.w, acc.w);
      }
      if (relu) { acc.x = fmaxf(acc.x, 0.f); acc.y = fmaxf(acc.y, 0.f); acc.z = fmaxf(acc.z, 0.f); acc.w = fmaxf(acc.w, 0.f); }
      *reinterpret_cast<float4*>(out + obase + (size_t)pp * cin + 4 * cq) = acc;
    }
    return;
  }
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int pp = i / cin, ci = i % cin;
    if (p0 + pp >= HW) break;
    float acc = 0.f;
#pragma unroll
    for (int t = 0; t < 9; ++t) acc = fmaf(tap[pp * 9 + t], wsl[t * cin + ci], acc);
    if (relu) acc = fmaxf(acc, 0.f);
    out[obase + i] = acc;
  }
}

void launch_rank1_deconv(const float* maps, const float* wd, int cout, int cin, const int* feat_dev, int n, int B,
                         int H, int W, int relu, float* out, cudaStream_t s) {
  if (n * B == 0) return;
  dim3 grid((H * W + 63) / 64, n * B);
  const size_t smem = (size_t)(9 * cin + 64 * 9) * sizeof(float);
  rank1_deconv_kernel<<<grid, 256, smem, s>>>(maps, wd, cout, cin, feat_dev, B, H, W, relu, out);
}

// ------------------------------------------------------------------------------------------------
// Weight preparation (DConvolution2D.__init__, deepdream.py:80-81): W (3,3,cin,cout) HWIO ->
// wd[kh][kw][co][ci] = W[2-kh][2-kw][ci][co], i.e. transpose(0,1,3,2) then flip both spatial axes.
__global__ void flip_transpose_kernel(const float* __restrict__ w, int cin, int cout, float* __restrict__ wd) {
  const size_t total = (size_t)9 * cin * cout;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int ci = (int)(i % cin);
    size_t p = i / cin;
    const int co = (int)(p % cout);
    const int t = (int)(p / cout);
    wd[i] = __ldg(w + ((size_t)(8 - t) * cin + ci) * cout + co);
  }
}

void launch_flip_transpose(const float* w_hwio, int cin, int cout, float* wd, cudaStream_t s) {
  flip_transpose_kernel<<<grid_for((size_t)9 * cin * cout, 256), 256, 0, s>>>(w_hwio, cin, cout, wd);
}

// ------------------------------------------------------------------------------------------------
// Dense layers (DDense, deepdream.py:264-321).  Batch is 1 in the API, so both directions are
// matrix-vector products bound by reading W once from HBM.
// forward: out[b][j] = act(sum_i x[b][i] W[i][j] + bias[j]); W rows are contiguous in j, so lanes
// run along j and blockIdx.y splits i; partials are reduced in a fixed order by stage 2.
__global__ void __launch_bounds__(256) dense_fwd_stage1(const float* __restrict__ x, const float* __restrict__ w,
                                                        int nin, int nout, int chunk, float* __restrict__ partial) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.z;
  const int i0 = blockIdx.y * chunk;
  const int i1 = min(nin, i0 + chunk);
  if (j >= nout) return;
  // fp64 accumulation is free here (the kernel waits on HBM) and keeps the cancellation-heavy dense
  // chains within the fp32 parity tolerance
  double acc = 0.0;
  const float* xb = x + (size_t)b * nin;
  for (int i = i0; i < i1; ++i) acc += (double)__ldg(xb + i) * (double)__ldg(w + (size_t)i * nout + j);
  partial[((size_t)blockIdx.y * gridDim.z + b) * nout + j] = (float)acc;
}

__global__ void dense_fwd_stage2(const float* __restrict__ partial, int nsplit, int B, int nout,
                                 const float* __restrict__ bias, int act, float* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (j >= nout) return;
  double accd = 0.0;
  for (int k = 0; k < nsplit; ++k) accd += (double)partial[((size_t)k * B + b) * nout + j];
  if (bias) accd += (double)bias[j];
  float acc = (float)accd;
  if (act == DCV_ACT_RELU) acc = fmaxf(acc, 0.f);
  out[(size_t)b * nout + j] = acc;
}

int dense_splits(int nin) {
  int s = (nin + 255) / 256;
  if (s > 128) s = 128;
  return s < 1 ? 1 : s;
}

void launch_dense_fwd(const float* x, const float* w, const float* bias, int B, int nin, int nout, int act,
                      float* partial, float* out, cudaStream_t s) {
  const int ns = dense_splits(nin);
  const int chunk = (nin + ns - 1) / ns;
  dim3 g1((nout + 255) / 256, ns, B);
  dense_fwd_stage1<<<g1, 256, 0, s>>>(x, w, nin, nout, chunk, partial);
  dim3 g2((nout + 255) / 256, B);
  dense_fwd_stage2<<<g2, 256, 0, s>>>(partial, ns, B, nout, bias, act, out);
}

// backward: out[b][i] = sum_j x[b][j] W[i][j]  (x W^T, deepdream.py:289-297).  One warp per row i.
__global__ void __launch_bounds__(256) dense_bwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                        int nin, int nout, float* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.y;
  if (warp >= nin) return;
  const float* wr = w + (size_t)warp * nout;
  const float* xb = x + (size_t)b * nout;
  double acc = 0.0;
  for (int j = lane; j < nout; j += 32) acc += (double)__ldg(xb + j) * (double)__ldg(wr + j);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) out[(size_t)b * nin + warp] = (float)acc;
}

void launch_dense_bwd(const float* x, const float* w, int B, int nin, int nout, float* out, cudaStream_t s) {
  dim3 g((nin + 7) / 8, B);
  dense_bwd_kernel<<<g, 256, 0, s>>>(x, w, nin, nout, out);
}

// row softmax (one block per row), used for 'predictions' in both directions (deepdream.py:226-235)
__global__ void __launch_bounds__(256) softmax_rows_kernel(const float* __restrict__ x, int n, float* __restrict__ out) {
  const float* r = x + (size_t)blockIdx.x * n;
  float* o = out + (size_t)blockIdx.x * n;
  __shared__ float red[8];
  float m = -INFINITY;
  for (int i = threadIdx.x; i < n; i += blockDim.x) m = fmaxf(m, r[i]);
#pragma unroll
  for (int k = 16; k > 0; k >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, k));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  m = red[0];
#pragma unroll
  for (int k = 1; k < 8; ++k) m = fmaxf(m, red[k]);
  __syncthreads();
  float sum = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) sum += expf(r[i] - m);
#pragma unroll
  for (int k = 16; k > 0; k >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, k);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sum;
  __syncthreads();
  sum = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) sum += red[k];
  for (int i = threadIdx.x; i < n; i += blockDim.x) o[i] = expf(r[i] - m) / sum;
}

void launch_softmax_rows(const float* x, int B, int n, float* out, cudaStream_t s) {
  if (B > 0) softmax_rows_kernel<<<B, 256, 0, s>>>(x, n, out);
}

__global__ void relu_kernel(const float* __restrict__ x, size_t n, float* __restrict__ out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = fmaxf(x[i], 0.f);
}

void launch_relu(const float* x, size_t n, float* out, cudaStream_t s) {
  if (n) relu_kernel<<<grid_for(n, 256), 256, 0, s>>>(x, n, out);
}

}  // namespace dcv
