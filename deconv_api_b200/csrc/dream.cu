// DeepDream gradient-ascent loop (a12 in SURVEY.md §8a).  The reference has NO such code (its
// app/deepdream.py is deconvnet-only); the algorithm is the derived spec of SURVEY.md Appendix A:
//   per octave (small -> large): img = resize_bilinear(img, shape); `iterations` times:
//       x = roll(img, (oy,ox)); loss = sum_l coeff_l * mean(act_l(x)^2); g = dloss/dx (true backprop: ReLU
//       masks by forward sign, max-pool routes by switch, conv backward = flipped-filter conv, no ReLU);
//       img += step * unroll(g) / max(sqrt(mean(g^2)), 1e-7)        (per image)
//     then img += resize(orig, shape) - resize(orig at the previous octave's shape, shape).
// Forward = the fp32 tape of dcv_forward; backward re-uses the backward-conv and unpool kernels of the
// projection path.  Streaming kernels here: bilinear resize (half-pixel centres, no antialias), jitter
// roll, gradient prep (seed + ReLU mask), RMS partials, fused normalise + un-roll + update.
#include "common.cuh"
#include "internal.cuh"
#include "tc_ptx.cuh"

#include <algorithm>
#include <cstdlib>
#include <vector>

using namespace dcv;

namespace {

inline int grid_for(size_t work, int block, int waves = 16) {
  size_t b = (work + block - 1) / block, cap = (size_t)kNumSMs * waves;
  if (b > cap) b = cap;
  return (int)(b < 1 ? 1 : b);
}

// Bilinear resize, NHWC, align_corners = false: src = (dst + 0.5) * (in/out) - 0.5 clamped at 0, neighbours
// x0 = floor(src), x1 = min(x0 + 1, in - 1).  One block row (blockIdx.y) = one output row of one image, so the
// vertical weights are computed once per block and there is no 64-bit div/mod per element; one thread = one
// output pixel (all C channels, C <= 4 in registers).
template <int CMAX>
__global__ void __launch_bounds__(256)
resize_bilinear_kernel(const float* __restrict__ x, int H, int W, int C, int OH, int OW, float* __restrict__ out) {
  const int oy = blockIdx.y % OH, b = blockIdx.y / OH;
  const float sh = (float)H / (float)OH, sw = (float)W / (float)OW;
  float fy = sh * ((float)oy + 0.5f) - 0.5f; if (fy < 0.f) fy = 0.f;
  int y0 = (int)fy; if (y0 > H - 1) y0 = H - 1;
  const int y1 = y0 + (y0 < H - 1 ? 1 : 0);
  const float ly = fy - (float)y0, hy = 1.f - ly;
  const float* r0 = x + ((size_t)b * H + y0) * W * C;
  const float* r1 = x + ((size_t)b * H + y1) * W * C;
  float* o = out + ((size_t)b * OH + oy) * OW * C;
  for (int ox = blockIdx.x * blockDim.x + threadIdx.x; ox < OW; ox += gridDim.x * blockDim.x) {
    float fx = sw * ((float)ox + 0.5f) - 0.5f; if (fx < 0.f) fx = 0.f;
    int x0 = (int)fx; if (x0 > W - 1) x0 = W - 1;
    const int x1 = x0 + (x0 < W - 1 ? 1 : 0);
    const float lx = fx - (float)x0, hx = 1.f - lx;
#pragma unroll
    for (int c = 0; c < CMAX; ++c) {
      if (c < C) {
        const float v00 = __ldg(r0 + x0 * C + c), v01 = __ldg(r0 + x1 * C + c);
        const float v10 = __ldg(r1 + x0 * C + c), v11 = __ldg(r1 + x1 * C + c);
        o[ox * C + c] = hy * (hx * v00 + lx * v01) + ly * (hx * v10 + lx * v11);
      }
    }
  }
}

// Strip form used for every shape (profiles/ub_stream.cu compares the candidates).  A block of 256 threads owns
// RESIZE_U * 256 consecutive elements f = ox*C + c of the flat output row and walks RESIZE_STRIP output rows:
//  * the x coordinates / weights of a thread's elements are computed once, the y coordinates of the strip once per
//    block (shared memory) - the per-element coordinate arithmetic, not memory, bounded the one-thread-per-pixel form;
//  * the horizontally interpolated values of the two source rows stay in registers and are re-used while the
//    source rows do not change (an enlargement by 1.4 touches 0.71 new source rows per output row; the branches are
//    warp-uniform), so ~1.4 loads + 1 store per output element instead of 4 + 1, all warp-coalesced.
// The arithmetic is the same expression in the same order as the per-pixel form above.
constexpr int RESIZE_STRIP = 16;
constexpr int RESIZE_U = 3;
__global__ void __launch_bounds__(256)
resize_bilinear_strip_kernel(const float* __restrict__ x, int H, int W, int C, int OH, int OW, float* __restrict__ out) {
  __shared__ int sy0[RESIZE_STRIP], sy1[RESIZE_STRIP];
  __shared__ float sly[RESIZE_STRIP];
  const int strips = (OH + RESIZE_STRIP - 1) / RESIZE_STRIP;
  const int b = blockIdx.y / strips, oy_begin = (blockIdx.y % strips) * RESIZE_STRIP;
  const int rows = min(RESIZE_STRIP, OH - oy_begin);
  if (threadIdx.x < RESIZE_STRIP) {
    const int oy = min(oy_begin + (int)threadIdx.x, OH - 1);
    const float sh = (float)H / (float)OH;
    float fy = sh * ((float)oy + 0.5f) - 0.5f; if (fy < 0.f) fy = 0.f;
    int y0 = (int)fy; if (y0 > H - 1) y0 = H - 1;
    sy0[threadIdx.x] = y0;
    sy1[threadIdx.x] = y0 + (y0 < H - 1 ? 1 : 0);
    sly[threadIdx.x] = fy - (float)y0;
  }
  const int n = OW * C;
  const int f0 = blockIdx.x * 256 * RESIZE_U + threadIdx.x;
  const float sw = (float)W / (float)OW;
  int i0[RESIZE_U], i1[RESIZE_U];
  float lx[RESIZE_U];
#pragma unroll
  for (int k = 0; k < RESIZE_U; ++k) {
    const int f = min(f0 + k * 256, n - 1);
    const int ox = f / C, c = f - ox * C;
    float fx = sw * ((float)ox + 0.5f) - 0.5f; if (fx < 0.f) fx = 0.f;
    int x0 = (int)fx; if (x0 > W - 1) x0 = W - 1;
    const int x1 = x0 + (x0 < W - 1 ? 1 : 0);
    lx[k] = fx - (float)x0;
    i0[k] = x0 * C + c; i1[k] = x1 * C + c;
  }
  __syncthreads();
  const float* img = x + (size_t)b * H * W * C;
  const size_t wc = (size_t)W * C;
  float* o = out + ((size_t)b * OH + oy_begin) * n;
  int cy0 = -1, cy1 = -1;                       // source rows whose horizontal interpolation is held in h0 / h1
  float h0[RESIZE_U], h1[RESIZE_U];
  for (int r = 0; r < rows; ++r, o += n) {
    const int y0 = sy0[r], y1 = sy1[r];
    const float ly = sly[r], hy = 1.f - ly;
    float n0[RESIZE_U], n1[RESIZE_U];
    if (y0 == cy0) {
#pragma unroll
      for (int k = 0; k < RESIZE_U; ++k) n0[k] = h0[k];
    } else if (y0 == cy1) {
#pragma unroll
      for (int k = 0; k < RESIZE_U; ++k) n0[k] = h1[k];
    } else {
      const float* rp = img + (size_t)y0 * wc;
      float a[RESIZE_U], bb[RESIZE_U];
#pragma unroll
      for (int k = 0; k < RESIZE_U; ++k) { a[k] = __ldg(rp + i0[k]); bb[k] = __ldg(rp + i1[k]); }
#pragma unroll
      for (int k = 0; k < RESIZE_U; ++k) n0[k] = (1.f - lx[k]) * a[k] + lx[k] * bb[k];
    }
    if (y1 == y0) {
#pragma unroll
      for (int k = 0; k < RESIZE_U; ++k) n1[k] = n0[k];
    } else if (y1 == cy1) {
#pragma unroll
      for (int k = 0; k < RESIZE_U; ++k) n1[k] = h1[k];
    } else {
      const float* rp = img + (size_t)y1 * wc;
      float a[RESIZE_U], bb[RESIZE_U];
#pragma unroll
      for (int k = 0; k < RESIZE_U; ++k) { a[k] = __ldg(rp + i0[k]); bb[k] = __ldg(rp + i1[k]); }
#pragma unroll
      for (int k = 0; k < RESIZE_U; ++k) n1[k] = (1.f - lx[k]) * a[k] + lx[k] * bb[k];
    }
#pragma unroll
    for (int k = 0; k < RESIZE_U; ++k) { h0[k] = n0[k]; h1[k] = n1[k]; }
    cy0 = y0; cy1 = y1;
#pragma unroll
    for (int k = 0; k < RESIZE_U; ++k)
      if (f0 + k * 256 < n) o[f0 + k * 256] = hy * h0[k] + ly * h1[k];
  }
}

// Parity-slot form (default).  The strip kernel above spends its time on instruction issue, not on memory (ncu: 81 %
// of the issue slots, DRAM at 52 %): every gathered load carried four instructions of 64-bit address arithmetic and
// every output row a round of register copies between the "held" and the "needed" rows.  Here
//  * a gathered load's address is the thread's pointer into source row 0 plus row * row bytes: one IMAD.WIDE.U32,
//  * the horizontally interpolated source row y lives in the register slot of y's parity (y0 and y1 = y0 + 1 never
//    share a slot), tagged with its row number: a row is loaded and interpolated once, nothing is copied, the blend
//    picks its two operands with a block-uniform select.
// A variant that staged the source rows through shared memory (coalesced loads, gathers from shared) was slower
// (0.215 against 0.152 ms at 8 x 1462^2 -> 2048^2): one barrier per source row serialises the block on the load latency.
// Same expression, same operation order as the two kernels above.
constexpr int RZ_THREADS = 256;
constexpr int RZ_U = 3;
constexpr int RZ_SEG = RZ_THREADS * RZ_U;
constexpr int RZ_STRIP = 32;

// *(const float*)(p + row * row_bytes), read-only path; the address is one mad.wide.u32
__device__ __forceinline__ float ldg_row(const char* p, unsigned row, unsigned row_bytes) {
  float v;
  asm("{\n"
      ".reg .u64 a;\n"
      "mad.wide.u32 a, %1, %2, %3;\n"
      "ld.global.nc.f32 %0, [a];\n"
      "}\n" : "=f"(v) : "r"(row), "r"(row_bytes), "l"(p));
  return v;
}

__global__ void __launch_bounds__(RZ_THREADS)
resize_bilinear_slots_kernel(const float* __restrict__ x, int H, int W, int C, int OH, int OW, float* __restrict__ out) {
  __shared__ int sy0[RZ_STRIP], sy1[RZ_STRIP];
  __shared__ float sly[RZ_STRIP];
  const int tid = threadIdx.x;
  const int strips = (OH + RZ_STRIP - 1) / RZ_STRIP;
  const int b = blockIdx.y / strips, oy_begin = (blockIdx.y % strips) * RZ_STRIP;
  const int rows = min(RZ_STRIP, OH - oy_begin);
  if (tid < RZ_STRIP) {
    const int oy = min(oy_begin + tid, OH - 1);
    const float sh = (float)H / (float)OH;
    float fy = sh * ((float)oy + 0.5f) - 0.5f; if (fy < 0.f) fy = 0.f;
    int y0 = (int)fy; if (y0 > H - 1) y0 = H - 1;
    sy0[tid] = y0;
    sy1[tid] = y0 + (y0 < H - 1 ? 1 : 0);
    sly[tid] = fy - (float)y0;
  }
  const int n = OW * C;
  const int f_begin = blockIdx.x * RZ_SEG;
  const float sw = (float)W / (float)OW;
  // per-thread byte pointers to the two neighbours in source row 0 of this image; row y is one 32 x 32 -> 64 bit
  // multiply-add away (IMAD.WIDE.U32)
  const unsigned row_bytes = (unsigned)W * (unsigned)C * 4u;
  const char* img = reinterpret_cast<const char*>(x + (size_t)b * H * W * C);
  const char *q0[RZ_U], *q1[RZ_U];
  float lx[RZ_U];
#pragma unroll
  for (int k = 0; k < RZ_U; ++k) {
    const int f = min(f_begin + tid + k * RZ_THREADS, n - 1);
    const int ox = f / C, c = f - ox * C;
    float fx = sw * ((float)ox + 0.5f) - 0.5f; if (fx < 0.f) fx = 0.f;
    int x0 = (int)fx; if (x0 > W - 1) x0 = W - 1;
    const int x1 = x0 + (x0 < W - 1 ? 1 : 0);
    lx[k] = fx - (float)x0;
    q0[k] = img + (size_t)(x0 * C + c) * 4; q1[k] = img + (size_t)(x1 * C + c) * 4;
  }
  __syncthreads();
  float* o = out + ((size_t)b * OH + oy_begin) * n + f_begin + tid;
  int tag_e = -1, tag_o = -1;                  // source rows held in the even / odd slot
  float he[RZ_U], ho[RZ_U];
#define RZ_LOAD_ROW(y, dst)                                                                  \
  {                                                                                          \
    float a[RZ_U], bb[RZ_U];                                                                 \
    _Pragma("unroll") for (int k = 0; k < RZ_U; ++k) {                                       \
      a[k] = ldg_row(q0[k], (unsigned)(y), row_bytes);                                       \
      bb[k] = ldg_row(q1[k], (unsigned)(y), row_bytes);                                      \
    }                                                                                        \
    _Pragma("unroll") for (int k = 0; k < RZ_U; ++k) dst[k] = (1.f - lx[k]) * a[k] + lx[k] * bb[k];           \
  }
  for (int r = 0; r < rows; ++r, o += n) {
    const int y0 = sy0[r], y1 = sy1[r];
    const float ly = sly[r], hy = 1.f - ly;
    if (y0 & 1) { if (tag_o != y0) { RZ_LOAD_ROW(y0, ho) tag_o = y0; } }
    else        { if (tag_e != y0) { RZ_LOAD_ROW(y0, he) tag_e = y0; } }
    if (y1 & 1) { if (tag_o != y1) { RZ_LOAD_ROW(y1, ho) tag_o = y1; } }
    else        { if (tag_e != y1) { RZ_LOAD_ROW(y1, he) tag_e = y1; } }
    const bool top_odd = y0 & 1, bot_odd = y1 & 1;
#pragma unroll
    for (int k = 0; k < RZ_U; ++k) {
      const float top = top_odd ? ho[k] : he[k], bot = bot_odd ? ho[k] : he[k];
      if (f_begin + tid + k * RZ_THREADS < n) o[k * RZ_THREADS] = hy * top + ly * bot;
    }
  }
#undef RZ_LOAD_ROW
}

// Pipelined form (default): the source rows arrive through a ring of shared-memory stages filled by bulk async copies
// (cp.async.bulk global -> shared, completion on an mbarrier), issued by a producer warp several rows ahead of the eight
// consumer warps.  The parity-slot kernel above still waits for its own gathered loads (ncu: 13-20 warps stalled on the
// long scoreboard per issued instruction, issue slots 41-58 % busy, DRAM 4.2-4.5 TB/s); here a consumer's loads are
// shared-memory gathers by 32-bit addresses and the only wait is for data that HBM has not delivered yet.
//  * A block owns RT_SEG consecutive elements of the flat output row over RT_STRIP output rows.  Every source row it
//    needs is ONE contiguous run of floats (block-uniform start / length); the copy starts at the 16-byte boundary at or
//    below the run and ends at the one at or above it (bulk copies move whole 16-byte units; both boundaries lie inside
//    the allocation because allocations are 16-byte granular).
//  * The rows are consumed in increasing order, each once: y0 / y1 of consecutive output rows never decrease, so "row
//    not seen yet" is y > last, for the producer and the consumers alike.  The interpolated row goes to the register
//    slot of its parity as in the kernel above.
// Same expression, same operation order as the other forms.
constexpr int RT_CONS_WARPS = 8;
constexpr int RT_THREADS = 32 * (RT_CONS_WARPS + 1);
constexpr int RT_U = 3;
constexpr int RT_SEG = 32 * RT_CONS_WARPS * RT_U;
constexpr int RT_STRIP = 32;
constexpr int RT_MAX_STAGES = 8;

__device__ __forceinline__ void bulk_load_1d(void* dst_smem, unsigned long long src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   tc::smem_u32(dst_smem)),
               "l"(src), "r"(bytes), "r"(tc::smem_u32(bar))
               : "memory");
}

__global__ void __launch_bounds__(RT_THREADS)
resize_bilinear_pipe_kernel(const float* __restrict__ x, int H, int W, int C, int OH, int OW, float* __restrict__ out,
                            int stage_bytes, int nstage) {
  extern __shared__ __align__(128) uint8_t rt_stage[];
  __shared__ int sy0[RT_STRIP], sy1[RT_STRIP];
  __shared__ float sly[RT_STRIP];
  __shared__ uint64_t full_bar[RT_MAX_STAGES], empty_bar[RT_MAX_STAGES];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int strips = (OH + RT_STRIP - 1) / RT_STRIP;
  const int b = blockIdx.y / strips, oy_begin = (blockIdx.y % strips) * RT_STRIP;
  const int rows = min(RT_STRIP, OH - oy_begin);
  if (tid < RT_STRIP) {
    const int oy = min(oy_begin + tid, OH - 1);
    const float sh = (float)H / (float)OH;
    float fy = sh * ((float)oy + 0.5f) - 0.5f; if (fy < 0.f) fy = 0.f;
    int y0 = (int)fy; if (y0 > H - 1) y0 = H - 1;
    sy0[tid] = y0;
    sy1[tid] = y0 + (y0 < H - 1 ? 1 : 0);
    sly[tid] = fy - (float)y0;
  }
  if (tid == 0) {
    for (int s = 0; s < nstage; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], RT_CONS_WARPS); }
    tc::fence_barrier_init();
  }
  const int n = OW * C;
  const int f_begin = blockIdx.x * RT_SEG;
  const float sw = (float)W / (float)OW;
  // the run of source floats of this segment: [s_lo, s_lo + len) of every source row
  int s_lo, len;
  {
    float fa = sw * ((float)(f_begin / C) + 0.5f) - 0.5f; if (fa < 0.f) fa = 0.f;
    int xa = (int)fa; if (xa > W - 1) xa = W - 1;
    float fb = sw * ((float)(min(f_begin + RT_SEG - 1, n - 1) / C) + 0.5f) - 0.5f; if (fb < 0.f) fb = 0.f;
    int xb = (int)fb; if (xb > W - 1) xb = W - 1;
    xb += (xb < W - 1 ? 1 : 0);
    s_lo = xa * C;
    len = (xb + 1) * C - s_lo;
  }
  const unsigned long long row_bytes = (unsigned long long)W * C * 4ull;
  const unsigned long long run0 = reinterpret_cast<unsigned long long>(x) + (unsigned long long)b * H * row_bytes + (unsigned long long)s_lo * 4ull;
  __syncthreads();
  if (warp == RT_CONS_WARPS) {
    // ===== producer: one lane walks the strip's source rows =====
    if (lane == 0) {
      int last = -1, s = 0;
      uint32_t ph = 0;
      for (int r = 0; r < rows; ++r) {
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          const int y = t ? sy1[r] : sy0[r];
          if (y > last) {
            last = y;
            const unsigned long long a = run0 + (unsigned long long)y * row_bytes;
            const unsigned long long a0 = a & ~15ull;
            const uint32_t bytes = (uint32_t)(((a + (unsigned long long)len * 4ull + 15ull) & ~15ull) - a0);
            tc::mbar_wait(&empty_bar[s], ph ^ 1u);
            tc::mbar_arrive_expect_tx(&full_bar[s], bytes);
            bulk_load_1d(rt_stage + (size_t)s * stage_bytes, a0, bytes, &full_bar[s]);
            if (++s == nstage) { s = 0; ph ^= 1u; }
          }
        }
      }
    }
    return;
  }
  // ===== consumers =====
  uint32_t g0[RT_U], g1[RT_U];                 // byte offsets of the two horizontal neighbours inside the run
  float lx[RT_U];
#pragma unroll
  for (int k = 0; k < RT_U; ++k) {
    const int f = min(f_begin + tid + k * 32 * RT_CONS_WARPS, n - 1);
    const int ox = f / C, c = f - ox * C;
    float fx = sw * ((float)ox + 0.5f) - 0.5f; if (fx < 0.f) fx = 0.f;
    int x0 = (int)fx; if (x0 > W - 1) x0 = W - 1;
    const int x1 = x0 + (x0 < W - 1 ? 1 : 0);
    lx[k] = fx - (float)x0;
    g0[k] = (uint32_t)(x0 * C + c - s_lo) * 4u; g1[k] = (uint32_t)(x1 * C + c - s_lo) * 4u;
  }
  float* o = out + ((size_t)b * OH + oy_begin) * n + f_begin + tid;
  const uint32_t stage0 = tc::smem_u32(rt_stage);
  float he[RT_U], ho[RT_U];
  int last = -1, s = 0;
  uint32_t ph = 0;
  for (int r = 0; r < rows; ++r, o += n) {
    const int y0 = sy0[r], y1 = sy1[r];
    const float ly = sly[r], hy = 1.f - ly;
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      const int y = t ? y1 : y0;
      if (y > last) {
        last = y;
        const uint32_t mis = (uint32_t)((run0 + (unsigned long long)y * row_bytes) & 15ull);
        const uint32_t base = stage0 + (uint32_t)s * (uint32_t)stage_bytes + mis;
        tc::mbar_wait(&full_bar[s], ph);
        float v[RT_U];
#pragma unroll
        for (int k = 0; k < RT_U; ++k) v[k] = (1.f - lx[k]) * tc::lds_f32(base + g0[k]) + lx[k] * tc::lds_f32(base + g1[k]);
        if (y & 1) {
#pragma unroll
          for (int k = 0; k < RT_U; ++k) ho[k] = v[k];
        } else {
#pragma unroll
          for (int k = 0; k < RT_U; ++k) he[k] = v[k];
        }
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&empty_bar[s]);
        if (++s == nstage) { s = 0; ph ^= 1u; }
      }
    }
    const bool top_odd = y0 & 1, bot_odd = y1 & 1;
#pragma unroll
    for (int k = 0; k < RT_U; ++k) {
      const float top = top_odd ? ho[k] : he[k], bot = bot_odd ? ho[k] : he[k];
      if (f_begin + tid + k * 32 * RT_CONS_WARPS < n) o[k * 32 * RT_CONS_WARPS] = hy * top + ly * bot;
    }
  }
}

// numpy.roll over the two spatial axes: out[b, y, x] = in[b, (y - oy) mod H, (x - ox) mod W].  One block row =
// one output row: a row is W*C contiguous floats rotated by ox*C, so the kernel is two contiguous segment copies
// per row with no per-element div/mod.
// jit (optional): device array {counter, oy0, ox0, oy1, ox1, ...}; the offsets of iteration `counter` are used, so
// that an iteration captured in a CUDA graph can be replayed with a different jitter every time
__global__ void __launch_bounds__(256)
roll_kernel(const float* __restrict__ x, int B, int H, int W, int C, int oy, int ox, const int* __restrict__ jit,
            float* __restrict__ out) {
  if (jit) { const int c = jit[0]; oy = jit[1 + 2 * c]; ox = jit[2 + 2 * c]; }
  const int y = blockIdx.y % H, b = blockIdx.y / H;
  int sy = (y - oy) % H; if (sy < 0) sy += H;
  int sx = ox % W; if (sx < 0) sx += W;
  const int wc = W * C, shift = sx * C;                  // out[f] = in[(f - shift) mod wc]
  const float* src = x + ((size_t)b * H + sy) * wc;
  float* dst = out + ((size_t)b * H + y) * wc;
  // 4-byte accesses that coalesce to full 128-byte lines on both sides (the rotation breaks 16-byte alignment of one
  // side for most shifts); eight independent elements per thread keep enough loads in flight.  Measured against the
  // 16-byte-store, shared-memory and shuffle forms in profiles/ub_stream.cu: this one matches a device-to-device copy.
  const int tstride = gridDim.x * blockDim.x;
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  for (; f + 7 * tstride < wc; f += 8 * tstride) {
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      int g = f + k * tstride - shift; if (g < 0) g += wc;
      v[k] = __ldg(src + g);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) dst[f + k * tstride] = v[k];
  }
  for (; f < wc; f += tstride) {
    int g = f - shift; if (g < 0) g += wc;
    dst[f] = __ldg(src + g);
  }
}

// gradient w.r.t. a layer's pre-activation: g = mask(act > 0) * (g_above + scale * act).  g_above may be absent.
__global__ void dream_prep_kernel(float* __restrict__ g, const float* __restrict__ act, size_t n, float scale, int has_g, int relu) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float a = __ldg(act + i);
    float v = (has_g ? g[i] : 0.f) + scale * a;
    if (relu && !(a > 0.f)) v = 0.f;
    g[i] = v;
  }
}

// img[b,y,x,c] += step / max(rms_b, 1e-7) * g[b,(y+oy) mod H,(x+ox) mod W,c]; rms from the fp64 block partials
__global__ void __launch_bounds__(256)
dream_update_kernel(float* __restrict__ img, const float* __restrict__ g, const double* __restrict__ partial, int np,
                    int H, int W, int C, float step, int oy, int ox, const int* __restrict__ jit) {
  if (jit) { const int c = jit[0]; oy = jit[1 + 2 * c]; ox = jit[2 + 2 * c]; }
  const size_t n = (size_t)H * W * C;
  double ts, tq;
  block_sum_partials(partial, np, ts, tq);
  const float rms = (float)sqrt(tq / (double)n);
  const float sc = step / fmaxf(rms, 1e-7f);
  // blocks walk whole rows (two contiguous segments of the rolled gradient row per image row): 32-bit arithmetic, no
  // division per element (the flat 64-bit index form took 20 us for a 512x512 image)
  int sxs = ox % W; if (sxs < 0) sxs += W;
  const unsigned wc = (unsigned)W * (unsigned)C, shift = (unsigned)sxs * (unsigned)C;    // g row index = (f + shift) mod wc
  for (int y = blockIdx.x; y < H; y += gridDim.x) {
    int sy = (y + oy) % H; if (sy < 0) sy += H;
    float* irow = img + (size_t)y * wc;
    const float* grow = g + (size_t)sy * wc;
    for (unsigned f = threadIdx.x; f < wc; f += blockDim.x) {
      unsigned sf = f + shift; if (sf >= wc) sf -= wc;
      irow[f] = fmaf(sc, __ldg(grow + sf), irow[f]);
    }
  }
}

__global__ void bump_counter_kernel(int* counter) { *counter += 1; }

__global__ void add_diff_kernel(float* __restrict__ img, const float* __restrict__ a, const float* __restrict__ b, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    img[i] += a[i] - b[i];
}

}  // namespace

void launch_resize_bilinear(const float* x, int B, int H, int W, int C, int OH, int OW, float* out, cudaStream_t s) {
  if ((size_t)B * OH * OW * C == 0) return;
  // DCV_RESIZE_PIXEL: 0 = the pipelined kernel (default), 1 = one thread per pixel, 2 = the strip kernel, 3 = the parity-slot kernel
  static const int per_pixel = getenv("DCV_RESIZE_PIXEL") ? atoi(getenv("DCV_RESIZE_PIXEL")) : 0;
  const int strips = (OH + RESIZE_STRIP - 1) / RESIZE_STRIP, zstrips = (OH + RZ_STRIP - 1) / RZ_STRIP;
  // pipelined kernel: one stage holds the source run of an output segment (+ 16-byte alignment slack on both sides)
  const double ratio = (double)W / (double)OW;
  const long long run_floats = ((long long)((RT_SEG / C + 2) * ratio) + 4) * C;
  const int stage_bytes = (int)((run_floats * 4 + 32 + 127) / 128 * 128);
  const int nstage = std::min(RT_MAX_STAGES, 36 * 1024 / std::max(stage_bytes, 1));
  const bool piped = per_pixel == 0 && run_floats < (1 << 20) && nstage >= 3 && zstrips == (OH + RT_STRIP - 1) / RT_STRIP;
  const bool staged = (per_pixel == 0 || per_pixel == 3) && (long long)H * W * C < (1ll << 30);     // row * row bytes is a 32 x 32 bit product
  int b0 = 0;
  while (b0 < B) {                                   // gridDim.y <= 65535 per launch
    const int nb = std::min(B - b0, std::max(1, 65535 / (per_pixel == 1 ? OH : strips)));
    const float* xi = x + (size_t)b0 * H * W * C;
    float* oi = out + (size_t)b0 * OH * OW * C;
    if (per_pixel == 1 && C <= 4) {
      dim3 grid((OW + 255) / 256, nb * OH);
      resize_bilinear_kernel<4><<<grid, 256, 0, s>>>(xi, H, W, C, OH, OW, oi);
    } else if (piped) {
      dim3 grid((OW * C + RT_SEG - 1) / RT_SEG, nb * zstrips);
      resize_bilinear_pipe_kernel<<<grid, RT_THREADS, (size_t)nstage * stage_bytes, s>>>(xi, H, W, C, OH, OW, oi, stage_bytes, nstage);
    } else if (staged) {
      dim3 grid((OW * C + RZ_SEG - 1) / RZ_SEG, nb * zstrips);
      resize_bilinear_slots_kernel<<<grid, RZ_THREADS, 0, s>>>(xi, H, W, C, OH, OW, oi);
    } else {
      dim3 grid((OW * C + 256 * RESIZE_U - 1) / (256 * RESIZE_U), nb * strips);
      resize_bilinear_strip_kernel<<<grid, 256, 0, s>>>(xi, H, W, C, OH, OW, oi);
    }
    b0 += nb;
  }
}

static void launch_roll_impl(const float* x, int B, int H, int W, int C, int oy, int ox, const int* jit, float* out, cudaStream_t s) {
  if ((size_t)B * H * W * C == 0) return;
  int b0 = 0;
  while (b0 < B) {
    const int nb = std::min(B - b0, std::max(1, 65535 / H));
    dim3 grid(std::max(1, std::min(W * C / 2048, 16)), nb * H);
    roll_kernel<<<grid, 256, 0, s>>>(x + (size_t)b0 * H * W * C, nb, H, W, C, oy, ox, jit, out + (size_t)b0 * H * W * C);
    b0 += nb;
  }
}

void launch_roll(const float* x, int B, int H, int W, int C, int oy, int ox, float* out, cudaStream_t s) {
  launch_roll_impl(x, B, H, W, C, oy, ox, nullptr, out, s);
}

// ---- host side ----------------------------------------------------------------------------------------
#define CUD(ctx, expr)                                                                            \
  do {                                                                                            \
    cudaError_t e_ = (expr);                                                                      \
    if (e_ != cudaSuccess) return (ctx)->fail(DCV_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
  } while (0)
#define TRYD(expr) do { int r_ = (expr); if (r_) return r_; } while (0)

namespace {

struct LossSpec {
  std::vector<float> coeff;      // per net layer, 0 = not a loss layer
  int deepest = -1;
};

int parse_losses(dcv_ctx* ctx, const int* layers, const float* coeffs, int nl, LossSpec* out) {
  if (!layers || !coeffs || nl <= 0) return ctx->fail(DCV_ERR_ARG, "dream: no loss layers given");
  out->coeff.assign(ctx->net.size(), 0.f);
  for (int k = 0; k < nl; ++k) {
    if (layers[k] < 0 || layers[k] >= (int)ctx->net.size()) return ctx->fail(DCV_ERR_ARG, "dream: loss layer %d out of range", layers[k]);
    if (ctx->net[layers[k]].kind != LK_CONV && ctx->net[layers[k]].kind != LK_POOL)
      return ctx->fail(DCV_ERR_UNSUPPORTED, "dream: loss layers must be conv or pool layers");
    out->coeff[layers[k]] += coeffs[k];
    out->deepest = std::max(out->deepest, layers[k]);
  }
  return 0;
}

// Backward pass of loss = sum_l coeff_l * mean(act_l^2) through the current tape; leaves dloss/dinput
// (B,H,W,C0) in *grad (one of ctx->dream[4], ctx->dream[5]).  If loss_out != null the loss value is computed too.
int dream_backward(dcv_ctx* ctx, const LossSpec& ls, float** grad, double* loss_out) {
  const int B = ctx->in_B;
  size_t max_elems = (size_t)B * ctx->in_H * ctx->in_W * ctx->in_C;
  for (int j = 0; j <= ls.deepest; ++j) max_elems = std::max(max_elems, (size_t)ctx->net[j].B * ctx->net[j].H * ctx->net[j].W * ctx->net[j].C);
  TRYD(ctx->dream[4].ensure(ctx, max_elems * sizeof(float)));
  TRYD(ctx->dream[5].ensure(ctx, max_elems * sizeof(float)));
  float* bufs[2] = {(float*)ctx->dream[4].p, (float*)ctx->dream[5].p};
  int cur = 0;
  bool has_g = false;
  double loss = 0.0;
  for (int j = ls.deepest; j >= 0; --j) {
    const Layer& l = ctx->net[j];
    const size_t n = (size_t)l.B * l.H * l.W * l.C;
    const float coeff = ls.coeff[j];
    if (coeff != 0.f && loss_out) {
      const int np = deprocess_partials(n);
      TRYD(ctx->scratch.ensure(ctx, (size_t)np * 2 * sizeof(double)));
      launch_sum_stats((const float*)l.act.p, n, (double*)ctx->scratch.p, np, ctx->stream); ctx->launches++;
      std::vector<double> h((size_t)np * 2);
      CUD(ctx, cudaMemcpyAsync(h.data(), ctx->scratch.p, h.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
      CUD(ctx, cudaStreamSynchronize(ctx->stream));
      double q = 0.0;
      for (int k = 0; k < np; ++k) q += h[2 * k + 1];
      loss += (double)coeff * q / (double)n;
    }
    const float scale = coeff != 0.f ? 2.f * coeff / (float)n : 0.f;
    if (l.kind == LK_CONV) {
      if (!has_g && scale == 0.f) continue;          // nothing flows yet (layers above the deepest loss)
      {
        ProfScope ps(ctx, DCV_K_OTHER, 0.0, 12.0 * n);
        dream_prep_kernel<<<grid_for(n, 256), 256, 0, ctx->stream>>>(bufs[cur], (const float*)l.act.p, n, scale, has_g ? 1 : 0, l.relu);
        ctx->launches++;
      }
      if (ctx->fp32_tc && l.wd) {
        // conv backward-data on the tensor cores, fp32-accurate (split operands, register accumulation)
        TRYD(tc_prepare_backward_exact(ctx, ctx->net[j]));
        TRYD(tc_conv_bwd_exact(ctx, l, bufs[cur], l.B, l.H, l.W, bufs[cur ^ 1], 0));
      } else {
        ConvEpilogue e;                                  // conv backward-data: no bias, no ReLU
        ProfScope ps(ctx, DCV_K_CONV_FP32, 2.0 * 9 * l.B * l.H * l.W * (double)l.cin * l.cout, 4.0 * l.B * l.H * l.W * ((double)l.cin + l.cout));
        launch_conv3x3_fp32(bufs[cur], l.wd, l.B, l.H, l.W, l.cout, l.cin, e, bufs[cur ^ 1], ctx->stream); ctx->launches++;
      }
      cur ^= 1;
      has_g = true;
    } else if (l.kind == LK_POOL) {
      if (!has_g && scale == 0.f) continue;
      {
        ProfScope ps(ctx, DCV_K_OTHER, 0.0, 12.0 * n);
        dream_prep_kernel<<<grid_for(n, 256), 256, 0, ctx->stream>>>(bufs[cur], (const float*)l.act.p, n, scale, has_g ? 1 : 0, 0);
        ctx->launches++;
      }
      // max-pool backward = unpool by switch; trailing row/column of an odd input gets zero gradient
      ProfScope ps(ctx, DCV_K_UNPOOL, 0.0, (double)l.B * l.in_H * l.in_W * l.C * (4.0 + 5.0 / 4.0));
      launch_unpool_switch_items(bufs[cur], (const uint8_t*)l.sw.p, l.B, l.B, l.in_H, l.in_W, l.C, 0, bufs[cur ^ 1], ctx->stream);
      ctx->launches++;
      cur ^= 1;
      has_g = true;
    } else {
      return ctx->fail(DCV_ERR_UNSUPPORTED, "dream: only conv/pool layers are supported below the loss layers");
    }
  }
  if (!has_g) return ctx->fail(DCV_ERR_ARG, "dream: all loss coefficients are zero");
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "dream backward: %s", cudaGetErrorString(e));
  *grad = bufs[cur];
  if (loss_out) *loss_out = loss;
  return 0;
}

// ---- bf16 / tensor-core variant: the tape lives in the padded-flat bf16 layout of conv_tc.cu -------------
// Forward: layer 0 (3 input channels, K = 27) stays on the FFMA kernel and is converted; every other conv is
// tc_conv_fwd (tcgen05, bias + ReLU epilogue), pools are tc_pool.  x is a device pointer (B,H,W,C0) fp32.
bool dream_first_on_tc(const dcv_ctx* ctx) {
  static const int first_tc = getenv("DCV_DREAM_FIRST_TC") ? atoi(getenv("DCV_DREAM_FIRST_TC")) : 1;
  return first_tc && !ctx->net.empty() && tc_first_layer_ok(ctx->net[0]);
}

// x: the image the forward pass sees; with jit != null (and the first layer on the tensor cores) x is the UN-rolled
// image and the jitter roll is applied by the first layer's gather
int dream_forward_bf16(dcv_ctx* ctx, const float* x, int B, int H, int W, int upto, const int* jit = nullptr) {
  if (ctx->net.empty() || ctx->net[0].kind != LK_CONV) return ctx->fail(DCV_ERR_STATE, "dream: network must start with a conv layer");
  ctx->tape_upto = -1;                      // the fp32 projection tape is not valid after this
  Layer& l0 = ctx->net[0];
  ctx->in_B = B; ctx->in_H = H; ctx->in_W = W; ctx->in_C = l0.cin;
  TRYD(l0.act_h.ensure(ctx, (size_t)B * tc_padded_elems(H, W, l0.cout) * sizeof(__nv_bfloat16)));
  if (dream_first_on_tc(ctx)) {
    // K = 27: im2col rows + one un-shifted tcgen05 GEMM writing the bf16 tape directly
    TRYD(tc_conv_first(ctx, l0, x, B, H, W, (__nv_bfloat16*)l0.act_h.p, jit));
  } else {
    if (jit) return ctx->fail(DCV_ERR_STATE, "dream: fused jitter needs the tensor-core first layer");
    TRYD(l0.act.ensure(ctx, (size_t)B * H * W * l0.cout * sizeof(float)));
    {
      ConvEpilogue e; e.bias = l0.bias; e.relu = l0.relu;
      ProfScope ps(ctx, DCV_K_CONV_FP32, 2.0 * 9 * B * H * W * (double)l0.cin * l0.cout, 4.0 * B * H * W * ((double)l0.cin + l0.cout));
      launch_conv3x3_fp32(x, l0.w, B, H, W, l0.cin, l0.cout, e, (float*)l0.act.p, ctx->stream); ctx->launches++;
    }
    TRYD(tc_pad_convert(ctx, (const float*)l0.act.p, B, H, W, l0.cout, (__nv_bfloat16*)l0.act_h.p));
  }
  l0.B = B; l0.H = H; l0.W = W; l0.C = l0.cout;
  int ch = H, cw = W, cc = l0.cout;
  for (int j = 1; j <= upto; ++j) {
    Layer& l = ctx->net[j];
    const __nv_bfloat16* prev = (const __nv_bfloat16*)ctx->net[j - 1].act_h.p;
    if (l.kind == LK_CONV) {
      if (l.cin != cc) return ctx->fail(DCV_ERR_ARG, "layer %d: expects %d input channels, got %d", j, l.cin, cc);
      TRYD(tc_prepare_forward(ctx, l));
      TRYD(l.act_h.ensure(ctx, (size_t)B * tc_padded_elems(ch, cw, l.cout) * sizeof(__nv_bfloat16)));
      TRYD(tc_conv_fwd(ctx, l, prev, B, ch, cw, (__nv_bfloat16*)l.act_h.p));
      cc = l.cout;
    } else if (l.kind == LK_POOL) {
      const int oh = ch / 2, ow = cw / 2;
      if (oh == 0 || ow == 0) return ctx->fail(DCV_ERR_ARG, "layer %d: input %dx%d too small to pool", j, ch, cw);
      TRYD(l.act_h.ensure(ctx, (size_t)B * tc_padded_elems(oh, ow, cc) * sizeof(__nv_bfloat16)));
      TRYD(l.sw.ensure(ctx, (size_t)B * oh * ow * cc));
      TRYD(tc_pool(ctx, prev, B, ch, cw, cc, (__nv_bfloat16*)l.act_h.p, (uint8_t*)l.sw.p));
      l.in_H = ch; l.in_W = cw;
      ch = oh; cw = ow;
    } else {
      return ctx->fail(DCV_ERR_UNSUPPORTED, "dream: only conv/pool layers are supported");
    }
    l.B = B; l.H = ch; l.W = cw; l.C = cc;
  }
  return 0;
}

int dream_backward_bf16(dcv_ctx* ctx, const LossSpec& ls, float** grad, double* loss_out) {
  const int B = ctx->in_B;
  size_t max_pad = (size_t)B * tc_padded_elems(ctx->in_H, ctx->in_W, ctx->in_C);
  for (int j = 0; j <= ls.deepest; ++j) max_pad = std::max(max_pad, (size_t)B * tc_padded_elems(ctx->net[j].H, ctx->net[j].W, ctx->net[j].C));
  TRYD(ctx->dream[4].ensure(ctx, max_pad * sizeof(__nv_bfloat16)));
  TRYD(ctx->dream[5].ensure(ctx, max_pad * sizeof(__nv_bfloat16)));
  TRYD(ctx->ws[0].ensure(ctx, (size_t)B * ctx->in_H * ctx->in_W * ctx->in_C * sizeof(float)));
  __nv_bfloat16* hb[2] = {(__nv_bfloat16*)ctx->dream[4].p, (__nv_bfloat16*)ctx->dream[5].p};
  static const int fuse_prep = getenv("DCV_DREAM_FUSE_PREP") ? atoi(getenv("DCV_DREAM_FUSE_PREP")) : 1;
  int cur = 0;
  bool has_g = false;
  bool prepped = false;              // hb[cur] already carries this layer's mask and loss seed (folded into its producer)
  double loss = 0.0;
  for (int j = ls.deepest; j >= 0; --j) {
    const Layer& l = ctx->net[j];
    const size_t n = (size_t)l.B * l.H * l.W * l.C;                          // real elements
    const size_t elems = (size_t)l.B * tc_padded_elems(l.H, l.W, l.C);      // padded bf16 elements
    const float coeff = ls.coeff[j];
    if (coeff != 0.f && loss_out) {
      const int np = tc_sumsq_partials(elems);
      TRYD(ctx->scratch.ensure(ctx, (size_t)np * 2 * sizeof(double)));
      TRYD(tc_sumsq(ctx, (const __nv_bfloat16*)l.act_h.p, elems, (double*)ctx->scratch.p, np));
      std::vector<double> h((size_t)np * 2);
      CUD(ctx, cudaMemcpyAsync(h.data(), ctx->scratch.p, h.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
      CUD(ctx, cudaStreamSynchronize(ctx->stream));
      double q = 0.0;
      for (int k = 0; k < np; ++k) q += h[2 * k + 1];
      loss += (double)coeff * q / (double)n;
    }
    const float scale = coeff != 0.f ? 2.f * coeff / (float)n : 0.f;
    if (!has_g && scale == 0.f) continue;
    const int prep_relu = l.kind == LK_CONV ? l.relu : 0;
    if (!prepped && (scale != 0.f || prep_relu))     // a pool that is not a loss layer passes the gradient through unchanged
      TRYD(tc_dream_prep(ctx, hb[cur], (const __nv_bfloat16*)l.act_h.p, elems, scale, has_g ? 1 : 0, prep_relu));
    prepped = false;
    // the same step for the layer below, folded into the kernel that produces its gradient
    TcPrep below{nullptr, 0.f, 0};
    const TcPrep* fold = nullptr;
    if (fuse_prep && j > 0) {
      const Layer& lo = ctx->net[j - 1];
      const size_t n_lo = (size_t)lo.B * lo.H * lo.W * lo.C;
      below.act = (const __nv_bfloat16*)lo.act_h.p;
      below.scale = ls.coeff[j - 1] != 0.f ? 2.f * ls.coeff[j - 1] / (float)n_lo : 0.f;
      below.relu = lo.kind == LK_CONV ? lo.relu : 0;
      if (below.scale != 0.f || below.relu) fold = &below;
    }
    if (l.kind == LK_CONV) {
      if (j == 0) {
        TRYD(tc_conv(ctx, l, hb[cur], l.B, l.H, l.W, nullptr, (float*)ctx->ws[0].p, 0));
      } else if (ctx->net[j - 1].kind == LK_POOL && ls.coeff[j - 1] == 0.f && l.tc.n_pad % 256 == 0 &&
                 (long long)l.B * (l.H + 1) * (l.W + 1) >= 256LL * ctx->num_sms / 2) {
        // max-pool backward of the layer below fused into this conv's epilogue (no seed to add at that pool)
        const Layer& pl = ctx->net[j - 1];
        TcUnpool up{(const uint8_t*)pl.sw.p, pl.B, pl.C, pl.in_H, pl.in_W};
        TRYD(tc_conv(ctx, l, hb[cur], l.B, l.H, l.W, hb[cur ^ 1], nullptr, 0, &up));
        cur ^= 1;
        --j;
      } else {
        TRYD(tc_conv(ctx, l, hb[cur], l.B, l.H, l.W, hb[cur ^ 1], nullptr, 0, nullptr, fold));
        cur ^= 1;
        prepped = fold != nullptr;
      }
    } else {
      TRYD(tc_unpool(ctx, hb[cur], (const uint8_t*)l.sw.p, l.B, l.B, l.H, l.W, l.in_H, l.in_W, l.C, hb[cur ^ 1], fold));
      cur ^= 1;
      prepped = fold != nullptr;
    }
    has_g = true;
  }
  if (!has_g) return ctx->fail(DCV_ERR_ARG, "dream: all loss coefficients are zero");
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "dream backward (bf16): %s", cudaGetErrorString(e));
  *grad = (float*)ctx->ws[0].p;
  if (loss_out) *loss_out = loss;
  return 0;
}

// one jittered ascent step on the device image img (B,H,W,C)
int dream_iter(dcv_ctx* ctx, float* img, int B, int H, int W, int C, const LossSpec& ls, float step, const int* jit_dev,
               double* loss_out, int precision) {
  const size_t n = (size_t)B * H * W * C;
  const bool fused_roll = precision == DCV_BF16 && dream_first_on_tc(ctx);   // the first layer's gather applies the jitter
  float* x = img;
  if (!fused_roll) {
    TRYD(ctx->dream[3].ensure(ctx, n * sizeof(float)));
    x = (float*)ctx->dream[3].p;
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 8.0 * n);
    launch_roll_impl(img, B, H, W, C, 0, 0, jit_dev, x, ctx->stream); ctx->launches++;
  }
  float* g = nullptr;
  if (precision == DCV_BF16) {
    struct Latency { dcv_ctx* c; Latency(dcv_ctx* c_) : c(c_) { c->latency_tiles = 1; } ~Latency() { c->latency_tiles = 0; } } lat(ctx);
    TRYD(dream_forward_bf16(ctx, x, B, H, W, ls.deepest, fused_roll ? jit_dev : nullptr));
    TRYD(dream_backward_bf16(ctx, ls, &g, loss_out));
  } else {
    TRYD(forward_impl(ctx, x, true, B, H, W, ls.deepest));
    TRYD(dream_backward(ctx, ls, &g, loss_out));
  }
  const size_t per = (size_t)H * W * C;
  const int np = deprocess_partials(per);
  TRYD(ctx->scratch.ensure(ctx, (size_t)np * 2 * sizeof(double)));
  for (int b = 0; b < B; ++b) {          // the RMS normalisation is per image
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 16.0 * per);
    launch_sum_stats(g + b * per, per, (double*)ctx->scratch.p, np, ctx->stream);
    // few, fat blocks: every block first reduces the block partials of the RMS
    dream_update_kernel<<<std::min(H, 4 * kNumSMs), 256, 0, ctx->stream>>>(img + b * per, g + b * per, (const double*)ctx->scratch.p, np,
                                                                         H, W, C, step, 0, 0, jit_dev);
    ctx->launches += 2;
  }
  bump_counter_kernel<<<1, 1, 0, ctx->stream>>>(const_cast<int*>(jit_dev)); ctx->launches++;
  return 0;
}

// ---- CUDA graph of one iteration ------------------------------------------------------------------------
// An iteration is ~80 small launches (plus two tensor-map encodes per conv); at 261x261 the host cannot issue them
// as fast as the GPU retires them.  The first iteration of an octave runs eagerly (it sizes every buffer), the
// second is captured, and the instantiated graph is replayed for the rest.  Graphs are cached per (shape, losses,
// step, precision) and dropped when any workspace buffer has been re-allocated since (ctx->alloc_epoch).
struct DreamGraphKey {
  int B, H, W, precision, deepest;
  float step;
  std::vector<float> coeff;
  bool operator==(const DreamGraphKey& o) const {
    return B == o.B && H == o.H && W == o.W && precision == o.precision && deepest == o.deepest && step == o.step && coeff == o.coeff;
  }
};
struct DreamGraph {
  DreamGraphKey key;
  cudaGraphExec_t exec = nullptr;
  unsigned long long epoch = 0;
  float* img = nullptr;
};
std::vector<DreamGraph>& graph_cache(dcv_ctx* ctx) {
  if (!ctx->dream_graphs) ctx->dream_graphs = new std::vector<DreamGraph>();
  return *static_cast<std::vector<DreamGraph>*>(ctx->dream_graphs);
}

int dream_iterations(dcv_ctx* ctx, float* img, int B, int H, int W, int C, const LossSpec& ls, float step, const int* jit_dev,
                     double* losses, int precision, int iterations) {
  static const int use_graph = getenv("DCV_DREAM_GRAPH") ? atoi(getenv("DCV_DREAM_GRAPH")) : 1;
  const bool graphable = use_graph && !losses && !ctx->prof_on && iterations >= 3;
  if (!graphable) {
    for (int it = 0; it < iterations; ++it)
      TRYD(dream_iter(ctx, img, B, H, W, C, ls, step, jit_dev, losses ? losses + it : nullptr, precision));
    return 0;
  }
  DreamGraphKey key{B, H, W, precision, ls.deepest, step, ls.coeff};
  auto& cache = graph_cache(ctx);
  DreamGraph* hit = nullptr;
  for (auto& g : cache)
    if (g.key == key && g.img == img) { hit = &g; break; }
  int done = 0;
  if (hit && hit->epoch != ctx->alloc_epoch) {            // a buffer moved: the captured pointers are stale
    cudaGraphExecDestroy(hit->exec);
    hit->exec = nullptr;
  }
  if (!hit || !hit->exec) {
    TRYD(dream_iter(ctx, img, B, H, W, C, ls, step, jit_dev, nullptr, precision));      // eager: sizes all buffers
    done = 1;
    const unsigned long long epoch = ctx->alloc_epoch;
    CUD(ctx, cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
    int r = dream_iter(ctx, img, B, H, W, C, ls, step, jit_dev, nullptr, precision);
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamEndCapture(ctx->stream, &graph);
    if (r) { if (graph) cudaGraphDestroy(graph); return r; }
    if (e != cudaSuccess || !graph) return ctx->fail(DCV_ERR_CUDA, "dream: stream capture failed: %s", cudaGetErrorString(e));
    if (ctx->alloc_epoch != epoch) { cudaGraphDestroy(graph); return ctx->fail(DCV_ERR_STATE, "dream: a buffer was re-allocated during capture"); }
    cudaGraphExec_t exec = nullptr;
    e = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "dream: cudaGraphInstantiate: %s", cudaGetErrorString(e));
    if (!hit) { cache.push_back(DreamGraph()); hit = &cache.back(); }
    hit->key = key; hit->exec = exec; hit->epoch = epoch; hit->img = img;
  }
  for (int it = done; it < iterations; ++it) {
    CUD(ctx, cudaGraphLaunch(hit->exec, ctx->stream));
    ctx->launches += ctx->net.size() * 4;       // approximate: launches inside a replayed graph are not counted one by one
  }
  return 0;
}

int dream_impl(dcv_ctx* ctx, const float* img_in, bool on_device, int B, int H, int W, const int* loss_layers, const float* coeffs,
               int nl, const int* shapes, int n_oct, int iterations, float step, const int* jitter, int precision, float* out,
               double* losses) {
  if (precision != DCV_FP32 && precision != DCV_BF16) return ctx->fail(DCV_ERR_ARG, "dcv_dream: unknown precision");
  if (!img_in || !out || B <= 0 || H <= 0 || W <= 0 || !shapes || n_oct <= 0 || iterations < 0 || (iterations > 0 && !jitter))
    return ctx->fail(DCV_ERR_ARG, "dcv_dream: bad arguments");
  if (ctx->net.empty() || ctx->net[0].kind != LK_CONV) return ctx->fail(DCV_ERR_STATE, "dcv_dream: network must start with a conv layer");
  LossSpec ls;
  TRYD(parse_losses(ctx, loss_layers, coeffs, nl, &ls));
  const int C = ctx->net[0].cin;
  if (shapes[2 * (n_oct - 1)] != H || shapes[2 * (n_oct - 1) + 1] != W)
    return ctx->fail(DCV_ERR_ARG, "dcv_dream: the last octave shape must be the image size %dx%d", H, W);
  size_t max_px = 0;
  for (int k = 0; k < n_oct; ++k) {
    if (shapes[2 * k] < 8 || shapes[2 * k + 1] < 8) return ctx->fail(DCV_ERR_ARG, "dcv_dream: octave %d is smaller than 8x8", k);
    max_px = std::max(max_px, (size_t)shapes[2 * k] * shapes[2 * k + 1]);
  }
  const size_t full = (size_t)B * H * W * C, maxn = (size_t)B * max_px * C;
  for (int k = 0; k < 3; ++k) TRYD(ctx->dream[k].ensure(ctx, std::max(full, maxn) * sizeof(float)));
  TRYD(ctx->stage.ensure(ctx, 2 * maxn * sizeof(float)));
  float* img = (float*)ctx->dream[0].p;
  float* orig = (float*)ctx->dream[1].p;
  float* shrunk = (float*)ctx->dream[2].p;
  float* tmp_a = (float*)ctx->stage.p;
  float* tmp_b = tmp_a + maxn;
  // jitter offsets live on the device behind an iteration counter (see roll_kernel)
  const size_t njit = (size_t)n_oct * iterations;
  TRYD(ctx->jit_dev.ensure(ctx, (1 + 2 * njit + 2) * sizeof(int)));
  CUD(ctx, cudaMemsetAsync(ctx->jit_dev.p, 0, sizeof(int), ctx->stream));
  if (njit) CUD(ctx, cudaMemcpyAsync((int*)ctx->jit_dev.p + 1, jitter, 2 * njit * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
  CUD(ctx, cudaMemcpyAsync(orig, img_in, full * sizeof(float), on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, ctx->stream));
  CUD(ctx, cudaMemcpyAsync(img, orig, full * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
  int ch = H, cw = W;                       // current size of img
  int sh = shapes[0], sw = shapes[1];       // size of `shrunk`
  launch_resize_bilinear(orig, B, H, W, C, sh, sw, shrunk, ctx->stream); ctx->launches++;
  for (int k = 0; k < n_oct; ++k) {
    const int oh = shapes[2 * k], ow = shapes[2 * k + 1];
    const size_t n = (size_t)B * oh * ow * C;
    if (oh != ch || ow != cw) {
      ProfScope ps(ctx, DCV_K_OTHER, 0.0, 4.0 * ((double)B * ch * cw * C + n));
      launch_resize_bilinear(img, B, ch, cw, C, oh, ow, tmp_a, ctx->stream); ctx->launches++;
      CUD(ctx, cudaMemcpyAsync(img, tmp_a, n * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
      ch = oh; cw = ow;
    }
    TRYD(dream_iterations(ctx, img, B, oh, ow, C, ls, step, (const int*)ctx->jit_dev.p, losses ? losses + (size_t)k * iterations : nullptr,
                          precision, iterations));
    // re-inject the detail lost by having worked at a lower resolution
    launch_resize_bilinear(shrunk, B, sh, sw, C, oh, ow, tmp_a, ctx->stream);          // upscaled shrunk original
    launch_resize_bilinear(orig, B, H, W, C, oh, ow, tmp_b, ctx->stream);              // original at this size
    add_diff_kernel<<<grid_for(n, 256), 256, 0, ctx->stream>>>(img, tmp_b, tmp_a, n);
    ctx->launches += 3;
    CUD(ctx, cudaMemcpyAsync(shrunk, tmp_b, n * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
    sh = oh; sw = ow;
  }
  CUD(ctx, cudaGetLastError());
  CUD(ctx, cudaMemcpyAsync(out, img, full * sizeof(float), on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, ctx->stream));
  return 0;
}

}  // namespace

void dream_release_graphs(dcv_ctx* ctx) {
  if (!ctx->dream_graphs) return;
  auto* v = static_cast<std::vector<DreamGraph>*>(ctx->dream_graphs);
  for (auto& g : *v)
    if (g.exec) cudaGraphExecDestroy(g.exec);
  delete v;
  ctx->dream_graphs = nullptr;
}

#define ENTERD(ctx)                                                       \
  if (!(ctx)) return DCV_ERR_ARG;                                          \
  CUD(ctx, cudaSetDevice((ctx)->device));

extern "C" int dcv_dream(dcv_ctx* ctx, const float* img, int B, int H, int W, const int* loss_layers, const float* coeffs, int nl,
                         const int* shapes, int n_oct, int iterations, float step, const int* jitter, int precision, float* out,
                         double* losses) {
  ENTERD(ctx);
  TRYD(dream_impl(ctx, img, false, B, H, W, loss_layers, coeffs, nl, shapes, n_oct, iterations, step, jitter, precision, out, losses));
  CUD(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_dream_dev(dcv_ctx* ctx, const float* img_dev, int B, int H, int W, const int* loss_layers, const float* coeffs,
                             int nl, const int* shapes, int n_oct, int iterations, float step, const int* jitter, int precision,
                             float* out_dev) {
  ENTERD(ctx);
  return dream_impl(ctx, img_dev, true, B, H, W, loss_layers, coeffs, nl, shapes, n_oct, iterations, step, jitter, precision, out_dev, nullptr);
}

extern "C" int dcv_dream_gradient(dcv_ctx* ctx, const float* img, int B, int H, int W, const int* loss_layers, const float* coeffs,
                                  int nl, int precision, float* grad_out, double* loss_out) {
  ENTERD(ctx);
  if (!img || !grad_out || B <= 0 || H <= 0 || W <= 0) return ctx->fail(DCV_ERR_ARG, "dcv_dream_gradient: bad arguments");
  if (precision != DCV_FP32 && precision != DCV_BF16) return ctx->fail(DCV_ERR_ARG, "dcv_dream_gradient: unknown precision");
  LossSpec ls;
  TRYD(parse_losses(ctx, loss_layers, coeffs, nl, &ls));
  float* g = nullptr;
  if (precision == DCV_BF16) {
    if (ctx->net.empty() || ctx->net[0].kind != LK_CONV) return ctx->fail(DCV_ERR_STATE, "dream: network must start with a conv layer");
    const size_t nimg = (size_t)B * H * W * ctx->net[0].cin;
    TRYD(ctx->dream[3].ensure(ctx, nimg * sizeof(float)));
    CUD(ctx, cudaMemcpyAsync(ctx->dream[3].p, img, nimg * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    TRYD(dream_forward_bf16(ctx, (const float*)ctx->dream[3].p, B, H, W, ls.deepest));
    TRYD(dream_backward_bf16(ctx, ls, &g, loss_out));
  } else {
    TRYD(forward_impl(ctx, img, false, B, H, W, ls.deepest));
    TRYD(dream_backward(ctx, ls, &g, loss_out));
  }
  CUD(ctx, cudaMemcpyAsync(grad_out, g, (size_t)B * H * W * ctx->in_C * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  CUD(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_resize_bilinear(dcv_ctx* ctx, const float* x, int B, int H, int W, int C, int OH, int OW, float* out) {
  ENTERD(ctx);
  if (!x || !out || B <= 0 || H <= 0 || W <= 0 || C <= 0 || OH <= 0 || OW <= 0) return ctx->fail(DCV_ERR_ARG, "dcv_resize_bilinear: bad arguments");
  const size_t nin = (size_t)B * H * W * C, nout = (size_t)B * OH * OW * C;
  const size_t o1 = (nin * 4 + 255) & ~(size_t)255;
  TRYD(ctx->stage.ensure(ctx, o1 + nout * 4));
  char* base = (char*)ctx->stage.p;
  CUD(ctx, cudaMemcpyAsync(base, x, nin * 4, cudaMemcpyHostToDevice, ctx->stream));
  {
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 4.0 * (nin + nout));
    launch_resize_bilinear((const float*)base, B, H, W, C, OH, OW, (float*)(base + o1), ctx->stream); ctx->launches++;
  }
  CUD(ctx, cudaGetLastError());
  CUD(ctx, cudaMemcpyAsync(out, base + o1, nout * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CUD(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int dcv_roll(dcv_ctx* ctx, const float* x, int B, int H, int W, int C, int oy, int ox, float* out) {
  ENTERD(ctx);
  if (!x || !out || B <= 0 || H <= 0 || W <= 0 || C <= 0) return ctx->fail(DCV_ERR_ARG, "dcv_roll: bad arguments");
  const size_t n = (size_t)B * H * W * C;
  const size_t o1 = (n * 4 + 255) & ~(size_t)255;
  TRYD(ctx->stage.ensure(ctx, o1 + n * 4));
  char* base = (char*)ctx->stage.p;
  CUD(ctx, cudaMemcpyAsync(base, x, n * 4, cudaMemcpyHostToDevice, ctx->stream));
  {
    ProfScope ps(ctx, DCV_K_OTHER, 0.0, 8.0 * n);
    launch_roll((const float*)base, B, H, W, C, oy, ox, (float*)(base + o1), ctx->stream); ctx->launches++;
  }
  CUD(ctx, cudaGetLastError());
  CUD(ctx, cudaMemcpyAsync(out, base + o1, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CUD(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}
