// "Strip" kernel for the 64-channel tail of a projection (included by conv_tc.cu only): the dx-grouped GEMM of
// tc_dxg.cuh walked DOWN the image instead of along it, so that everything a tile shares with the tile below it stays
// in shared memory — and with that the two neighbours of the 64 -> 64 conv can be fused into it:
//
//   UNPOOL  the conv's input is the switch-driven unpool of a pooled tensor (DPooling.down, deepdream.py:191-209):
//           producer warps expand pooled rows + switch bytes straight into the swizzled A tiles, the 4x larger
//           unpooled tensor is never written to or read from HBM;
//   LAST    the conv's output feeds the last backward conv (64 -> <= 3 channels, deepdream.py:102-111 of block1_conv1):
//           the bf16 output tile stays in shared memory as the A operand of a second GEMM (all 9 taps stacked along N,
//           the "GEMM + col2im" of tc_last.cuh), three consecutive P tiles give one finished row of the projection;
//           the conv's 64-channel output never leaves the SM.
//
// Geometry.  The padded-flat tensor of P items x (h+1) x (w+1) rows is one tall image of T = P*(h+1) "tall rows" of
// width wp = w+1 (pad rows / columns are zero, so a 3x3 'same' conv needs no other halo).  A tile is 120 consecutive
// positions of ONE tall row (4 lane quarters x 30, as in tc_dxg.cuh); a strip is a column block (118 apart) times R
// consecutive tall rows, processed top to bottom by one CTA.  The input rows live in a ring of IN_SLOTS shared-memory
// A tiles: tile t multiplies rows t-1, t, t+1 (dy = -1, 0, +1) and only row t+1 is new - one 16 KB A tile per tile
// instead of three.  Strips cost two extra input rows (and, with LAST, two extra tiles) at their top and bottom.
//
// Roles: warp 0 TMA (weights; input rows unless UNPOOL), warp 1 MMA issue, then 8 conv epilogue warps (TMEM lane quarter
// x 32 output channels, two 16-channel passes), four "tail" warps (LAST: P tiles out of TMEM, col2im, fp32 stores) and
// eight unpool producer warps (UNPOOL).  Every stage hands over through mbarriers only; measured with the col2im inside
// the conv epilogue's lock-step (two 512-thread barriers per tile) the fused kernel was no faster than the two kernels it
// replaces.  Warp split measured on the 512-projection step (epilogue / producer warps): 16/4 12.1 ms, 16/8 12.0, 8/4
// 11.9, 8/16 11.9, 4/16 11.7, 8/8 and 4/8 11.3 (profiles/build_variants.sh): the producers need the issue slots.
#pragma once
#include "tc_kernel.cuh"

namespace dcv {
namespace tc {

constexpr int STRIP_N = 192;                 // three dx taps x 64 output channels
constexpr int STRIP_SEG = 30;                // output positions per lane quarter
constexpr int STRIP_POS = 120;               // positions per tile
constexpr int STRIP_CB = 118;                // column-block pitch (LAST loses one position at each end of a tile)
constexpr int STRIP_IN_SLOTS = 4;
#ifndef STRIP_EPI_WARPS_CFG
#define STRIP_EPI_WARPS_CFG 8
#endif
#ifndef STRIP_PROD_WARPS_CFG
#define STRIP_PROD_WARPS_CFG 8
#endif
constexpr int STRIP_EPI_WARPS = STRIP_EPI_WARPS_CFG;    // 16: one 16-channel group per warp; 8: two groups per warp, one after the other
constexpr int STRIP_EPI_PASSES = 16 / STRIP_EPI_WARPS;  // 16-channel groups per epilogue warp
constexpr int STRIP_PROD_WARPS = STRIP_PROD_WARPS_CFG;  // 4: four pooled pixels per thread; 8: two
constexpr int STRIP_NPX = 16 / STRIP_PROD_WARPS;        // pooled pixels per producer thread
constexpr int STRIP_PXSTEP = 4 * STRIP_PROD_WARPS;      // distance between a thread's pooled pixels
constexpr int STRIP_W_BYTES = 3 * STRIP_N * 128;        // resident conv weights: [dy][192][64] bf16
constexpr int STRIP_WL_BYTES = 32 * 128;                // resident all-tap weights of the last layer: [32][64] bf16
#ifndef STRIP_BACKOFF_NS
#define STRIP_BACKOFF_NS 0
#endif
// how the waiting roles wait: 0 = try_wait with a suspend hint, N > 0 = test + nanosleep(N) (tc_ptx.cuh)
__device__ __forceinline__ void strip_wait(uint64_t* bar, uint32_t parity) {
  if constexpr (STRIP_BACKOFF_NS > 0) mbar_wait_backoff(bar, parity, STRIP_BACKOFF_NS);
  else mbar_wait_parked(bar, parity);
}
constexpr int STRIP_PS = 29;                            // P row pitch in floats (27 used, odd: conflict-free)
constexpr int STRIP_P_BYTES = 128 * STRIP_PS * 4;
constexpr int STRIP_P_SLOTS = 3;

constexpr int STRIP_TAIL_WARPS = 4;
// tail warps: 27 floats of a P row to shared memory / the three channels of tap dx of a P row, offsets as immediates
template <int Q>
__device__ __forceinline__ void strip_store27(uint32_t dst, const uint32_t (&v)[32]) {
  sts32_at<4 * Q>(dst, v[Q]);
  if constexpr (Q + 1 < 27) strip_store27<Q + 1>(dst, v);
}
template <int DX>
__device__ __forceinline__ void strip_tap3(uint32_t ps, float& a0, float& a1, float& a2) {
  a0 += lds_f32_at<(DX * STRIP_PS + DX * 3) * 4>(ps);
  a1 += lds_f32_at<(DX * STRIP_PS + DX * 3 + 1) * 4>(ps);
  a2 += lds_f32_at<(DX * STRIP_PS + DX * 3 + 2) * 4>(ps);
}
__host__ __device__ constexpr int strip_threads(bool unpool, bool last) {
  return 64 + 32 * STRIP_EPI_WARPS + (last ? 32 * STRIP_TAIL_WARPS : 0) + (unpool ? 32 * STRIP_PROD_WARPS : 0);
}
constexpr int STRIP_LUT_BYTES = 4 * 256 * 8;            // UNPOOL: switch bytes -> lane masks, per quad corner
__host__ __device__ constexpr int strip_smem_bytes(bool unpool, bool last) {
  return STRIP_W_BYTES + STRIP_IN_SLOTS * A_TILE_BYTES + (unpool ? STRIP_LUT_BYTES : 0) +
         (last ? STRIP_WL_BYTES + 2 * A_TILE_BYTES + STRIP_P_SLOTS * STRIP_P_BYTES : A_TILE_BYTES) + 1024 + 256;
}

struct StripParams {
  int rows;                  // P * (h+1) * (w+1)
  int h, w;                  // image size at the conv (unpooled size with UNPOOL)
  int T;                     // tall rows: P * (h+1)
  int n_cb;                  // column blocks per tall row
  int R;                     // tall rows per strip
  int total_strips;
  int relu;                  // activation of the 64 -> 64 conv
  // UNPOOL: pooled padded-flat input (P, ph+1, pw+1, 64) bf16 and the pool layer's switches (Bsw, ph, pw, C) bytes
  const __nv_bfloat16* pooled;
  const uint8_t* sw;
  int Bsw, C, ph, pw;
  // LAST: the 64 -> cout3 conv; fp32 compact output (P, h, w, cout3)
  int cout3, relu_last;
  float* out;
};

// TMEM -> registers: this warp's 32 lanes x 8 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
}

// what every role needs to know about a strip; all roles walk the same sequence
struct StripWalk {
  int c0;        // first position of the column block
  int tfirst;    // tall row of the strip's first tile
  int nt;        // tiles in the strip
};
template <bool LAST>
__device__ __forceinline__ StripWalk strip_walk(const StripParams& p, int s) {
  StripWalk k;
  const int rc = s / p.n_cb, cb = s % p.n_cb;
  const int t0 = rc * p.R;
  const int rr = min(p.R, p.T - t0);
  // without LAST the tile is stored as it is: the last block is pulled back so that it ends with the row (the store's
  // box is min(120, wp) positions); with LAST positions beyond the row are never used
  k.c0 = LAST ? cb * STRIP_CB : min(cb * STRIP_CB, max(0, p.w + 1 - STRIP_POS));
  k.tfirst = t0 - (LAST ? 1 : 0);
  k.nt = rr + (LAST ? 2 : 0);
  return k;
}

template <bool UNPOOL, bool LAST>
__global__ void __launch_bounds__(strip_threads(UNPOOL, LAST), 1)
deconv_strip_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const __grid_constant__ CUtensorMap tmap_bl, const __grid_constant__ CUtensorMap tmap_out, StripParams p) {
  constexpr int TMEM_COLS = 512;                      // conv accumulators at columns 0 / 256 (192 wide), P accumulators at 192 / 448 (32 wide)
  constexpr int ACC_STRIDE = 256;
  constexpr int PACC_OFF = 192;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* wres = smem;                                             // [dy][192 rows][128 B]
  uint8_t* in_ring = wres + STRIP_W_BYTES;                          // IN_SLOTS A tiles (one input row of the strip each)
  uint8_t* xs_base = in_ring + STRIP_IN_SLOTS * A_TILE_BYTES;       // !LAST: staging tile of the TMA store; LAST: two X tiles
  uint8_t* wlast = xs_base + (LAST ? 2 : 1) * A_TILE_BYTES;         // LAST: [32 rows][128 B]
  float* ps_base = reinterpret_cast<float*>(wlast + (LAST ? STRIP_WL_BYTES : 0));
  uint2* lut = reinterpret_cast<uint2*>(reinterpret_cast<uint8_t*>(ps_base) + (LAST ? STRIP_P_SLOTS * STRIP_P_BYTES : 0));   // UNPOOL: [corner][256]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(lut) + (UNPOOL ? STRIP_LUT_BYTES : 0));
  uint64_t* in_full = bars;                           // [IN_SLOTS]
  uint64_t* in_empty = in_full + STRIP_IN_SLOTS;      // [IN_SLOTS]
  uint64_t* tfull_bar = in_empty + STRIP_IN_SLOTS;    // [2] conv accumulator ready
  uint64_t* tempty_bar = tfull_bar + 2;               // [2] conv accumulator drained
  uint64_t* xfull_bar = tempty_bar + 2;               // [2] X tile written by the epilogue
  uint64_t* xempty_bar = xfull_bar + 2;               // [2] X tile consumed by the second GEMM
  uint64_t* pfull_bar = xempty_bar + 2;               // [2] P accumulator ready
  uint64_t* pempty_bar = pfull_bar + 2;               // [2] P accumulator drained
  uint64_t* wres_bar = pempty_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wres_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_b);
    if (!UNPOOL) tma_prefetch_desc(&tmap_a);
    if (LAST) tma_prefetch_desc(&tmap_bl); else tma_prefetch_desc(&tmap_out);
    for (int s = 0; s < STRIP_IN_SLOTS; ++s) { mbar_init(&in_full[s], UNPOOL ? STRIP_PROD_WARPS : 1); mbar_init(&in_empty[s], 1); }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], STRIP_EPI_WARPS);
      mbar_init(&xfull_bar[b], STRIP_EPI_WARPS); mbar_init(&xempty_bar[b], 1);
      mbar_init(&pfull_bar[b], 1); mbar_init(&pempty_bar[b], STRIP_TAIL_WARPS);
    }
    mbar_init(wres_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch_dependents();

  const int wp = p.w + 1;

  if (warp == 0) {
    // ===== TMA: resident weights once; without UNPOOL also one input row (four 32-row boxes, 30 apart) per tile =====
    if (lane == 0) {
      mbar_arrive_expect_tx(wres_bar, (uint32_t)(STRIP_W_BYTES + (LAST ? STRIP_WL_BYTES : 0)));
      for (int dy = 0; dy < 3; ++dy) tma_load_2d(wres + dy * STRIP_N * 128, &tmap_b, wres_bar, 0, dy * STRIP_N);
      if (LAST) tma_load_2d(wlast, &tmap_bl, wres_bar, 0, 0);
      if (!UNPOOL) {
        griddep_wait();
        uint32_t gi = 0;                              // input rows loaded by this CTA so far
        for (int s = blockIdx.x; s < p.total_strips; s += gridDim.x) {
          const StripWalk k = strip_walk<LAST>(p, s);
          for (int i = 0; i < k.nt + 2; ++i, ++gi) {
            const uint32_t slot = gi % STRIP_IN_SLOTS, use = gi / STRIP_IN_SLOTS;
            mbar_wait(&in_empty[slot], (use & 1) ^ 1);
            uint8_t* dst = in_ring + slot * A_TILE_BYTES;
            mbar_arrive_expect_tx(&in_full[slot], (uint32_t)A_TILE_BYTES);
            const int row0 = (k.tfirst - 1 + i) * wp + k.c0 - 1;
#pragma unroll
            for (int q = 0; q < 4; ++q) tma_load_2d(dst + q * 32 * 128, &tmap_a, &in_full[slot], 0, row0 + STRIP_SEG * q);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issue: conv tile g (12 MMAs, N = 192) then, with LAST, the second GEMM of tile g-1 (4 MMAs, N = 32) =====
    constexpr uint32_t idesc = make_idesc_bf16(BLOCK_M, STRIP_N);
    constexpr uint32_t idesc_l = make_idesc_bf16(BLOCK_M, 32);
    mbar_wait(wres_bar, 0);
    tc_fence_after();
    const uint64_t bl_desc = make_sw128_kmajor_desc(smem_u32(wlast), 0);
    auto second_gemm = [&](uint32_t g) {              // P(g) = X(g) * Wall
      const uint32_t b = g & 1, use = g >> 1;
      mbar_wait(&xfull_bar[b], use & 1);
      mbar_wait(&pempty_bar[b], (use & 1) ^ 1);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t a_desc = make_sw128_kmajor_desc(smem_u32(xs_base + b * A_TILE_BYTES), 0);
        const uint32_t d = tmem_base + (uint32_t)(b * ACC_STRIDE + PACC_OFF);
#pragma unroll
        for (int kk = 0; kk < BLOCK_K / 16; ++kk) umma_bf16(d, a_desc + (uint64_t)(2 * kk), bl_desc + (uint64_t)(2 * kk), idesc_l, kk != 0);
        umma_commit(&xempty_bar[b]);
        umma_commit(&pfull_bar[b]);
      }
      __syncwarp();
    };
    uint32_t gi0 = 0;                                 // global index of the strip's first input row
    uint32_t g = 0;                                   // tiles issued by this CTA so far
    for (int s = blockIdx.x; s < p.total_strips; s += gridDim.x) {
      const StripWalk k = strip_walk<LAST>(p, s);
      for (int j = 0; j < k.nt; ++j, ++g) {
        const uint32_t buf = g & 1, use = g >> 1;
        mbar_wait(&tempty_bar[buf], (use & 1) ^ 1);
        for (int dy = (j == 0 ? 0 : 2); dy < 3; ++dy) {   // rows j, j+1 arrived with the previous tiles
          const uint32_t gi = gi0 + j + dy;
          mbar_wait(&in_full[gi % STRIP_IN_SLOTS], (gi / STRIP_IN_SLOTS) & 1);
        }
        tc_fence_after();
        if (elect_one()) {
          const uint32_t d_tmem = tmem_base + buf * ACC_STRIDE;
#pragma unroll
          for (int dy = 0; dy < 3; ++dy) {
            const uint32_t gi = gi0 + j + dy;
            const uint64_t a_desc = make_sw128_kmajor_desc(smem_u32(in_ring + (gi % STRIP_IN_SLOTS) * A_TILE_BYTES), 0);
            const uint64_t b_desc = make_sw128_kmajor_desc(smem_u32(wres + dy * STRIP_N * 128), 0);
#pragma unroll
            for (int kk = 0; kk < BLOCK_K / 16; ++kk)
              umma_bf16(d_tmem, a_desc + (uint64_t)(2 * kk), b_desc + (uint64_t)(2 * kk), idesc, (dy | kk) != 0);
          }
          umma_commit(&tfull_bar[buf]);
          umma_commit(&in_empty[(gi0 + j) % STRIP_IN_SLOTS]);          // row j is not needed below this tile
          if (j == k.nt - 1) {                                          // bottom of the strip: its last two rows too
            umma_commit(&in_empty[(gi0 + j + 1) % STRIP_IN_SLOTS]);
            umma_commit(&in_empty[(gi0 + j + 2) % STRIP_IN_SLOTS]);
          }
        }
        __syncwarp();
        if (LAST && g > 0) second_gemm(g - 1);
      }
      gi0 += (uint32_t)k.nt + 2;
    }
    if (LAST && g > 0) second_gemm(g - 1);
  } else if (warp < 2 + STRIP_EPI_WARPS) {
    // ===== conv epilogue warps: accumulator -> dx combination (lane shuffles) -> ReLU -> bf16 tile in shared memory;
    // !LAST: one TMA store per tile; LAST: the tile is the A operand of the second GEMM =====
    const int lg = warp & 3;                                  // TMEM lane quarter this warp may read
    const int cc = ((warp - 2) >> 2) * 16 * STRIP_EPI_PASSES; // its first output channel (16 per pass)
    const float relu_lo = p.relu ? 0.f : -INFINITY;
    const bool issuer = (warp == 2) && (lane == 0);
    const int row_local = STRIP_SEG * lg + lane;
    const int jch = cc >> 3;                                  // 16-byte chunk of this thread's first 8 channels
    const uint32_t xs_u32 = smem_u32(xs_base);
    uint32_t g = 0;
    for (int s = blockIdx.x; s < p.total_strips; s += gridDim.x) {
      const StripWalk k = strip_walk<LAST>(p, s);
      // a position is a pixel when its column is one (positions past the row's end belong to the next tall row and are
      // never used) and its tall row is not a pad row; the row's y is kept incrementally (no division per tile)
      const bool col_ok = (lane < STRIP_SEG) && (k.c0 + row_local < p.w);
      int tall = k.tfirst;
      int y = tall >= 0 ? tall % (p.h + 1) : p.h;
      for (int j = 0; j < k.nt; ++j, ++g, ++tall) {
        const uint32_t buf = g & 1, use = g >> 1;
        const bool is_pixel = col_ok && tall >= 0 && tall < p.T && y < p.h;
        if (tall < 0) y = 0; else if (++y > p.h) y = 0;       // y of the next tile
        strip_wait(&tfull_bar[buf], use & 1);
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(buf * ACC_STRIDE);
        uint32_t packed[STRIP_EPI_PASSES][8];
#pragma unroll
        for (int ps = 0; ps < STRIP_EPI_PASSES; ++ps) {
          const int cp = cc + 16 * ps;
          uint32_t v0[32], v1[32], v2[32];                    // 16 used
          tmem_ld16(t_row + (uint32_t)cp, v0);
          tmem_ld16(t_row + (uint32_t)(64 + cp), v1);
          tmem_ld16(t_row + (uint32_t)(128 + cp), v2);
          tmem_ld_wait();
          if (ps == STRIP_EPI_PASSES - 1) {                   // the accumulator is in registers: hand the TMEM buffer back
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar[buf]);
          }
#pragma unroll
          for (int c = 0; c < 16; c += 2) {
            float sv[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              float t = __uint_as_float(v0[c + e]);
              t += __shfl_down_sync(0xffffffffu, __uint_as_float(v1[c + e]), 1) + __shfl_down_sync(0xffffffffu, __uint_as_float(v2[c + e]), 2);
              sv[e] = fmaxf(t, relu_lo);
            }
            __nv_bfloat162 h2 = __floats2bfloat162_rn(sv[0], sv[1]);
            packed[ps][c / 2] = *reinterpret_cast<uint32_t*>(&h2);
          }
        }
        if (!is_pixel) {                                      // pad positions store zeros
#pragma unroll
          for (int ps = 0; ps < STRIP_EPI_PASSES; ++ps)
#pragma unroll
            for (int c = 0; c < 8; ++c) packed[ps][c] = 0u;
        }
        if constexpr (!LAST) {
          if (issuer) tma_store_wait_read();                  // the previous tile's store has read the staging tile
          named_barrier_sync(1, 32 * STRIP_EPI_WARPS);
          if (lane < STRIP_SEG) {
            const uint32_t rowp = xs_u32 + row_local * 128;
#pragma unroll
            for (int ps = 0; ps < STRIP_EPI_PASSES; ++ps) {
              sts128(rowp + (((jch + 2 * ps) ^ (row_local & 7)) << 4), make_uint4(packed[ps][0], packed[ps][1], packed[ps][2], packed[ps][3]));
              sts128(rowp + (((jch + 2 * ps + 1) ^ (row_local & 7)) << 4), make_uint4(packed[ps][4], packed[ps][5], packed[ps][6], packed[ps][7]));
            }
          }
          fence_proxy_async_smem();
          named_barrier_sync(1, 32 * STRIP_EPI_WARPS);
          if (issuer && tall >= 0) {
            tma_store_2d(&tmap_out, xs_base, 0, tall * wp + k.c0);
            tma_store_commit();
          }
        } else {
          // X tile (rows = positions, 128 B = 64 channels, 128-byte swizzle); no barrier among the warps: each one
          // credits the tile's mbarrier when its part is visible to the async proxy
          strip_wait(&xempty_bar[buf], (use & 1) ^ 1);
          if (lane < STRIP_SEG) {
            const uint32_t rowp = xs_u32 + buf * A_TILE_BYTES + row_local * 128;
#pragma unroll
            for (int ps = 0; ps < STRIP_EPI_PASSES; ++ps) {
              sts128(rowp + (((jch + 2 * ps) ^ (row_local & 7)) << 4), make_uint4(packed[ps][0], packed[ps][1], packed[ps][2], packed[ps][3]));
              sts128(rowp + (((jch + 2 * ps + 1) ^ (row_local & 7)) << 4), make_uint4(packed[ps][4], packed[ps][5], packed[ps][6], packed[ps][7]));
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&xfull_bar[buf]);
        }
      }
    }
    if constexpr (!LAST) {
      if (issuer) tma_store_wait_read();
    }
  } else if (LAST && warp < 2 + STRIP_EPI_WARPS + STRIP_TAIL_WARPS) {
    // ===== tail warps: P(g) = X(g) * Wall out of TMEM into a ring of three shared-memory tiles, then the finished
    // projection row of tile g-1:  out[x][c] = act( sum_{dy,dx} P(g-2+dy)[x+dx-1][(dy*3+dx)*cout3 + c] ) =====
    const int lg = warp & 3;
    const int tt = threadIdx.x - 32 * (2 + STRIP_EPI_WARPS);  // 0 .. 127
    int hj = -1, hnt = 0, ht = 0, hc0 = 0;                    // the previous tile (the middle of the col2im)
    const uint32_t ps_u32 = smem_u32(ps_base);
    uint32_t g = 0;
    for (int s = blockIdx.x; s < p.total_strips; s += gridDim.x) {
      const StripWalk k = strip_walk<LAST>(p, s);
      for (int j = 0; j < k.nt; ++j, ++g) {
        const uint32_t b = g & 1, use = g >> 1;
        strip_wait(&pfull_bar[b], use & 1);
        tc_fence_after();
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(b * ACC_STRIDE + PACC_OFF), v);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&pempty_bar[b]);
        // st.shared / ld.shared by 32-bit address in this role too (generic-pointer LD / ST carry 64-bit address math and
        // take the generic path through the LSU)
        const uint32_t dst = ps_u32 + (uint32_t)(((g % STRIP_P_SLOTS) * (128 * STRIP_PS) + (lg * 32 + lane) * STRIP_PS) * 4);
        strip_store27<0>(dst, v);
        named_barrier_sync(2, 32 * STRIP_TAIL_WARPS);         // P(g) complete in shared memory
        if (hj >= 1 && hj <= hnt - 2) {                       // top and bottom tiles of a strip are halo only
          const int y = ht % (p.h + 1), item = ht / (p.h + 1);
          const int xs = hc0 == 0 ? 0 : hc0 + 1;
          const int xe = min(hc0 + STRIP_POS - 1, p.w);
          if (y < p.h) {
            const float* p0 = ps_base + ((g + 1) % STRIP_P_SLOTS) * (128 * STRIP_PS);   // P(g-2): (g-2) % 3 == (g+1) % 3
            const float* p1 = ps_base + ((g + 2) % STRIP_P_SLOTS) * (128 * STRIP_PS);   // P(g-1)
            const float* p2 = ps_base + (g % STRIP_P_SLOTS) * (128 * STRIP_PS);         // P(g)
            float* orow = p.out + (((size_t)item * p.h + y) * p.w) * p.cout3;
            if (p.cout3 == 3) {
              // the usual case with every offset a compile-time constant (shared-memory reads with immediate offsets)
              for (int x = xs + tt; x < xe; x += 32 * STRIP_TAIL_WARPS) {
                const int i = x - hc0;
                float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
                for (int dy = 0; dy < 3; ++dy) {
                  // row i-1 of P(g-2+dy), columns dy*9..: every further offset is an immediate of the load
                  const uint32_t ps = ps_u32 + (uint32_t)((((g + 1 + dy) % STRIP_P_SLOTS) * (128 * STRIP_PS) + (i - 1) * STRIP_PS + dy * 9) * 4);
                  if (i > 0) strip_tap3<0>(ps, a0, a1, a2);
                  strip_tap3<1>(ps, a0, a1, a2);
                  strip_tap3<2>(ps, a0, a1, a2);
                }
                if (p.relu_last) { a0 = fmaxf(a0, 0.f); a1 = fmaxf(a1, 0.f); a2 = fmaxf(a2, 0.f); }
                float* o3 = orow + (size_t)x * 3;
                o3[0] = a0; o3[1] = a1; o3[2] = a2;
              }
            } else
            for (int x = xs + tt; x < xe; x += 32 * STRIP_TAIL_WARPS) {
              const int i = x - hc0;                          // position inside the tile; i-1 = -1 only at x = 0 (zero pad)
              float acc[3] = {0.f, 0.f, 0.f};
#pragma unroll
              for (int dy = 0; dy < 3; ++dy) {
                const float* ps = dy == 0 ? p0 : (dy == 1 ? p1 : p2);
#pragma unroll
                for (int dx = 0; dx < 3; ++dx) {
                  const int ii = i + dx - 1;
                  if (ii >= 0) {
#pragma unroll
                    for (int c = 0; c < 3; ++c)
                      if (c < p.cout3) acc[c] += ps[ii * STRIP_PS + (dy * 3 + dx) * p.cout3 + c];
                  }
                }
              }
#pragma unroll
              for (int c = 0; c < 3; ++c)
                if (c < p.cout3) orow[(size_t)x * p.cout3 + c] = p.relu_last ? fmaxf(acc[c], 0.f) : acc[c];
            }
          }
        }
        named_barrier_sync(2, 32 * STRIP_TAIL_WARPS);         // the slot of P(g-2) may be overwritten by P(g+1)
        hj = j; hnt = k.nt; ht = k.tfirst + j; hc0 = k.c0;
      }
    }
  } else if (UNPOOL) {
    // ===== unpool producers: one input row of the strip = positions x = c0-1 .. c0+120 of tall row tin, as 128 A rows
    // (four 32-row segments 30 positions apart).  A position holds the pooled pixel (y/2, x/2) where the switch names
    // this corner of the 2x2 quad and zero elsewhere (DPooling.down).  A thread owns one 8-channel chunk of four pooled
    // pixels; the pooled row serves two input rows (y even / odd) and the next one is prefetched into registers while
    // the current one is expanded.  The expansion is a table lookup: four switch bytes (two bits each) index a
    // 256-entry table per corner that holds the four 16-bit lane masks (measured: with the masks computed by byte
    // compares the four producer warps needed 570 instructions per row and paced the whole kernel) =====
    const int pt = threadIdx.x - 32 * (2 + STRIP_EPI_WARPS + (LAST ? STRIP_TAIL_WARPS : 0));  // 0 .. 127
    const int chunk = pt & 7, pp = pt >> 3;
    const int pwp = p.pw + 1, prp = (p.ph + 1) * pwp;
    for (int e = pt; e < 4 * 256; e += 32 * STRIP_PROD_WARPS) {
      const uint32_t qc = (uint32_t)e >> 8, idx = (uint32_t)e & 255u;
      uint32_t m[4];
#pragma unroll
      for (int b4 = 0; b4 < 4; ++b4) m[b4] = (((idx >> (2 * b4)) & 3u) == qc) ? 0xFFFFu : 0u;
      lut[e] = make_uint2(m[0] | (m[1] << 16), m[2] | (m[3] << 16));
    }
    named_barrier_sync(3, 32 * STRIP_PROD_WARPS);
    const uint32_t lut_u32 = smem_u32(lut), in_ring_u32 = smem_u32(in_ring);
    griddep_wait();
    uint4 pv[STRIP_NPX], npv[STRIP_NPX];
    uint32_t si[STRIP_NPX], nsi[STRIP_NPX];                                   // per pooled pixel: table index of channels 0-3 | 4-7 << 8
    int curid = -1, nid = -1;                                 // pooled row (item * ph + py) the registers hold / will hold
    auto load_row = [&](int item, int py, int px0, uint4 (&dv)[STRIP_NPX], uint32_t (&ds)[STRIP_NPX]) {
      const __nv_bfloat16* prow = p.pooled + ((size_t)item * prp + (size_t)py * pwp) * 64 + chunk * 8;
      const uint8_t* srow = p.sw + ((size_t)(item % p.Bsw) * p.ph + py) * p.pw * p.C + chunk * 8;
#pragma unroll
      for (int q = 0; q < STRIP_NPX; ++q) {
        const int px = px0 + pp + STRIP_PXSTEP * q;
        if (px >= 0 && px < p.pw) {
          dv[q] = __ldg(reinterpret_cast<const uint4*>(prow + (size_t)px * 64));
          const uint2 s2 = __ldg(reinterpret_cast<const uint2*>(srow + (size_t)px * p.C));
          // bytes b0..b3 in 0..3 -> b0 | b1 << 2 | b2 << 4 | b3 << 6 (the products of the other byte pairs fall below bit 24 or overflow)
          ds[q] = ((s2.x * 0x01041040u) >> 24) | (((s2.y * 0x01041040u) >> 24) << 8);
        } else {
          dv[q] = make_uint4(0u, 0u, 0u, 0u);
          ds[q] = 0u;
        }
      }
    };
    uint32_t gi = 0;
    for (int s = blockIdx.x; s < p.total_strips; s += gridDim.x) {
      const StripWalk k = strip_walk<LAST>(p, s);
      const int px0 = (k.c0 - 1) >> 1;
      // byte offsets inside the A tile of this thread's eight positions (pooled pixel q, corner column xb): the row of
      // its segment, and for the two positions that two segments share also the row in the segment before; 0xFFFF = none
      uint32_t offs[2 * STRIP_NPX];
#pragma unroll
      for (int e = 0; e < 2 * STRIP_NPX; ++e) {
        const int xr = 2 * (px0 + pp + STRIP_PXSTEP * (e >> 1)) + (e & 1) - (k.c0 - 1);      // 0 .. 121 are positions of the tile
        uint32_t o1 = 0xFFFFu, o2 = 0xFFFFu;
        if (xr >= 0 && xr <= 121) {
          const int seg = min(3, xr / STRIP_SEG), l = xr - STRIP_SEG * seg;
          int rr = 32 * seg + l;
          o1 = (uint32_t)(rr * 128 + ((chunk ^ (rr & 7)) << 4));
          if (seg > 0 && l < 2) {
            rr = 32 * (seg - 1) + l + STRIP_SEG;
            o2 = (uint32_t)(rr * 128 + ((chunk ^ (rr & 7)) << 4));
          }
        }
        offs[e] = o1 | (o2 << 16);
      }
      curid = nid = -1;                                       // the cached pixels belong to the previous strip's column block
      // (item, y) of the input row, kept incrementally: no division per row.  Rows above the tensor count as pad rows.
      int tin = k.tfirst - 1;
      int item = tin >= 0 ? tin / (p.h + 1) : 0, y = tin >= 0 ? tin % (p.h + 1) : p.h;
      for (int i = 0; i < k.nt + 2; ++i, ++gi, ++tin) {
        const uint32_t slot = gi % STRIP_IN_SLOTS, use = gi / STRIP_IN_SLOTS;
        const bool live = tin >= 0 && tin < p.T && y < p.h;   // otherwise a zero row (pad row, or outside the tensor)
        const int key = live ? 0 : -1;
        if (live) {
          const int id = item * p.ph + (y >> 1);
          if (id != curid) {
            if (id == nid) {
#pragma unroll
              for (int q = 0; q < STRIP_NPX; ++q) { pv[q] = npv[q]; si[q] = nsi[q]; }
            } else {
              load_row(item, y >> 1, px0, pv, si);
            }
            curid = id;
          }
        }
        // the next pooled row of this item, wanted two input rows from now at the latest; issued AFTER this row's
        // proxy fence (below): the fence is a MEMBAR that would wait for loads in flight
        const int py2 = (y >> 1) + 1;
        const bool want_next = live && py2 < p.ph && curid + 1 != nid && tin + 2 - (y & 1) < p.T && i + 2 - (y & 1) < k.nt + 2;
        const int item_now = item;
        const uint32_t yodd = (uint32_t)(y & 1);
        if (tin < 0) y = (tin == -1) ? 0 : p.h;              // state of the NEXT row; rows above the tensor are pad rows
        else if (++y > p.h) { y = 0; ++item; }
        strip_wait(&in_empty[slot], (use & 1) ^ 1);
        const uint32_t dst = in_ring_u32 + slot * A_TILE_BYTES;
        if (key >= 0) {
          const uint32_t lq = lut_u32 + yodd * 4096u;                                  // corners (y & 1) * 2 + {0, 1}
#pragma unroll
          for (int e = 0; e < 2 * STRIP_NPX; ++e) {
            const uint32_t o1 = offs[e] & 0xFFFFu, o2 = offs[e] >> 16;
            if (o1 != 0xFFFFu) {
              const uint32_t le = lq + (e & 1) * 2048u;
              const uint2 ma = lds64(le + (si[e >> 1] & 255u) * 8u), mb = lds64(le + (si[e >> 1] >> 8) * 8u);
              const uint4 val = make_uint4(pv[e >> 1].x & ma.x, pv[e >> 1].y & ma.y, pv[e >> 1].z & mb.x, pv[e >> 1].w & mb.y);
              sts128(dst + o1, val);
              if (o2 != 0xFFFFu) sts128(dst + o2, val);
            }
          }
        } else {
          const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
          for (int e = 0; e < 2 * STRIP_NPX; ++e) {
            const uint32_t o1 = offs[e] & 0xFFFFu, o2 = offs[e] >> 16;
            if (o1 != 0xFFFFu) sts128(dst + o1, zero);
            if (o2 != 0xFFFFu) sts128(dst + o2, zero);
          }
        }
        fence_proxy_async_smem();                             // the MMA reads the tile through the async proxy
        __syncwarp();
        if (lane == 0) mbar_arrive(&in_full[slot]);
        if (want_next) {
          load_row(item_now, py2, px0, npv, nsi);
          nid = curid + 1;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace tc
}  // namespace dcv
