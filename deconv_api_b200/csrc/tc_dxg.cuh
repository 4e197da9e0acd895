// Backward conv for layers with 64 output channels (block1_conv2, block2_conv1): "dx-grouped" implicit GEMM.
//
// With N = 64 the kernels of tc_kernel.cuh are bound by fetching the UMMA operands from shared memory (~64 B/clk):
// each of the 9 taps re-reads its own 128x64 A tile for only 64 output columns (32 % tensor-pipe utilisation,
// profiles/r1_summary.md).  Here the three dx taps of one dy share ONE un-shifted A tile by being stacked along N:
//   D[i][dx*64 + c] = sum_dy sum_k X[row(i) + (dy-1)*(w+1)][k] * wd[dy*3+dx][k][c]            (N = 192, 3 x fewer A reads)
// and the dx shift is applied to the RESULT instead of the operand:
//   out[row(j) + 1][c] = D[j][c] + D[j+1][64 + c] + D[j+2][128 + c]
// i.e. one and two TMEM lanes further down = warp shuffles.  To keep every shuffle inside one warp (a warp can only
// read its own 32 TMEM lanes) the 128-row A tile is built from FOUR 32-row TMA boxes that start 30 rows apart: warp k
// owns input rows m0-1+30k .. m0+30k+30 and produces the 30 output rows m0+30k .. m0+30k+29.  No cross-warp
// exchange of accumulator data; a tile yields 120 output rows (94 % of the MMA rows).  The packed weights
// [tap][64][K_pad] already have the three dx blocks of a dy contiguous (192 rows): B is one TMA box per (dy, k block).
// With only 12 MMAs per tile the epilogue sets the pace: 16 epilogue warps (lane quarter x 16-column group), the TMEM
// buffer is released as soon as the accumulator is in registers, optional steps are compiled out (EPI), and the
// 120 x 64 bf16 result goes through a swizzled shared-memory tile and ONE TMA store per tile.
#pragma once
#include "tc_kernel.cuh"

namespace dcv {
namespace tc {

constexpr int DXG_N = 192;
constexpr int DXG_ROWS_OUT = 120;             // 4 warps x 30 rows
constexpr int DXG_SEG = 30;
constexpr int DXG_EPI_WARPS = 16;              // 4 lane quarters x 4 column chunks of 16: one warp per SMSP would be latency-bound
constexpr int DXG_THREADS = 64 + 32 * DXG_EPI_WARPS;
constexpr int DXG_B_BYTES = DXG_N * 128;              // 24 KB per (dy, k block)
constexpr int DXG_OUT_BYTES = DXG_ROWS_OUT * 128;     // output tile staged in shared memory: 120 rows x 64 bf16

__host__ __device__ inline int dxg_resident_bytes(int resident, int k_blocks) { return resident ? 3 * k_blocks * DXG_B_BYTES : 0; }
__host__ __device__ inline int dxg_stage_bytes(int resident) { return A_TILE_BYTES + (resident ? 0 : DXG_B_BYTES); }

// EPI bits: 1 = add the bias (forward conv), 2 = fold the DeepDream gradient prep of the next layer in.
// GROUPED = 0 turns the kernel into a plain [rows x K] x [K x 64] GEMM (one un-shifted tap, N = 64, four contiguous
// 32-row segments = 128 output rows per tile, no shuffles): the first conv layer over im2col rows, which with 4 MMAs
// per tile is all epilogue - in tc_kernel.cuh's 4-warp epilogue it ran at 3 us per tile (11 cycles per instruction).
template <int RESIDENT, int EPI, int GROUPED = 1>
__global__ void __launch_bounds__(DXG_THREADS, 1)
deconv_dxg_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ CUtensorMap tmap_out, TcParams p) {
  constexpr int TMEM_COLS = 512;                      // 2 accumulators x 192 columns
  constexpr int ACC_STRIDE = 256;
  constexpr int MAX_STAGES = 8;
  constexpr int N_TILE = GROUPED ? DXG_N : 64;
  constexpr int SEG = GROUPED ? DXG_SEG : 32;         // output rows per lane quarter
  constexpr int ROWS_OUT = 4 * SEG;
  constexpr int B_BYTES = N_TILE * 128;
  constexpr int NDY = GROUPED ? 3 : 1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = A_TILE_BYTES + (RESIDENT ? 0 : B_BYTES);
  const int res_bytes = RESIDENT ? NDY * p.k_blocks * B_BYTES : 0;
  uint8_t* wres = smem;
  uint8_t* ring = smem + res_bytes;
  uint8_t* out_tile = ring + p.stages * stage_bytes;          // [ROWS_OUT][128 B], 128-byte swizzle, source of the TMA store
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(out_tile + ROWS_OUT * 128);
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tfull_bar = empty_bar + MAX_STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* wres_bar = tempty_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wres_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    tma_prefetch_desc(&tmap_out);
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], DXG_EPI_WARPS); }
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
  griddep_launch_dependents();                        // programmatic dependent launch, see tc_kernel.cuh

  const int total_tiles = p.m_tiles;                  // tiles of ROWS_OUT output rows; n_tiles == 1
  const int wp = p.w + 1;
  const int total_st = NDY * p.k_blocks;

  if (warp == 0) {
    if (lane == 0) {
      if constexpr (RESIDENT) {
        mbar_arrive_expect_tx(wres_bar, (uint32_t)res_bytes);
        for (int dy = 0; dy < NDY; ++dy)
          for (int kc = 0; kc < p.k_blocks; ++kc)
            tma_load_2d(wres + (dy * p.k_blocks + kc) * B_BYTES, &tmap_b, wres_bar, kc * BLOCK_K, dy * N_TILE);
      }
      griddep_wait();                                 // the activations come from the previous kernel of the stream
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int m0 = tile * ROWS_OUT;
        for (int st = 0; st < total_st; ++st) {
          const int dy = st / p.k_blocks, kc = st % p.k_blocks;
          // grouped: input rows m0-1 .. of image row dy-1; plain GEMM: the tile's own rows
          const int row0 = GROUPED ? m0 - 1 + (dy - 1) * wp : m0;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = ring + stage * stage_bytes;
          mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)stage_bytes);
#pragma unroll
          for (int k = 0; k < 4; ++k)      // four 32-row segments, SEG rows apart (grouped: 2 rows of overlap per warp boundary)
            tma_load_2d(a_dst + k * 32 * 128, &tmap_a, &full_bar[stage], kc * BLOCK_K, row0 + SEG * k);
          if constexpr (!RESIDENT) tma_load_2d(a_dst + A_TILE_BYTES, &tmap_b, &full_bar[stage], kc * BLOCK_K, dy * N_TILE);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc_bf16(BLOCK_M, N_TILE);
    if constexpr (RESIDENT) {
      mbar_wait(wres_bar, 0);
      tc_fence_after();
    }
    int stage = 0;
    uint32_t phase = 0;
    int local = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      const int buf = local & 1;
      const uint32_t use = (uint32_t)(local >> 1);
      mbar_wait(&tempty_bar[buf], (use & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(buf * ACC_STRIDE);
      for (int st = 0; st < total_st; ++st) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(ring + stage * stage_bytes);
        const uint32_t b_addr = RESIDENT ? smem_u32(wres + st * B_BYTES) : a_addr + A_TILE_BYTES;
        if (elect_one()) {
          const uint64_t a_desc = make_sw128_kmajor_desc(a_addr, 0);
          const uint64_t b_desc = make_sw128_kmajor_desc(b_addr, 0);
#pragma unroll
          for (int k = 0; k < BLOCK_K / 16; ++k)
            umma_bf16(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc, (st | k) != 0);
          umma_commit(&empty_bar[stage]);
          if (st == total_st - 1) umma_commit(&tfull_bar[buf]);
        }
        __syncwarp();
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    const int lg = warp & 3;                                  // TMEM lane quarter this warp may read
    const int c0 = ((warp - 2) >> 2) * 16;                    // its 16 output columns
    const int rp = (p.h + 1) * wp;
    const float relu_lo = p.relu ? 0.f : -INFINITY;
    float bias_r[16];
    if constexpr (EPI & 1) {
#pragma unroll
      for (int i = 0; i < 16; ++i) bias_r[i] = (c0 + i < p.n_real) ? __ldg(p.bias + c0 + i) : 0.f;
    }
    int local = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
      const int m0 = tile * ROWS_OUT;
      const int buf = local & 1;
      const uint32_t use = (uint32_t)(local >> 1);
      const int r = m0 + SEG * lg + lane;                     // output row of this thread (grouped: lanes 30, 31 are helpers only)
      const int q = r % rp;
      const int y = q / wp, x = q % wp;
      const bool in_range = (lane < SEG) && (r < p.rows);
      const bool is_pixel = in_range && (x < p.w) && (y < p.h);
      // gradient prep: the activations this tile will be masked / seeded with do not depend on the accumulator - fetch
      // them BEFORE waiting for the MMAs (behind the wait the load latency sat on the epilogue's critical path:
      // block1_conv2's backward ran at 493 TFLOP/s against 862 for the same conv without the prep)
      uint32_t actw[8];
      if constexpr (EPI & 2) {
        if (in_range) {
          const uint4* ap = reinterpret_cast<const uint4*>(p.prep_act + (size_t)r * p.n_pad + c0);
          const uint4 t0 = __ldg(ap), t1 = __ldg(ap + 1);
          actw[0] = t0.x; actw[1] = t0.y; actw[2] = t0.z; actw[3] = t0.w; actw[4] = t1.x; actw[5] = t1.y; actw[6] = t1.z; actw[7] = t1.w;
        }
      }
      mbar_wait(&tfull_bar[buf], use & 1);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(buf * ACC_STRIDE);
      uint32_t v0[32], v1[32], v2[32];                        // 16 used
      tmem_ld16(t_row + (uint32_t)c0, v0);
      if constexpr (GROUPED) {
        tmem_ld16(t_row + (uint32_t)(64 + c0), v1);
        tmem_ld16(t_row + (uint32_t)(128 + c0), v2);
      }
      tmem_ld_wait();
      // the accumulator is in registers: hand the TMEM buffer back before the arithmetic and the stores
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[buf]);
      // Measured: this epilogue, not the tensor pipe, paces the kernel (12 MMAs per tile; skipping the body took
      // block1_conv2 from 2.4 to 1.35 ms), and it is instruction-bound - so the optional steps are compiled out
      // (EPI) and the ReLU / pad masking are branch-free.
      uint32_t packed[8];
#pragma unroll
      for (int c = 0; c < 16; c += 2) {
        float s[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          float t = __uint_as_float(v0[c + e]);
          if constexpr (GROUPED)
            t += __shfl_down_sync(0xffffffffu, __uint_as_float(v1[c + e]), 1) + __shfl_down_sync(0xffffffffu, __uint_as_float(v2[c + e]), 2);
          if constexpr (EPI & 1) t += bias_r[c + e];
          s[e] = fmaxf(t, relu_lo);
        }
        if constexpr (EPI & 2) { if (in_range) prep_pair(p, actw[c / 2], s[0], s[1]); }
        __nv_bfloat162 h2 = __floats2bfloat162_rn(s[0], s[1]);
        packed[c / 2] = *reinterpret_cast<uint32_t*>(&h2);
      }
      if (!is_pixel) {                                        // pad positions store zeros (one predicated block, not a select per pair)
#pragma unroll
        for (int c = 0; c < 8; ++c) packed[c] = 0u;
      }
      // Output through shared memory + ONE TMA store per tile: a direct store has every thread write 32 bytes of its
      // own 128-byte row, 32 lines per warp instruction (measured: 0.6 ms of block1_conv2's 2.4 ms).
      const bool issuer = (warp == 2) && (lane == 0);
      if (issuer) tma_store_wait_read();                      // the previous tile's store has read the staging tile
      named_barrier_sync(1, 32 * DXG_EPI_WARPS);
      if (lane < SEG) {
        // st.shared by 32-bit address (a generic-pointer store costs a dozen instructions of address conversion)
        const int row_local = SEG * lg + lane;
        const int j = c0 >> 3;                                // 16-byte chunk index of this thread's first 8 channels
        const uint32_t rowp = smem_u32(out_tile) + (uint32_t)(row_local * 128);
        sts128(rowp + (uint32_t)(((j) ^ (row_local & 7)) << 4), make_uint4(packed[0], packed[1], packed[2], packed[3]));
        sts128(rowp + (uint32_t)(((j + 1) ^ (row_local & 7)) << 4), make_uint4(packed[4], packed[5], packed[6], packed[7]));
      }
      fence_proxy_async_smem();
      named_barrier_sync(1, 32 * DXG_EPI_WARPS);
      if (issuer) {
        tma_store_2d(&tmap_out, out_tile, 0, m0);             // rows beyond p.rows are clipped by the tensor map
        tma_store_commit();
      }
    }
    if (warp == 2 && lane == 0) tma_store_wait_read();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace tc
}  // namespace dcv
