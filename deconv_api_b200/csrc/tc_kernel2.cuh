// Two-CTA (cta_group::2) variant of the tap-mode kernel for wide layers (N tile 256).
//
// With operands read from shared memory a single CTA tops out at ~72 % of the tensor peak for M=128, N=256
// (profiles/r1_summary.md): every 128x256x16 MMA has to pull 4 KB of A and 8 KB of B through one SM's shared
// memory port.  A CTA pair on one TPC issues ONE 256x256x16 MMA instead: each CTA supplies its own 128 rows of A
// and HALF of the B tile (128 of the 256 weight rows), so the shared-memory traffic per SM and per flop drops by a
// third, and each stage is 32 KB instead of 48 KB (6 stages instead of 4).
//
// Protocol (cluster of 2, rank 0 = leader): both CTAs run a TMA producer; their loads credit the LEADER's full
// barrier (cta_group::2 TMA with the peer bit of the barrier address cleared; arrival count 2 = one arrive per CTA).
// Only the leader's MMA thread issues tcgen05.mma.cta_group::2; its commits are multicast to both CTAs' empty /
// accumulator-full barriers.  Both CTAs run the epilogue on their own 128 TMEM lanes and arrive remotely on the
// leader's accumulator-empty barrier.
#pragma once
#include "tc_kernel.cuh"

namespace dcv {
namespace tc {

// per-CTA bytes of one pipeline stage / of the resident weights: as in tc_kernel.cuh, but each CTA of the pair
// holds only HALF of the B tile (block_n / 2 weight rows)
__host__ __device__ inline int tc2_stage_bytes(int mode, int block_n) {
  if (mode == MODE_TAP) return A_TILE_BYTES + (block_n / 2) * 128;
  if (mode == MODE_BAND) return A_BAND_BYTES + 3 * (block_n / 2) * 128;
  return A_BAND_BYTES;
}
__host__ __device__ inline int tc2_resident_bytes(int mode, int block_n, int k_blocks) {
  return mode == MODE_BAND_RES ? 9 * k_blocks * (block_n / 2) * 128 : 0;
}

__device__ __forceinline__ void tmem_alloc2(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish2() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t addr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols) : "memory");
}
// TMA load whose completion bytes go to the barrier at the same offset in the LEADER CTA (peer bit 24 cleared)
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// commit: arrive on the barrier at this offset in BOTH CTAs of the pair once the MMAs issued so far have retired
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}

template <int BLOCK_N, int MODE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
deconv_tc2_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, TcParams p) {
  constexpr int HALF_N = BLOCK_N / 2;
  constexpr int B_HALF_BYTES = HALF_N * 128;
  constexpr int TMEM_COLS = (2 * BLOCK_N) < 32 ? 32 : (2 * BLOCK_N);
  constexpr int MAX_STAGES = 8;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = tc2_stage_bytes(MODE, BLOCK_N);
  const int res_bytes = tc2_resident_bytes(MODE, BLOCK_N, p.k_blocks);
  uint8_t* wres = smem;
  uint8_t* ring = smem + res_bytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ring + p.stages * stage_bytes);
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tfull_bar = empty_bar + MAX_STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* wres_bar = tempty_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wres_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 2); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], 2 * TC2_EPI_WARPS); }
    mbar_init(wres_bar, 2);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc2(tmem_slot, TMEM_COLS);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_tiles = p.m_tiles * p.n_tiles;          // m_tiles counts 256-row tiles here
  const int npairs = gridDim.x >> 1, pair = blockIdx.x >> 1;
  const int wp = p.w + 1;
  const int groups = (MODE == MODE_TAP) ? 9 : 3;
  const int total_st = groups * p.k_blocks;

  if (warp == 0) {
    // ===== TMA producer (both CTAs; completion bytes go to the leader's barriers) =====
    if (lane == 0) {
      if constexpr (MODE == MODE_BAND_RES) {
        // this CTA's half of every weight tile, once; n_tiles == 1 in this mode
        if (leader) mbar_arrive_expect_tx(wres_bar, 2u * (uint32_t)res_bytes);
        for (int t = 0; t < 9; ++t)
          for (int kc = 0; kc < p.k_blocks; ++kc)
            tma_load_2d_pair(wres + (t * p.k_blocks + kc) * B_HALF_BYTES, &tmap_b, wres_bar, kc * BLOCK_K,
                             t * p.n_pad + (int)rank * HALF_N);
        if (!leader) mbar_arrive_remote(wres_bar, 0);
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < total_tiles; tile += npairs) {
        const int m0 = (tile / p.n_tiles) * 256 + (int)rank * 128;
        const int n0 = (tile % p.n_tiles) * BLOCK_N + (int)rank * HALF_N;
        for (int st = 0; st < total_st; ++st) {
          const int g = st / p.k_blocks, kc = st % p.k_blocks;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = ring + stage * stage_bytes;
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2u * (uint32_t)stage_bytes);
          if constexpr (MODE == MODE_TAP) {
            const int shift = (g / 3 - 1) * wp + (g % 3 - 1);
            tma_load_2d_pair(a_dst, &tmap_a, &full_bar[stage], kc * BLOCK_K, m0 + shift);
            tma_load_2d_pair(a_dst + A_TILE_BYTES, &tmap_b, &full_bar[stage], kc * BLOCK_K, g * p.n_pad + n0);
          } else {
            tma_load_2d_pair(a_dst, &tmap_a, &full_bar[stage], kc * BLOCK_K, m0 + (g - 1) * wp - 1);
            if constexpr (MODE == MODE_BAND) {
#pragma unroll
              for (int dx = 0; dx < 3; ++dx)
                tma_load_2d_pair(a_dst + A_BAND_BYTES + dx * B_HALF_BYTES, &tmap_b, &full_bar[stage], kc * BLOCK_K,
                                 (g * 3 + dx) * p.n_pad + n0);
            }
          }
          if (!leader) mbar_arrive_remote(&full_bar[stage], 0);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one thread of the LEADER issues 256 x BLOCK_N x 16 MMAs for the pair =====
    if (lane == 0 && leader) {
      constexpr uint32_t idesc = make_idesc_bf16(256, BLOCK_N);
      if constexpr (MODE == MODE_BAND_RES) {
        mbar_wait(wres_bar, 0);
        tc_fence_after();
      }
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int tile = pair; tile < total_tiles; tile += npairs, ++local) {
        const int buf = local & 1;
        const uint32_t use = (uint32_t)(local >> 1);
        mbar_wait(&tempty_bar[buf], (use & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(buf * BLOCK_N);
        for (int st = 0; st < total_st; ++st) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(ring + stage * stage_bytes);
          if constexpr (MODE == MODE_TAP) {
            const uint64_t a_desc = make_sw128_kmajor_desc(a_addr, 0);
            const uint64_t b_desc = make_sw128_kmajor_desc(a_addr + A_TILE_BYTES, 0);
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k)
              umma_bf16_pair(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc, (st | k) != 0);
          } else {
            const int g = st / p.k_blocks, kc = st % p.k_blocks;
#pragma unroll
            for (int dx = 0; dx < 3; ++dx) {
              const uint64_t a_desc = make_sw128_kmajor_desc(a_addr + dx * 128, 0);
              uint32_t b_addr;
              if constexpr (MODE == MODE_BAND) b_addr = a_addr + A_BAND_BYTES + dx * B_HALF_BYTES;
              else b_addr = smem_u32(wres + ((g * 3 + dx) * p.k_blocks + kc) * B_HALF_BYTES);
              const uint64_t b_desc = make_sw128_kmajor_desc(b_addr, 0);
#pragma unroll
              for (int k = 0; k < BLOCK_K / 16; ++k)
                umma_bf16_pair(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc, (st | dx | k) != 0);
            }
          }
          umma_commit_pair(&empty_bar[stage]);
          if (st == total_st - 1) umma_commit_pair(&tfull_bar[buf]);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
      // drain: the peer's epilogue arrives remotely on these barriers; do not leave before it has
      for (int b = 0; b < 2; ++b) {
        const int uses = (local + 1 - b) >> 1;              // tiles that used accumulator buffer b
        if (uses > 0) mbar_wait(&tempty_bar[b], (uint32_t)((uses - 1) & 1));
      }
    }
  } else {
    // ===== epilogue (both CTAs, own 128 TMEM lanes = own 128 rows of the 256-row tile) =====
    const int lg = warp & 3;
    const int sub = (warp - 2) >> 2;                   // which of the quarter's warps: it takes every NSUB-th column chunk
    constexpr int NSUB = TC2_EPI_WARPS / 4;
    const int rp = (p.h + 1) * wp;
    int local = 0;
    for (int tile = pair; tile < total_tiles; tile += npairs, ++local) {
      const int m0 = (tile / p.n_tiles) * 256 + (int)rank * 128;
      const int n0 = (tile % p.n_tiles) * BLOCK_N;
      const int buf = local & 1;
      const uint32_t use = (uint32_t)(local >> 1);
      mbar_wait(&tfull_bar[buf], use & 1);
      tc_fence_after();
      const int r = m0 + lg * 32 + lane;
      const int q = r % rp;
      const int y = q / wp, x = q % wp;
      const bool in_range = r < p.rows;
      const bool is_pixel = in_range && (x < p.w) && (y < p.h);
      const uint32_t t_row = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(buf * BLOCK_N);
#pragma unroll 1
      for (int c0 = sub * 32; c0 < BLOCK_N; c0 += NSUB * 32) {
        uint32_t v[32];
        tmem_ld32(t_row + (uint32_t)c0, v);
        tmem_ld_wait();
        if (p.out_bf16) {
          if (in_range) {
            uint32_t packed[16];
            uint32_t actw[16];
            if (p.prep_act) {
              const uint4* ap = reinterpret_cast<const uint4*>(p.prep_act + (size_t)r * p.n_pad + n0 + c0);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 t = __ldg(ap + i);
                actw[4 * i] = t.x; actw[4 * i + 1] = t.y; actw[4 * i + 2] = t.z; actw[4 * i + 3] = t.w;
              }
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              float a = __uint_as_float(v[2 * i]), b = __uint_as_float(v[2 * i + 1]);
              if (p.bias) {
                const int c = n0 + c0 + 2 * i;
                if (c < p.n_real) a += __ldg(p.bias + c);
                if (c + 1 < p.n_real) b += __ldg(p.bias + c + 1);
              }
              if (p.relu) { a = fmaxf(a, 0.f); b = fmaxf(b, 0.f); }
              if (p.prep_act) prep_pair(p, actw[i], a, b);
              if (!is_pixel) { a = 0.f; b = 0.f; }
              __nv_bfloat162 h2 = __floats2bfloat162_rn(a, b);
              packed[i] = *reinterpret_cast<uint32_t*>(&h2);
            }
            store_bf16_chunk<32>(p, packed, r, is_pixel, r / rp, y, x, n0 + c0);
          }
        } else if (is_pixel) {
          const int item = r / rp;
          float* dst = p.out_f32 + (((size_t)item * p.h + y) * p.w + x) * p.n_real;
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int c = n0 + c0 + i;
            if (c < p.n_real) {
              float a = __uint_as_float(v[i]);
              if (p.bias) a += __ldg(p.bias + c);
              if (p.relu) a = fmaxf(a, 0.f);
              dst[c] = a;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(&tempty_bar[buf], 0);
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc2(tmem_base, TMEM_COLS);
}

}  // namespace tc
}  // namespace dcv
