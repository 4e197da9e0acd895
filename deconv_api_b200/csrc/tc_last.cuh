// Last backward conv of a projection (64 -> 3 channels in VGG16): "GEMM + col2im" on tcgen05.
//
// In the implicit-GEMM form of tc_kernel.cuh this layer has N = 3 output channels, so each 128x16x16 MMA is
// spent on reading its A slice from shared memory (8 % tensor-pipe utilisation, profiles/r1_summary.md).  Here
// the 9 taps are stacked along N instead:  P[r][tap*cout + c] = sum_k X[r][k] * wd[tap][k][c]  is ONE GEMM with
// N = 9*cout <= 32 and NO shifted operands, and the convolution is finished in the epilogue as
//   out[r'][c] = relu( sum_tap P[r' + shift_tap][tap*cout + c] ).
// Tiles are 2-D patches so that the shifted rows stay inside the tile: 14x14 output pixels need the 16x16 halo
// patch = 256 rows = two M=128 MMAs per 64-channel block; the patch is ONE 3-D TMA box (64 ch, 16 x, 16 y) over
// the padded-flat tensor viewed as (C_pad, w+1, P*(h+1)).  4*k_blocks*2 MMAs per 196 pixels instead of 36 per 128.
//
// EXACT = 1 (fp32 projection mode): the operands are the three-plane bf16 splits of conv_tc.cu (activations: columns
// [hi | mid | lo] of the padded-flat tensor; weights: three 32-row blocks).  One ring stage holds ONE plane's patch; the
// planes arrive lo, mid, hi and the MMA warp issues the products each plane takes part in (lo*hi | mid*mid, mid*hi |
// hi*lo, hi*mid, hi*hi), so the small correction products enter the TMEM accumulator first and only the four hi*hi
// MMAs of a 64-channel block are added while it is large (the tensor core truncates on accumulate).
#pragma once
#include "common.cuh"
#include "tc_ptx.cuh"

namespace dcv {
namespace tc {

constexpr int LAST_T = 14;                       // output patch edge
constexpr int LAST_H = LAST_T + 2;               // halo patch edge (16)
constexpr int LAST_ROWS = LAST_H * LAST_H;       // 256 rows = 2 x M128
constexpr int LAST_N = 32;                       // 9 * cout rounded up
constexpr int LAST_A_BYTES = LAST_ROWS * 128;    // 32 KB per 64-channel block
constexpr int LAST_STAGES = 4;
constexpr int LAST_PS = 29;                      // P row pitch in shared memory (floats), odd -> conflict-free
constexpr int LAST_THREADS = 320;                // TMA warp, MMA warp, 8 epilogue warps
constexpr int LAST_EPI = 256;

struct LastParams {
  int P, h, w, cout;          // items, image size, real output channels (9*cout <= 32)
  int k_blocks;
  int tiles_x, tiles_y, total_tiles;
  int relu;
  float* out;                 // compact (P,h,w,cout) fp32
  // EXACT = 2: fp16 two-plane operands scaled by powers of two (tc_exact.cuh, SCHEME 1): the largest |activation| and
  // |weight| as float bit patterns (device scalars)
  const uint32_t* amax_a;
  const uint32_t* amax_b;
};

__host__ __device__ inline int last_smem_bytes(int k_blocks, int exact = 0) {
  return (exact == 1 ? 3 : (exact == 2 ? 2 : 1)) * k_blocks * LAST_N * 128 + LAST_STAGES * LAST_A_BYTES + 2 * LAST_ROWS * LAST_PS * 4 + 1024 + 256;
}

__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

template <int EXACT>
__global__ void __launch_bounds__(LAST_THREADS, 1)
deconv_last_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, LastParams p) {
  constexpr int NPLANES = EXACT == 1 ? 3 : (EXACT == 2 ? 2 : 1);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int w_bytes = NPLANES * p.k_blocks * LAST_N * 128;   // resident weights [plane][k_block][32 rows][128 B]
  uint8_t* wres = smem;                                       // resident all-tap weights [k_block][32 rows][128 B]
  uint8_t* ring = smem + w_bytes;                             // A patches
  float* ps_base = reinterpret_cast<float*>(ring + LAST_STAGES * LAST_A_BYTES);  // two P tiles [256][LAST_PS]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(ps_base) + 2 * LAST_ROWS * LAST_PS * 4);
  uint64_t* empty_bar = full_bar + LAST_STAGES;
  uint64_t* tfull_bar = empty_bar + LAST_STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* wres_bar = tempty_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wres_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr int TMEM_COLS = 128;                              // 2 buffers x 2 halves x 32 columns

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < LAST_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], 8); }
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
  const int tiles_per_item = p.tiles_x * p.tiles_y;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(wres_bar, (uint32_t)w_bytes);
      for (int pl = 0; pl < NPLANES; ++pl)
        for (int kc = 0; kc < p.k_blocks; ++kc)
          tma_load_2d(wres + (pl * p.k_blocks + kc) * LAST_N * 128, &tmap_b, wres_bar, kc * 64, pl * LAST_N);
      griddep_wait();                                 // the activations come from the previous kernel of the stream
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const int item = tile / tiles_per_item, t2 = tile % tiles_per_item;
        const int y0 = (t2 / p.tiles_x) * LAST_T, x0 = (t2 % p.tiles_x) * LAST_T;
        for (int kc = 0; kc < p.k_blocks; ++kc) {
          for (int pi = 0; pi < NPLANES; ++pi) {
            const int pl = NPLANES - 1 - pi;                 // exact: lo, mid, hi
            mbar_wait(&empty_bar[stage], phase ^ 1);
            mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)LAST_A_BYTES);
            tma_load_3d(ring + stage * LAST_A_BYTES, &tmap_a, &full_bar[stage], (pl * p.k_blocks + kc) * 64, x0 - 1,
                        item * (p.h + 1) + y0 - 1);
            if (++stage == LAST_STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = EXACT == 2 ? make_idesc_f16(128, LAST_N) : make_idesc_bf16(128, LAST_N);
      mbar_wait(wres_bar, 0);
      tc_fence_after();
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++local) {
        const int buf = local & 1;
        const uint32_t use = (uint32_t)(local >> 1);
        mbar_wait(&tempty_bar[buf], (use & 1) ^ 1);
        tc_fence_after();
        for (int kc = 0; kc < p.k_blocks; ++kc) {
          for (int pi = 0; pi < NPLANES; ++pi) {
            const int pa = NPLANES - 1 - pi;                 // activation plane in this stage (exact: lo, mid, hi)
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(ring + stage * LAST_A_BYTES);
            // weight planes this activation plane is multiplied with, smallest product first:
            // lo -> {hi}; mid -> {mid, hi}; hi -> {lo, mid, hi}   (plain mode: the one plane)
            const int nb = EXACT ? NPLANES - pa : 1;
            for (int q = 0; q < nb; ++q) {
              const int pb = EXACT ? nb - 1 - q : 0;
              const uint64_t b_desc = make_sw128_kmajor_desc(smem_u32(wres + (pb * p.k_blocks + kc) * LAST_N * 128), 0);
#pragma unroll
              for (int half = 0; half < 2; ++half) {
                const uint64_t a_desc = make_sw128_kmajor_desc(a_addr + half * 128 * 128, 0);
                const uint32_t d_tmem = tmem_base + (uint32_t)(buf * 64 + half * 32);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  umma_bf16(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc, (kc | pi | q | k) != 0);
              }
            }
            umma_commit(&empty_bar[stage]);
            if (kc == p.k_blocks - 1 && pi == NPLANES - 1) umma_commit(&tfull_bar[buf]);
            if (++stage == LAST_STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else {
    // ===== 8 epilogue warps: warps 2-5 drain accumulator half 0, warps 6-9 half 1 (TMEM lane group = warp % 4);
    // then all 256 threads finish the convolution from the P tile in shared memory, one output pixel each =====
    const int lg = warp & 3;
    const int half = (warp - 2) >> 2;
    const int et = threadIdx.x - 64;                          // 0..255 among the epilogue threads
    const int ncol = 9 * p.cout;
    const int ly = et / LAST_T, lx = et % LAST_T;             // this thread's output pixel inside the patch
    const bool has_px = et < LAST_T * LAST_T;
    const float wscale_inv = EXACT == 2 ? 1.f / pow2_scale_from_amax(__ldg(p.amax_b)) : 1.f;
    int local = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++local) {
      const int item = tile / tiles_per_item, t2 = tile % tiles_per_item;
      const int y0 = (t2 / p.tiles_x) * LAST_T, x0 = (t2 % p.tiles_x) * LAST_T;
      const int buf = local & 1;
      const uint32_t use = (uint32_t)(local >> 1);
      float* ps = ps_base + buf * (LAST_ROWS * LAST_PS);       // P tiles alternate, so one barrier per tile suffices
      mbar_wait(&tfull_bar[buf], use & 1);
      tc_fence_after();
      {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(buf * 64 + half * 32), v);
        tmem_ld_wait();
        // st.shared / ld.shared by 32-bit address (generic LD / ST carry 64-bit address math and take the generic path)
        const uint32_t dst = smem_u32(ps) + (uint32_t)((half * 128 + lg * 32 + lane) * LAST_PS * 4);
#pragma unroll
        for (int j = 0; j < 27; ++j)
          if (j < ncol) sts32(dst + 4 * j, v[j]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[buf]);          // TMEM buffer is free: next tile's MMAs may start
      asm volatile("bar.sync 1, 256;" ::: "memory");          // P tile complete in shared memory
      const int y = y0 + ly, x = x0 + lx;
      if (has_px && y < p.h && x < p.w) {
        float* dst = p.out + (((size_t)item * p.h + y) * p.w + x) * p.cout;
        const float* pp = ps + (ly * LAST_H + lx) * LAST_PS;
        if (p.cout == 3) {
          // the usual case with every offset a compile-time constant; same summation order as the loop below
          const uint32_t pp_u32 = smem_u32(pp);
          float a3[3] = {0.f, 0.f, 0.f};
#pragma unroll
          for (int t = 0; t < 9; ++t) {
#pragma unroll
            for (int c = 0; c < 3; ++c) a3[c] += lds_f32(pp_u32 + (uint32_t)((((t / 3) * LAST_H + (t % 3)) * LAST_PS + t * 3 + c) * 4));
          }
          float sc = 1.f;
          if (EXACT == 2) sc = wscale_inv / pow2_scale_from_amax(__ldg(p.amax_a + item));        // per-item activation scale
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            float acc = a3[c];
            if (EXACT == 2) acc *= sc;
            if (p.relu) acc = fmaxf(acc, 0.f);
            dst[c] = acc;
          }
        } else
        for (int c = 0; c < p.cout; ++c) {
          float acc = 0.f;
#pragma unroll
          for (int t = 0; t < 9; ++t) acc += pp[((t / 3) * LAST_H + (t % 3)) * LAST_PS + t * p.cout + c];
          if (EXACT == 2) acc *= wscale_inv / pow2_scale_from_amax(__ldg(p.amax_a + item));      // per-item activation scale
          if (p.relu) acc = fmaxf(acc, 0.f);
          dst[c] = acc;
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
