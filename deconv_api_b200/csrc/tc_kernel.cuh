// The tcgen05 implicit-GEMM kernel of the backward conv (included by conv_tc.cu only).
//
// Three operand-staging modes share one pipeline and one epilogue:
//   MODE_TAP       one smem stage = (tap, 64-channel block): A box of 128 rows at row m0 + dy*(w+1) + dx,
//                  B box of BLOCK_N weight rows.  Best when N is large (N tile 256: block3/block4 layers).
//   MODE_BAND      one smem stage = (dy, 64-channel block): ONE A box of 136 rows starting at
//                  m0 + dy*(w+1) - 1 serves the three dx taps (UMMA descriptors start 0, 1, 2 rows into
//                  the box), plus the three B boxes.  Cuts the A traffic from L2 by 2.8x; for N <= 128.
//   MODE_BAND_RES  as MODE_BAND, but ALL weights of the layer (9 x K x N) are loaded into shared memory
//                  once per persistent CTA and stay resident; stages then hold only A bands.  For the
//                  64-channel block1 layers, which are otherwise bound by the L2 -> SM operand stream.
#pragma once
#include "common.cuh"
#include "tc_ptx.cuh"

namespace dcv {
namespace tc {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;                 // bf16 elements = 128 bytes = one swizzle row
constexpr int A_TILE_BYTES = BLOCK_M * 128;
constexpr int BAND_ROWS = 136;              // 128 + 2 halo rows, rounded up to the 8-row swizzle atom
constexpr int A_BAND_BYTES = BAND_ROWS * 128;
#ifndef TC2_EPI_WARPS_CFG
#define TC2_EPI_WARPS_CFG 8
#endif
constexpr int TC2_EPI_WARPS = TC2_EPI_WARPS_CFG;   // CTA-pair kernel (tc_kernel2.cuh): epilogue warps per CTA, 4 or 8 (two per TMEM lane quarter, interleaved column chunks)
constexpr int NUM_THREADS = 64 + 32 * TC2_EPI_WARPS;   // + TMA warp, MMA warp
constexpr int TC1_EPI_WARPS = 8;            // deconv_tc_kernel: two epilogue warps per TMEM lane quarter, interleaved column chunks (16 measured equal)
constexpr int TC1_THREADS = 64 + 32 * TC1_EPI_WARPS;
constexpr int SMEM_BUDGET = 200 * 1024;

enum { MODE_TAP = 0, MODE_BAND = 1, MODE_BAND_RES = 2 };

struct TcParams {
  int rows;          // P * (h+1) * (w+1)
  int h, w;          // image size (unpadded)
  int k_blocks;      // K_pad / 64
  int n_pad;         // padded output channels (weights / bf16 output pitch)
  int n_real;        // real output channels (fp32 compact output)
  int m_tiles, n_tiles;
  int relu;
  int stages;        // smem ring depth (host-computed from the budget)
  int base_off;      // 1: put (start_addr >> 7) & 7 into the descriptor's base-offset field
  int ntaps;         // 9 = 3x3 conv; 1 = plain GEMM over un-shifted rows (first layer through im2col), MODE_TAP only
  int nprod_planes;  // weight planes per tap in the packed weights: 1, or 3 for the split-operand kernels (tc_exact.cuh)
  int nprod;         // 1 = plain bf16 product (the kernels of this file); 6 = fp32-accurate mode: operands split into
                     // three bf16 planes (hi, mid, lo), six cross products - conv_exact_kernel only
  // fp16 two-plane scheme of the exact kernels (tc_exact.cuh, SCHEME 1): bit patterns of the largest |activation| and
  // |weight| (device scalars); both operands were scaled by pow2_scale_from_amax() before they were split
  const uint32_t* amax_a;
  const uint32_t* amax_b;
  uint32_t* amax_out;        // optional: per-item maximum of |output| (float bits, atomicMax; zeroed by the host): the scale of the NEXT conv
  // conv_exact_kernel SCHEME 1, optional: instead of the fp32 result write it as the NEXT conv's operand - two fp16 planes
  // [hi | lo] per padded-flat row, n_pad columns each, scaled by pow2_scale_from_amax(split_scale[item]); split_scale is
  // an upper bound of the item's largest |output| that is known before this conv runs (largest |input| x the layer's
  // largest L1 norm of an output channel's weights), so no pass over the finished output is needed to find a scale
  uint16_t* split_out;
  const uint32_t* split_scale;
  const float* bias;         // n_real entries added before the ReLU (forward conv), or null
  // DeepDream backward: fold the NEXT layer's gradient prep into this epilogue (bf16 plain store only):
  //   out = mask(act > 0) * (acc + prep_scale * act), act = that layer's bf16 tape in the output's layout
  const __nv_bfloat16* prep_act;
  float prep_scale;
  int prep_relu;
  // fused switch-driven unpool of the result (the conv's output feeds a pool layer's .down): out_bf16 is then the
  // padded-flat tensor at the UNPOOLED size (up_h, up_w) and every output row is expanded into its 2x2 quad
  const uint8_t* up_sw;      // switches (up_B, h, w, up_C) bytes, or null = plain store
  int up_B, up_C, up_h, up_w;
  __nv_bfloat16* out_bf16;   // padded-flat (rows, n_pad), or null
  float* out_f32;            // compact (P,h,w,n_real), or null
};

__host__ __device__ inline int tc_stage_bytes(int mode, int block_n) {
  if (mode == MODE_TAP) return A_TILE_BYTES + block_n * 128;
  if (mode == MODE_BAND) return A_BAND_BYTES + 3 * block_n * 128;
  return A_BAND_BYTES;
}
__host__ __device__ inline int tc_resident_bytes(int mode, int block_n, int k_blocks) {
  return mode == MODE_BAND_RES ? 9 * k_blocks * block_n * 128 : 0;
}

// Stores CHUNK bf16 values (packed pairs) of output row r, columns [col, col + CHUNK).  Plain: one row of the
// padded-flat output.  With p.up_sw: the row is a POOLED pixel (item, y, x) whose values go to the position its
// switch names inside the 2x2 quad of the unpooled output, zeros to the other three (DPooling.down fused into the
// producing conv's epilogue: the pooled tensor is never written).  Pad rows of the pooled grid write the zero pad
// row / column of the unpooled tensor.
// gradient prep on a pair of values; actbits = two bf16 activations of the layer the gradient is flowing into
__device__ __forceinline__ void prep_pair(const TcParams& p, uint32_t actbits, float& a, float& b) {
  const float a0 = __uint_as_float(actbits << 16), a1 = __uint_as_float(actbits & 0xFFFF0000u);
  a = fmaf(p.prep_scale, a0, a);
  b = fmaf(p.prep_scale, a1, b);
  if (p.prep_relu) {
    if (!(a0 > 0.f)) a = 0.f;
    if (!(a1 > 0.f)) b = 0.f;
  }
}

template <int CHUNK>
__device__ __forceinline__ void store_bf16_chunk(const TcParams& p, const uint32_t (&packed)[CHUNK / 2], int r, bool is_pixel,
                                                 int item, int y, int x, int col) {
  if (!p.up_sw) {
    uint4* dst = reinterpret_cast<uint4*>(p.out_bf16 + (size_t)r * p.n_pad + col);
#pragma unroll
    for (int i = 0; i < CHUNK / 8; ++i) dst[i] = make_uint4(packed[4 * i], packed[4 * i + 1], packed[4 * i + 2], packed[4 * i + 3]);
    return;
  }
  uint32_t sidx[CHUNK / 4];                      // 4 switch bytes per word
  if (is_pixel) {
    const uint8_t* sp = p.up_sw + (((size_t)(item % p.up_B) * p.h + y) * p.w + x) * p.up_C + col;
    if ((p.up_C & 15) == 0 && col + CHUNK <= p.up_C) {
#pragma unroll
      for (int i = 0; i < CHUNK / 16; ++i) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(sp) + i);
        sidx[4 * i] = q.x; sidx[4 * i + 1] = q.y; sidx[4 * i + 2] = q.z; sidx[4 * i + 3] = q.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < CHUNK / 4; ++i) {
        uint32_t wd = 0xFFFFFFFFu;
#pragma unroll
        for (int b = 0; b < 4; ++b)
          if (col + 4 * i + b < p.up_C) wd = (wd & ~(0xFFu << (8 * b))) | ((uint32_t)__ldg(sp + 4 * i + b) << (8 * b));
        sidx[i] = wd;
      }
    }
  }
  const int owp = p.up_w + 1;
  const size_t obase = (size_t)item * (p.up_h + 1) * owp;
#pragma unroll
  for (int qd = 0; qd < 4; ++qd) {
    const int oy = 2 * y + (qd >> 1), ox = 2 * x + (qd & 1);
    if (oy > p.up_h || ox > p.up_w) continue;
    uint32_t o[CHUNK / 2];
#pragma unroll
    for (int i = 0; i < CHUNK / 2; ++i) {
      uint32_t v = 0u;
      if (is_pixel) {
        const uint32_t s2 = sidx[i >> 1] >> (16 * (i & 1));          // switches of elements 2i, 2i+1
        v = (((s2 & 0xFFu) == (uint32_t)qd) ? (packed[i] & 0xFFFFu) : 0u) |
            ((((s2 >> 8) & 0xFFu) == (uint32_t)qd) ? (packed[i] & 0xFFFF0000u) : 0u);
      }
      o[i] = v;
    }
    uint4* dst = reinterpret_cast<uint4*>(p.out_bf16 + (obase + (size_t)oy * owp + ox) * p.n_pad + col);
#pragma unroll
    for (int i = 0; i < CHUNK / 8; ++i) dst[i] = make_uint4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
  }
}

// SPLIT > 1 (DeepDream's small layers, MODE_BAND): the kernel runs as clusters of SPLIT CTAs that share one output tile
// and each take 1/SPLIT of the 64-channel blocks of K.  CTAs 1.. of the cluster write their fp32 accumulators into the
// shared memory of CTA 0 (st.shared::cluster), which adds them to its own in the epilogue - a 16 x 16 x 512 -> 512 conv
// is 24 tiles of 24 K steps, i.e. 24 busy SMs out of 148 for 13 us; split four ways it is 96 SMs for 6 steps each.
template <int BLOCK_N, int MODE, int SPLIT = 1>
__global__ void __launch_bounds__(TC1_THREADS, 1)
deconv_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, TcParams p) {
  constexpr int B_TILE_BYTES = BLOCK_N * 128;
  // One TMEM accumulator per tile buffer.  (Rotating the K steps over several accumulators was measured to give no
  // speed-up - the ~85 cycles per small-N MMA are the shared-memory operand fetch, not a dependency chain - and only
  // cost epilogue time; the fp32-accurate mode that needed them now accumulates in registers, tc_exact.cuh.)
  constexpr int KSPLIT = 1;
  constexpr int ACC_COLS = BLOCK_N * KSPLIT;              // TMEM columns of one accumulator buffer
  constexpr int TMEM_COLS = (2 * ACC_COLS) < 32 ? 32 : (2 * ACC_COLS);
  constexpr int MAX_STAGES = 8;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = tc_stage_bytes(MODE, BLOCK_N);
  const int res_bytes = tc_resident_bytes(MODE, BLOCK_N, p.k_blocks);
  uint8_t* wres = smem;                               // resident weights (MODE_BAND_RES)
  uint8_t* ring = smem + res_bytes;                   // stage ring
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ring + p.stages * stage_bytes);
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tfull_bar = empty_bar + MAX_STAGES;       // [2] accumulator ready
  uint64_t* tempty_bar = tfull_bar + 2;               // [2] accumulator drained
  uint64_t* wres_bar = tempty_bar + 2;                // resident weights landed
  uint64_t* red_full = wres_bar + 1;                  // SPLIT: the peers' partial sums of a tile have landed (CTA 0)
  uint64_t* red_free = red_full + 1;                  // SPLIT: CTA 0 has read this CTA's partial sums (peers)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(red_free + 1);
  // SPLIT: (SPLIT - 1) partial accumulators of 128 x BLOCK_N floats behind the barriers, as [peer][column / 4][row][4]
  // (one 16-byte vector per row and column quad: rows - the lanes - are consecutive, so the writes and reads are
  // free of bank conflicts)
  float4* red = reinterpret_cast<float4*>(reinterpret_cast<uint8_t*>(full_bar) + 256);
  const uint32_t crank = SPLIT > 1 ? cluster_ctarank() : 0u;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], TC1_EPI_WARPS); }
    mbar_init(wres_bar, 1);
    mbar_init(red_full, (SPLIT - 1) * TC1_EPI_WARPS);
    mbar_init(red_free, TC1_EPI_WARPS);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  if constexpr (SPLIT > 1) cluster_sync_all();          // the peers' barriers exist before anyone arrives on them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Programmatic dependent launch (DeepDream chains of small convs): the next kernel may run ITS prologue (barrier
  // init, TMEM allocation, descriptor prefetch) on an SM as soon as this CTA has left it; it blocks in griddep_wait()
  // below, before its first activation load, until this whole grid has finished.
  griddep_launch_dependents();

  const int total_tiles = p.m_tiles * p.n_tiles;
  const int wp = p.w + 1;
  const int groups = (MODE == MODE_TAP) ? p.ntaps : 3; // stages per 64-channel block
  const int kcl = p.k_blocks / SPLIT;                 // 64-channel blocks this CTA multiplies
  const int kc0 = (int)crank * kcl;
  const int total_st = groups * p.nprod * kcl;        // stages per tile
  const int tile0 = blockIdx.x / SPLIT, tile_step = gridDim.x / SPLIT;
  const int kcols = p.k_blocks * BLOCK_K;             // columns of one operand plane

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      if constexpr (MODE == MODE_BAND_RES) {
        // all 9 x k_blocks weight tiles, once; n_tiles == 1 in this mode
        mbar_arrive_expect_tx(wres_bar, (uint32_t)res_bytes);
        for (int t = 0; t < 9; ++t)
          for (int kc = 0; kc < p.k_blocks; ++kc)
            tma_load_2d(wres + (t * p.k_blocks + kc) * B_TILE_BYTES, &tmap_b, wres_bar, kc * BLOCK_K, t * p.n_pad);
      }
      // One stage's loads.  which: 1 = the weight tiles (they never depend on the previous kernel), 2 = the activation box,
      // 3 = both; the barrier is armed with the whole stage's bytes when the weights go out.
      auto load_stage = [&](int st, int m0, int n0, int stage, int which) {
        const int kc = kc0 + st % kcl;
        const int gp = st / kcl;
        const int g = gp / p.nprod, prod = gp % p.nprod;
        // operand planes of product `prod` (hi*hi, hi*mid, mid*hi, hi*lo, lo*hi, mid*mid); plane 0 when nprod == 1
        const int pa = (0x120100 >> (4 * prod)) & 3;       // A plane: 0,0,1,0,2,1
        const int pb = (0x102010 >> (4 * prod)) & 3;       // B plane: 0,1,0,2,0,1
        const int a_col = pa * kcols + kc * BLOCK_K;
        uint8_t* a_dst = ring + stage * stage_bytes;
        if (which & 1) mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)stage_bytes);
        if constexpr (MODE == MODE_TAP) {
          const int shift = p.ntaps == 1 ? 0 : (g / 3 - 1) * wp + (g % 3 - 1);
          if (which & 2) tma_load_2d(a_dst, &tmap_a, &full_bar[stage], a_col, m0 + shift);
          if (which & 1) tma_load_2d(a_dst + A_TILE_BYTES, &tmap_b, &full_bar[stage], kc * BLOCK_K, (g * p.nprod_planes + pb) * p.n_pad + n0);
        } else {
          if (which & 2) tma_load_2d(a_dst, &tmap_a, &full_bar[stage], a_col, m0 + (g - 1) * wp - 1);
          if constexpr (MODE == MODE_BAND) {
            if (which & 1) {
#pragma unroll
              for (int dx = 0; dx < 3; ++dx)
                tma_load_2d(a_dst + A_BAND_BYTES + dx * B_TILE_BYTES, &tmap_b, &full_bar[stage], kc * BLOCK_K,
                            ((g * 3 + dx) * p.nprod_planes + pb) * p.n_pad + n0);
            }
          }
        }
      };
      // Before the wait for the previous kernel: the weight tiles of the first stages of this CTA's first tile (the ring is
      // empty, no slot to wait for).  Small DeepDream layers are one tile per CTA - the weights' trip from L2 then overlaps
      // the predecessor's tail instead of following it.
      int pre = 0;
      if (tile0 < total_tiles) {
        pre = total_st < p.stages ? total_st : p.stages;
        const int m0 = (tile0 / p.n_tiles) * BLOCK_M, n0 = (tile0 % p.n_tiles) * BLOCK_N;
        for (int st = 0; st < pre; ++st) load_stage(st, m0, n0, st, 1);
      }
      griddep_wait();                 // the activations were written by the previous kernel of the stream
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = tile0; tile < total_tiles; tile += tile_step) {
        const int m0 = (tile / p.n_tiles) * BLOCK_M;
        const int n0 = (tile % p.n_tiles) * BLOCK_N;
        for (int st = 0; st < total_st; ++st) {
          if (tile == tile0 && st < pre) {
            load_stage(st, m0, n0, stage, 2);         // the weights of this stage are already on their way
          } else {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            load_stage(st, m0, n0, stage, 3);
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the whole warp walks the pipeline (uniform control flow), one elected lane issues =====
    {
      constexpr uint32_t idesc = make_idesc_bf16(BLOCK_M, BLOCK_N);
      if constexpr (MODE == MODE_BAND_RES) {
        mbar_wait(wres_bar, 0);
        tc_fence_after();
      }
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int tile = tile0; tile < total_tiles; tile += tile_step, ++local) {
        const int buf = local & 1;
        const uint32_t use = (uint32_t)(local >> 1);          // how many times this buffer was used before
        mbar_wait(&tempty_bar[buf], (use & 1) ^ 1);           // epilogue has drained it
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(buf * ACC_COLS);
        for (int st = 0; st < total_st; ++st) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(ring + stage * stage_bytes);
          const int kc = kc0 + st % kcl;
          const int gp = st / kcl;
          const int g = gp / p.nprod;
          if (elect_one()) {
            if constexpr (MODE == MODE_TAP) {
              const uint64_t a_desc = make_sw128_kmajor_desc(a_addr, 0);
              const uint64_t b_desc = make_sw128_kmajor_desc(a_addr + A_TILE_BYTES, 0);
#pragma unroll
              for (int k = 0; k < BLOCK_K / 16; ++k) {  // +32 bytes along K inside the swizzle atom = +2 in addr>>4
                const bool accumulate = (st != 0) || (k != 0);
                umma_bf16(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc, accumulate);
              }
            } else {
#pragma unroll
              for (int dx = 0; dx < 3; ++dx) {
                // tap (g-1, dx-1): the A operand starts dx rows (dx * 128 bytes) into the band
                const uint64_t a_desc = make_sw128_kmajor_desc(a_addr + dx * 128, p.base_off ? (uint32_t)dx : 0u);
                uint32_t b_addr;
                if constexpr (MODE == MODE_BAND) b_addr = a_addr + A_BAND_BYTES + dx * B_TILE_BYTES;
                else b_addr = smem_u32(wres + ((g * 3 + dx) * p.k_blocks + kc) * B_TILE_BYTES);
                const uint64_t b_desc = make_sw128_kmajor_desc(b_addr, 0);
#pragma unroll
                for (int k = 0; k < BLOCK_K / 16; ++k) {
                  const bool accumulate = ((st | dx) != 0) || (k != 0);
                  umma_bf16(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc, accumulate);
                }
              }
            }
            umma_commit(&empty_bar[stage]);                     // smem slot free once these MMAs retire
            if (st == total_st - 1) umma_commit(&tfull_bar[buf]);
          }
          __syncwarp();
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===== epilogue warps 2..9: TMEM lane group = warp % 4; warps 2-5 take the even column chunks, 6-9 the odd ones.
    // One warp per sub-partition runs at ~11 cycles per instruction (latency-bound), which made the epilogue - not the
    // MMAs - the pace of every layer with few K steps per tile (forward convs with bias, K = 64 layers) =====
    const int lg = warp & 3;
    const int sub = (warp - 2) >> 2;
    constexpr int NSUB = TC1_EPI_WARPS / 4;
    const int rp = (p.h + 1) * wp;
    const float relu_lo = p.relu ? 0.f : -INFINITY;
    int local = 0;
    for (int tile = tile0; tile < total_tiles; tile += tile_step, ++local) {
      const int m0 = (tile / p.n_tiles) * BLOCK_M;
      const int n0 = (tile % p.n_tiles) * BLOCK_N;
      const int buf = local & 1;
      const uint32_t use = (uint32_t)(local >> 1);
      // DeepDream backward: the gradient-prep activations of this thread's first column chunk are fetched before the wait
      // for the MMAs (a small layer is one tile per CTA: the load's latency would otherwise sit at the end of the kernel)
      constexpr int PRE_CHUNK = BLOCK_N >= 32 * NSUB ? 32 : 16;
      uint4 pre_act[PRE_CHUNK / 8];
      if (p.prep_act && crank == 0 && m0 + lg * 32 + lane < p.rows) {
        const uint4* ap = reinterpret_cast<const uint4*>(p.prep_act + (size_t)(m0 + lg * 32 + lane) * p.n_pad + n0 + sub * PRE_CHUNK);
#pragma unroll
        for (int i = 0; i < PRE_CHUNK / 8; ++i) pre_act[i] = __ldg(ap + i);
      }
      mbar_wait(&tfull_bar[buf], use & 1);
      tc_fence_after();
      if constexpr (SPLIT > 1) {
        constexpr int CH = BLOCK_N >= 32 * NSUB ? 32 : 16;
        const uint32_t t_acc = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(buf * ACC_COLS);
        if (crank != 0) {
          // a peer: hand the partial sums to CTA 0 and go on to the next tile
          mbar_wait_cluster(red_free, (uint32_t)(local & 1) ^ 1u);      // CTA 0 has read the previous tile's
          const uint32_t dst0 = mapa_u32(smem_u32(red), 0) +
                                (uint32_t)((((int)crank - 1) * (BLOCK_N / 4) * BLOCK_M + lg * 32 + lane) * 16);
#pragma unroll 1
          for (int c0 = sub * CH; c0 < BLOCK_N; c0 += NSUB * CH) {
            uint32_t v[32];
            if constexpr (CH == 32) tmem_ld32(t_acc + (uint32_t)c0, v);
            else tmem_ld16(t_acc + (uint32_t)c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < CH / 4; ++i)
              st_cluster_v4(dst0 + (uint32_t)((c0 / 4 + i) * BLOCK_M * 16), v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
          }
          fence_acq_rel_cluster();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            mbar_arrive_remote_release(red_full, 0);
            mbar_arrive(&tempty_bar[buf]);
          }
          continue;
        }
        mbar_wait_cluster(red_full, (uint32_t)(local & 1));
      }
      const int r = m0 + lg * 32 + lane;                      // this thread's output row
      const int q = r % rp;
      const int y = q / wp, x = q % wp;
      const bool in_range = r < p.rows;
      const bool is_pixel = in_range && (x < p.w) && (y < p.h);
      const uint32_t t_row = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(buf * ACC_COLS);
      constexpr int CHUNK = BLOCK_N >= 32 * NSUB ? 32 : 16;
#pragma unroll 1
      for (int c0 = sub * CHUNK; c0 < BLOCK_N; c0 += NSUB * CHUNK) {
        uint32_t v[32];
        if constexpr (CHUNK == 32) tmem_ld32(t_row + (uint32_t)c0, v);
        else tmem_ld16(t_row + (uint32_t)c0, v);
        tmem_ld_wait();
        if constexpr (SPLIT > 1) {               // add the peers' partial sums, in CTA order
          const uint32_t rp4 = smem_u32(red) + (uint32_t)(((c0 / 4) * BLOCK_M + lg * 32 + lane) * 16);
#pragma unroll
          for (int sp = 0; sp < SPLIT - 1; ++sp) {
#pragma unroll
            for (int i = 0; i < CHUNK / 4; ++i) {
              const float4 u = lds128_f32(rp4 + (uint32_t)((sp * (BLOCK_N / 4) + i) * BLOCK_M * 16));
              v[4 * i] = __float_as_uint(__uint_as_float(v[4 * i]) + u.x);
              v[4 * i + 1] = __float_as_uint(__uint_as_float(v[4 * i + 1]) + u.y);
              v[4 * i + 2] = __float_as_uint(__uint_as_float(v[4 * i + 2]) + u.z);
              v[4 * i + 3] = __float_as_uint(__uint_as_float(v[4 * i + 3]) + u.w);
            }
          }
        }
        if constexpr (KSPLIT > 1) {              // add the other K-split accumulators
#pragma unroll
          for (int sp = 1; sp < KSPLIT; ++sp) {
            uint32_t u[32];
            if constexpr (CHUNK == 32) tmem_ld32(t_row + (uint32_t)(sp * BLOCK_N + c0), u);
            else tmem_ld16(t_row + (uint32_t)(sp * BLOCK_N + c0), u);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < CHUNK; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) + __uint_as_float(u[i]));
          }
        }
        if (p.out_bf16) {
          if (in_range) {
            // optional steps as separate, warp-uniform passes over the chunk (nothing per element when they are off);
            // ReLU and pad masking branch-free.  One epilogue warp per sub-partition is latency-bound, so its
            // instruction count is the epilogue's duration.
            if (p.bias) {
              if ((p.n_real & 3) == 0 && n0 + c0 + CHUNK <= p.n_real) {
                const float4* bp = reinterpret_cast<const float4*>(p.bias + n0 + c0);
#pragma unroll
                for (int i = 0; i < CHUNK / 4; ++i) {
                  const float4 b4 = __ldg(bp + i);
                  v[4 * i] = __float_as_uint(__uint_as_float(v[4 * i]) + b4.x);
                  v[4 * i + 1] = __float_as_uint(__uint_as_float(v[4 * i + 1]) + b4.y);
                  v[4 * i + 2] = __float_as_uint(__uint_as_float(v[4 * i + 2]) + b4.z);
                  v[4 * i + 3] = __float_as_uint(__uint_as_float(v[4 * i + 3]) + b4.w);
                }
              } else {
#pragma unroll
                for (int i = 0; i < CHUNK; ++i) {
                  const int c = n0 + c0 + i;
                  if (c < p.n_real) v[i] = __float_as_uint(__uint_as_float(v[i]) + __ldg(p.bias + c));
                }
              }
            }
#pragma unroll
            for (int i = 0; i < CHUNK; ++i) v[i] = __float_as_uint(fmaxf(__uint_as_float(v[i]), relu_lo));
            if (p.prep_act) {
              const uint4* ap = reinterpret_cast<const uint4*>(p.prep_act + (size_t)r * p.n_pad + n0 + c0);
#pragma unroll
              for (int i = 0; i < CHUNK / 8; ++i) {
                const uint4 t = c0 == sub * CHUNK ? pre_act[i] : __ldg(ap + i);
                const uint32_t aw[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  float a = __uint_as_float(v[8 * i + 2 * k]), b = __uint_as_float(v[8 * i + 2 * k + 1]);
                  prep_pair(p, aw[k], a, b);
                  v[8 * i + 2 * k] = __float_as_uint(a); v[8 * i + 2 * k + 1] = __float_as_uint(b);
                }
              }
            }
            uint32_t packed[CHUNK / 2];
#pragma unroll
            for (int i = 0; i < CHUNK / 2; ++i) {
              __nv_bfloat162 h2 = __floats2bfloat162_rn(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
              packed[i] = is_pixel ? *reinterpret_cast<uint32_t*>(&h2) : 0u;
            }
            store_bf16_chunk<CHUNK>(p, packed, r, is_pixel, r / rp, y, x, n0 + c0);
          }
        } else if (is_pixel) {
          const int item = r / rp;
          float* dst = p.out_f32 + (((size_t)item * p.h + y) * p.w + x) * p.n_real;
          if ((p.n_real & 3) == 0 && n0 + c0 + CHUNK <= p.n_real) {      // 16-byte stores for whole chunks
#pragma unroll
            for (int i = 0; i < CHUNK; i += 4) {
              float a[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                a[k] = __uint_as_float(v[i + k]);
                if (p.bias) a[k] += __ldg(p.bias + n0 + c0 + i + k);
                if (p.relu) a[k] = fmaxf(a[k], 0.f);
              }
              *reinterpret_cast<float4*>(dst + n0 + c0 + i) = make_float4(a[0], a[1], a[2], a[3]);
            }
          } else {
#pragma unroll
            for (int i = 0; i < CHUNK; ++i) {
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
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&tempty_bar[buf]);
        if constexpr (SPLIT > 1) {               // the peers may overwrite their part of `red`
          for (uint32_t c = 1; c < (uint32_t)SPLIT; ++c) mbar_arrive_remote_release(red_free, c);
        }
      }
    }
  }

  tc_fence_before();
  if constexpr (SPLIT > 1) cluster_sync_all();          // no CTA leaves while a peer may still write to it or arrive on it
  else __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace tc
}  // namespace dcv
