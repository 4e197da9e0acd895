// fp32-accurate 3x3 conv on tcgen05 (included by conv_tc.cu only): the kernel behind the forward pass and the
// DCV_FP32 projection mode.
//
// Arithmetic.  Both operands are split into three bf16 planes (x = hi + mid + lo, 8 mantissa bits each) and the six
// significant cross products hi*hi, hi*mid, mid*hi, hi*lo, lo*hi, mid*mid are formed on the tensor cores; every bf16
// product is exact in fp32, the dropped terms are ~2^-24 relative.  What is NOT exact is the tensor core's own
// accumulation: it truncates when it adds into a TMEM accumulator, so a long K chain inside TMEM drifts towards zero
// (measured: 1e-5 relative after 13 chained layers with three rotating accumulators).  This kernel therefore never
// chains more than one pipeline stage (12 MMAs of K = 16) inside TMEM: each stage starts a fresh accumulator slot,
// and the eight accumulate warps - idle during the main loop of the bf16 kernels - drain finished slots into fp32
// REGISTER accumulators with round-to-nearest adds while the next stages run.  Result: error below the FFMA
// kernel's (sequential fp32 sums), at several times its speed.
//
// Pipeline.  warp 0: TMA producer (one A band of 136 rows serving the three dx taps + three weight tiles per
// stage, as MODE_BAND of tc_kernel.cuh); warp 1: MMA issuer; warps 2-9: accumulate + epilogue.  Barriers:
// full/empty per smem stage, slot_full/slot_empty per TMEM slot.
#pragma once
#include "tc_kernel.cuh"

namespace dcv {
namespace tc {

constexpr int EXACT_SLOTS = 4;

__host__ __device__ inline int exact_stage_bytes(int block_n) { return A_BAND_BYTES + 3 * block_n * 128; }

constexpr int EXACT_THREADS = 320;   // TMA warp, MMA warp, 8 accumulate warps (two per TMEM lane quarter, half of the columns each)

// SCHEME 0: three bf16 planes, six products (hh, hm, mh, hl, lh, mm), ~2^-24: the forward tape.
// SCHEME 1: two fp16 planes of the operands scaled into fp16's range by a power of two taken from the tensor's largest
// magnitude (x*s = hi + lo, 11 significand bits each), three products (hh, hl, lh): ~2^-21 per product at HALF the MMA
// work - the backward convs of the DCV_FP32 chain (measured on the 512-map step against the oracle: see DESIGN.md §4).
// Everything else is shared: one fresh TMEM slot per stage, fp32 register accumulation, the result rescaled once.
// SPLIT > 1 (forward pass of single images: 14x14 .. 28x28 layers are 16 - 56 tiles of 144 stages): the kernel runs as
// clusters of SPLIT CTAs that share one tile and each take 1/SPLIT of the 64-channel blocks of K; CTAs 1.. send their
// fp32 register accumulators to CTA 0's shared memory (st.shared::cluster), CTA 0 adds them in rank order and runs the
// epilogue.  The partial sums are plain fp32 adds: the result differs from the unsplit kernel's in the last place only.
template <int BLOCK_N, int SCHEME = 0, int SPLIT = 1>
__global__ void __launch_bounds__(EXACT_THREADS, 1)
conv_exact_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, TcParams p) {
  constexpr int B_TILE_BYTES = BLOCK_N * 128;
  constexpr int TMEM_COLS = EXACT_SLOTS * BLOCK_N < 32 ? 32 : EXACT_SLOTS * BLOCK_N;
  constexpr int MAX_STAGES = 8;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = exact_stage_bytes(BLOCK_N);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ring + p.stages * stage_bytes);
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* sfull_bar = empty_bar + MAX_STAGES;         // [EXACT_SLOTS] slot holds a finished stage
  uint64_t* sempty_bar = sfull_bar + EXACT_SLOTS;       // [EXACT_SLOTS] slot drained by the eight accumulate warps
  uint64_t* red_full = sempty_bar + EXACT_SLOTS;        // SPLIT: the peers' partial sums of a tile have landed (CTA 0)
  uint64_t* red_free = red_full + 1;                    // SPLIT: CTA 0 has read this CTA's partial sums (peers)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(red_free + 1);
  // SPLIT: (SPLIT - 1) partial accumulators of 128 x BLOCK_N floats behind the barriers, as [peer][column / 4][row][4]
  float4* red = reinterpret_cast<float4*>(reinterpret_cast<uint8_t*>(full_bar) + 256);
  const uint32_t crank = SPLIT > 1 ? cluster_ctarank() : 0u;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < EXACT_SLOTS; ++s) { mbar_init(&sfull_bar[s], 1); mbar_init(&sempty_bar[s], 8); }
    mbar_init(red_full, (SPLIT - 1) * 8);
    mbar_init(red_free, 8);
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

  const int total_tiles = p.m_tiles * p.n_tiles;
  const int wp = p.w + 1;
  constexpr int NPROD = SCHEME == 0 ? 6 : 3;
  constexpr int NPLANE = SCHEME == 0 ? 3 : 2;
  const int kcl = p.k_blocks / SPLIT;                 // 64-channel blocks this CTA multiplies
  const int kc0 = (int)crank * kcl;
  const int total_st = 3 * NPROD * kcl;               // (dy, product, 64-channel block) stages per tile
  const int tile0 = blockIdx.x / SPLIT, tile_step = gridDim.x / SPLIT;
  const int kcols = p.k_blocks * BLOCK_K;             // columns of one operand plane

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = tile0; tile < total_tiles; tile += tile_step) {
        const int m0 = (tile / p.n_tiles) * BLOCK_M;
        const int n0 = (tile % p.n_tiles) * BLOCK_N;
        for (int st = 0; st < total_st; ++st) {
          const int kc = kc0 + st % kcl;
          const int gp = st / kcl;
          const int g = gp / NPROD, prod = gp % NPROD;
          const int pa = SCHEME == 0 ? (0x120100 >> (4 * prod)) & 3 : (0x100 >> (4 * prod)) & 3;   // A plane: 0,0,1,0,2,1 | 0,0,1
          const int pb = SCHEME == 0 ? (0x102010 >> (4 * prod)) & 3 : (0x010 >> (4 * prod)) & 3;   // B plane: 0,1,0,2,0,1 | 0,1,0
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = ring + stage * stage_bytes;
          mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)stage_bytes);
          tma_load_2d(a_dst, &tmap_a, &full_bar[stage], pa * kcols + kc * BLOCK_K, m0 + (g - 1) * wp - 1);
#pragma unroll
          for (int dx = 0; dx < 3; ++dx)
            tma_load_2d(a_dst + A_BAND_BYTES + dx * B_TILE_BYTES, &tmap_b, &full_bar[stage], kc * BLOCK_K,
                        ((g * 3 + dx) * NPLANE + pb) * p.n_pad + n0);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    constexpr uint32_t idesc = SCHEME == 0 ? make_idesc_bf16(BLOCK_M, BLOCK_N) : make_idesc_f16(BLOCK_M, BLOCK_N);
    int stage = 0;
    uint32_t phase = 0;
    uint32_t chunk = 0;                                 // running stage count of this CTA: slot = chunk % EXACT_SLOTS
    for (int tile = tile0; tile < total_tiles; tile += tile_step) {
      for (int st = 0; st < total_st; ++st, ++chunk) {
        const uint32_t slot = chunk % EXACT_SLOTS, suse = chunk / EXACT_SLOTS;
        mbar_wait(&sempty_bar[slot], (suse & 1) ^ 1);   // the accumulate warps have read this slot's previous stage
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(ring + stage * stage_bytes);
        const uint32_t d_tmem = tmem_base + slot * BLOCK_N;
        if (elect_one()) {
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) {
            const uint64_t a_desc = make_sw128_kmajor_desc(a_addr + dx * 128, 0u);
            const uint64_t b_desc = make_sw128_kmajor_desc(a_addr + A_BAND_BYTES + dx * B_TILE_BYTES, 0u);
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k)
              umma_bf16(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc, (dx | k) != 0);
          }
          umma_commit(&empty_bar[stage]);
          umma_commit(&sfull_bar[slot]);
        }
        __syncwarp();
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===== accumulate + epilogue warps 2..9: TMEM lane group = warp % 4, one output row per thread =====
    // (one warp per sub-partition runs at ~11 cycles per instruction; with four warps the adds of an N = 128 stage
    // took longer than its MMAs)
    const int lg = warp & 3;
    constexpr int COLS = BLOCK_N >= 32 ? BLOCK_N / 2 : BLOCK_N;      // columns owned by this warp
    const int cbase = BLOCK_N >= 32 ? ((warp - 2) >> 2) * COLS : 0;
    const bool owner = BLOCK_N >= 32 || warp < 6;                    // N = 16: warps 6-9 only keep the barriers balanced
    const int rp = (p.h + 1) * wp;
    // SCHEME 1: both operands carry a power-of-two scale; undo it once, on the fp32 sum (exact)
    const float wscale_inv = SCHEME == 0 ? 1.f : 1.f / pow2_scale_from_amax(__ldg(p.amax_b));
    uint32_t chunk = 0;
    int local = 0;
    for (int tile = tile0; tile < total_tiles; tile += tile_step, ++local) {
      const int m0 = (tile / p.n_tiles) * BLOCK_M;
      const int n0 = (tile % p.n_tiles) * BLOCK_N + cbase;
      float acc[COLS];
#pragma unroll
      for (int i = 0; i < COLS; ++i) acc[i] = 0.f;
      for (int st = 0; st < total_st; ++st, ++chunk) {
        const uint32_t slot = chunk % EXACT_SLOTS, suse = chunk / EXACT_SLOTS;
        mbar_wait(&sfull_bar[slot], suse & 1);
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(lg * 32) << 16) + slot * BLOCK_N + (uint32_t)cbase;
        if constexpr (BLOCK_N >= 64) {
#pragma unroll
          for (int c0 = 0; c0 < COLS; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(t_row + (uint32_t)c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) acc[c0 + i] += __uint_as_float(v[i]);
          }
        } else if (owner) {
          uint32_t v[32];
          tmem_ld16(t_row, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) acc[i] += __uint_as_float(v[i]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sempty_bar[slot]);
      }
      if constexpr (SPLIT > 1) {
        // [peer][column / 4][row][4]: one 16-byte vector per row and column quad, rows (lanes) consecutive
        const uint32_t my_off = (uint32_t)(((cbase / 4) * BLOCK_M + lg * 32 + lane) * 16);
        if (crank != 0) {
          mbar_wait_cluster(red_free, (uint32_t)(local & 1) ^ 1u);       // CTA 0 has read the previous tile's
          if (owner) {
            const uint32_t dst = mapa_u32(smem_u32(red), 0) + (uint32_t)(((int)crank - 1) * (BLOCK_N / 4) * BLOCK_M * 16) + my_off;
#pragma unroll
            for (int i = 0; i < COLS; i += 4)
              st_cluster_v4(dst + (uint32_t)((i / 4) * BLOCK_M * 16), __float_as_uint(acc[i]), __float_as_uint(acc[i + 1]),
                            __float_as_uint(acc[i + 2]), __float_as_uint(acc[i + 3]));
          }
          fence_acq_rel_cluster();
          __syncwarp();
          if (lane == 0) mbar_arrive_remote_release(red_full, 0);
          continue;
        }
        mbar_wait_cluster(red_full, (uint32_t)(local & 1));
        if (owner) {
          const uint32_t src = smem_u32(red) + my_off;
#pragma unroll
          for (int sp = 0; sp < SPLIT - 1; ++sp) {
#pragma unroll
            for (int i = 0; i < COLS; i += 4) {
              const float4 u = lds128_f32(src + (uint32_t)((sp * (BLOCK_N / 4) + i / 4) * BLOCK_M * 16));
              acc[i] += u.x; acc[i + 1] += u.y; acc[i + 2] += u.z; acc[i + 3] += u.w;
            }
          }
        }
        __syncwarp();
        if (lane == 0) {                                   // the peers may overwrite their part of `red`
          for (uint32_t c = 1; c < (uint32_t)SPLIT; ++c) mbar_arrive_remote_release(red_free, c);
        }
      }
      // epilogue: bias, ReLU, fp32 compact store (P,h,w,n_real)
      const int r = m0 + lg * 32 + lane;
      const int q = r % rp;
      const int y = q / wp, x = q % wp;
      float vmax = 0.f;                                 // largest |output| of this thread's row part (amax_out)
      const int item_of_row = r / rp;
      if (SCHEME == 1 && p.split_out) {
        // the result leaves as the next conv's two fp16 planes (all rows of the tile: pad rows and pad columns are zeros)
        if (owner && r < p.rows) {
          const bool is_pixel = x < p.w && y < p.h;
          const float oscale = wscale_inv / pow2_scale_from_amax(__ldg(p.amax_a + item_of_row));
          const float nscale = pow2_scale_from_amax(__ldg(p.split_scale + item_of_row));
          uint16_t* dst = p.split_out + (size_t)r * (2 * (size_t)p.n_pad) + n0;
#pragma unroll
          for (int i = 0; i < COLS; i += 8) {
            uint32_t hi[4], lo[4];
#pragma unroll
            for (int k = 0; k < 8; k += 2) {
              float a0 = acc[i + k] * oscale, a1 = acc[i + k + 1] * oscale;
              if (p.relu) { a0 = fmaxf(a0, 0.f); a1 = fmaxf(a1, 0.f); }
              if (!is_pixel) { a0 = 0.f; a1 = 0.f; }
              vmax = fmaxf(vmax, fmaxf(fabsf(a0), fabsf(a1)));
              a0 *= nscale; a1 *= nscale;
              const __half2 h2 = __floats2half2_rn(a0, a1);
              const float2 hf = __half22float2(h2);
              const __half2 l2 = __floats2half2_rn(a0 - hf.x, a1 - hf.y);
              hi[k >> 1] = *reinterpret_cast<const uint32_t*>(&h2);
              lo[k >> 1] = *reinterpret_cast<const uint32_t*>(&l2);
            }
            *reinterpret_cast<uint4*>(dst + i) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(dst + p.n_pad + i) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
          }
        }
      } else if (owner && r < p.rows && x < p.w && y < p.h) {
        const int item = item_of_row;
        const float oscale = SCHEME == 0 ? 1.f : wscale_inv / pow2_scale_from_amax(__ldg(p.amax_a + item));   // per-item activation scale
        float* dst = p.out_f32 + (((size_t)item * p.h + y) * p.w + x) * p.n_real + n0;
        if ((p.n_real & 3) == 0 && n0 + COLS <= p.n_real) {
#pragma unroll
          for (int i = 0; i < COLS; i += 4) {
            float a[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              a[k] = SCHEME == 0 ? acc[i + k] : acc[i + k] * oscale;
              if (p.bias) a[k] += __ldg(p.bias + n0 + i + k);
              if (p.relu) a[k] = fmaxf(a[k], 0.f);
              vmax = fmaxf(vmax, fabsf(a[k]));
            }
            *reinterpret_cast<float4*>(dst + i) = make_float4(a[0], a[1], a[2], a[3]);
          }
        } else {
#pragma unroll
          for (int i = 0; i < COLS; ++i) {
            if (n0 + i < p.n_real) {
              float a = SCHEME == 0 ? acc[i] : acc[i] * oscale;
              if (p.bias) a += __ldg(p.bias + n0 + i);
              if (p.relu) a = fmaxf(a, 0.f);
              vmax = fmaxf(vmax, fabsf(a));
              dst[i] = a;
            }
          }
        }
      }
      if (SCHEME == 0 && p.split_out) {
        // forward tape: besides the fp32 activation, the result goes out as the NEXT conv's operand - three bf16 planes
        // [hi | mid | lo] per padded-flat row (the arithmetic of split3_pad_kernel, which this replaces); pad rows and
        // pad channels are written as zeros
        if (owner && r < p.rows) {
          const bool is_px = x < p.w && y < p.h;
          uint16_t* dst = p.split_out + (size_t)r * (3 * (size_t)p.n_pad) + n0;
#pragma unroll
          for (int i = 0; i < COLS; i += 8) {
            uint32_t hi[4], mi[4], lo[4];
#pragma unroll
            for (int k = 0; k < 8; k += 2) {
              float a2[2];
#pragma unroll
              for (int e2 = 0; e2 < 2; ++e2) {
                float a = acc[i + k + e2];
                const int c = n0 + i + k + e2;
                if (p.bias && c < p.n_real) a += __ldg(p.bias + c);
                if (p.relu) a = fmaxf(a, 0.f);
                a2[e2] = (is_px && c < p.n_real) ? a : 0.f;
              }
              const __nv_bfloat162 h2 = __floats2bfloat162_rn(a2[0], a2[1]);
              const float2 hf = __bfloat1622float2(h2);
              const float r0 = a2[0] - hf.x, r1 = a2[1] - hf.y;
              const __nv_bfloat162 m2 = __floats2bfloat162_rn(r0, r1);
              const float2 mf = __bfloat1622float2(m2);
              const __nv_bfloat162 l2 = __floats2bfloat162_rn(r0 - mf.x, r1 - mf.y);
              hi[k >> 1] = *reinterpret_cast<const uint32_t*>(&h2);
              mi[k >> 1] = *reinterpret_cast<const uint32_t*>(&m2);
              lo[k >> 1] = *reinterpret_cast<const uint32_t*>(&l2);
            }
            *reinterpret_cast<uint4*>(dst + i) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(dst + p.n_pad + i) = make_uint4(mi[0], mi[1], mi[2], mi[3]);
            *reinterpret_cast<uint4*>(dst + 2 * p.n_pad + i) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
          }
        }
      }
      if (p.amax_out) {
        // one atomic per warp when the whole warp sits in one item (the rule; tiles straddle items only at their ends)
        const int item0 = __shfl_sync(0xffffffffu, item_of_row, 0);
        if (__all_sync(0xffffffffu, item_of_row == item0)) {
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
          if (lane == 0 && vmax > 0.f) atomicMax(p.amax_out + item0, __float_as_uint(vmax));
        } else if (vmax > 0.f) {
          atomicMax(p.amax_out + item_of_row, __float_as_uint(vmax));
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
