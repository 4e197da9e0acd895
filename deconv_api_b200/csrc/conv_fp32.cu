// fp32 direct 3x3 'same' convolution on CUDA cores (FFMA), NHWC activations, [9][Cin][Cout] weights.
// This is the PARITY-mode arithmetic (rel-L2 <= 1e-5 against the oracle) and the arithmetic of the
// forward tape in every mode: the switches recorded by the pools must come from fp32-accurate
// activations (bf16 activations flip ~0.5 % of the switches and cost 10-20 % rel-L2).
// The tensor-core (tcgen05) kernel for the backward chain lives in conv_tc.cu.
//
// Tiling: one CTA = 8x16 output pixels x 64 output channels, K streamed in chunks of 8 input
// channels x 9 taps through shared memory; one thread = 8 consecutive pixels of a row x 4 channels
// (32 accumulators).  Per (ci, kh) a thread loads 10 inputs + 3 float4 of weights for 96 FFMA, which
// keeps the kernel FFMA-issue-bound rather than LDS-bound (16 smem wavefronts vs 24 issue cycles).
#include "common.cuh"

namespace dcv {

namespace {

constexpr int TH = 8, TW = 16;        // output tile (pixels)
// TC (output channels per CTA) is a template parameter: 64, or 32 for launches that would otherwise leave most
// of the 148 SMs idle (the B=1 forward of the 28x28 / 14x14 layers).
constexpr int KC = 8;                 // input channels per smem chunk
constexpr int IN_W = TW + 2 + 2;      // 20: row pitch of the input patch (18 used)
constexpr int IN_PLANE = (TH + 2) * IN_W;

template <int TC>
__global__ void __launch_bounds__(256, 2)
conv3x3_fp32_kernel(const float* __restrict__ x, const float* __restrict__ w9, int H, int W, int cin, int cout,
                    int tiles_w, ConvEpilogue epi, float* __restrict__ out) {
  __shared__ __align__(16) float in_s[KC * IN_PLANE];        // [ci][row][col]
  __shared__ __align__(16) float w_s[9 * KC * TC];           // [tap][ci][co]

  const int tid = threadIdx.x;
  constexpr int NTX = TC / 4;         // threads along the channel axis (4 couts each)
  constexpr int NTY = 256 / NTX;      // pixel groups
  constexpr int PXT = (TH * TW) / NTY;  // consecutive pixels of one row per thread: 8 (TC=64) or 4 (TC=32)
  constexpr int GPR = TW / PXT;       // groups per tile row
  const int tx = tid % NTX;           // channel group: couts tx*4 .. tx*4+3
  const int ty = tid / NTX;
  const int row = ty / GPR;           // 0..7
  const int col0 = (ty % GPR) * PXT;
  const int tile_y = blockIdx.x / tiles_w, tile_x = blockIdx.x % tiles_w;
  const int y0 = tile_y * TH, x0 = tile_x * TW;
  const int co0 = blockIdx.y * TC;
  const size_t img_off = (size_t)blockIdx.z * H * W;
  const float* xin = x + img_off * cin;

  float acc[PXT][4];
#pragma unroll
  for (int p = 0; p < PXT; ++p)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[p][c] = 0.f;

  const bool vec_in = (cin % 4) == 0;
  const bool vec_w = (cout % 4) == 0;

  for (int c0 = 0; c0 < cin; c0 += KC) {
    // ---- stage the input patch: (TH+2) x (TW+2) pixels x KC channels, zero outside the image ----
    if (vec_in) {
      for (int i = tid; i < 2 * (TH + 2) * (TW + 2); i += 256) {
        const int half = i / ((TH + 2) * (TW + 2));
        const int pix = i % ((TH + 2) * (TW + 2));
        const int r = pix / (TW + 2), c = pix % (TW + 2);
        const int y = y0 + r - 1, xx = x0 + c - 1;
        const int ci = c0 + half * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (y >= 0 && y < H && xx >= 0 && xx < W && ci < cin)
          v = __ldg(reinterpret_cast<const float4*>(xin + ((size_t)y * W + xx) * cin + ci));
        float* d = in_s + (half * 4) * IN_PLANE + r * IN_W + c;
        d[0] = v.x; d[IN_PLANE] = v.y; d[2 * IN_PLANE] = v.z; d[3 * IN_PLANE] = v.w;
      }
    } else {
      for (int i = tid; i < KC * (TH + 2) * (TW + 2); i += 256) {
        const int k = i / ((TH + 2) * (TW + 2));
        const int pix = i % ((TH + 2) * (TW + 2));
        const int r = pix / (TW + 2), c = pix % (TW + 2);
        const int y = y0 + r - 1, xx = x0 + c - 1;
        const int ci = c0 + k;
        float v = 0.f;
        if (y >= 0 && y < H && xx >= 0 && xx < W && ci < cin) v = __ldg(xin + ((size_t)y * W + xx) * cin + ci);
        in_s[k * IN_PLANE + r * IN_W + c] = v;
      }
    }
    // ---- stage the weights: 9 taps x KC channels x TC couts -------------------------------------
    if (vec_w) {
      for (int i = tid; i < 9 * KC * (TC / 4); i += 256) {
        const int c4 = i % (TC / 4);
        const int k = (i / (TC / 4)) % KC;
        const int t = i / ((TC / 4) * KC);
        const int ci = c0 + k, co = co0 + c4 * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ci < cin && co < cout) v = __ldg(reinterpret_cast<const float4*>(w9 + ((size_t)t * cin + ci) * cout + co));
        *reinterpret_cast<float4*>(w_s + (t * KC + k) * TC + c4 * 4) = v;
      }
    } else {
      for (int i = tid; i < 9 * KC * TC; i += 256) {
        const int c = i % TC;
        const int k = (i / TC) % KC;
        const int t = i / (TC * KC);
        const int ci = c0 + k, co = co0 + c;
        float v = 0.f;
        if (ci < cin && co < cout) v = __ldg(w9 + ((size_t)t * cin + ci) * cout + co);
        w_s[(t * KC + k) * TC + c] = v;
      }
    }
    __syncthreads();

    // ---- FFMA ------------------------------------------------------------------------------------
#pragma unroll 2
    for (int k = 0; k < KC; ++k) {
#pragma unroll
      for (int kh = 0; kh < 3; ++kh) {
        const float* ip = in_s + k * IN_PLANE + (row + kh) * IN_W + col0;
        float iv[PXT + 2];
#pragma unroll
        for (int j = 0; j < PXT + 2; ++j) iv[j] = ip[j];
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
          const float4 wv = *reinterpret_cast<const float4*>(w_s + ((kh * 3 + kw) * KC + k) * TC + tx * 4);
#pragma unroll
          for (int p = 0; p < PXT; ++p) {
            acc[p][0] = fmaf(iv[p + kw], wv.x, acc[p][0]);
            acc[p][1] = fmaf(iv[p + kw], wv.y, acc[p][1]);
            acc[p][2] = fmaf(iv[p + kw], wv.z, acc[p][2]);
            acc[p][3] = fmaf(iv[p + kw], wv.w, acc[p][3]);
          }
        }
      }
    }
    __syncthreads();
  }

  // ---- epilogue: bias, ReLU / sign mask, seed add, store ------------------------------------------
  const int y = y0 + row;
  if (y >= H) return;
  const int co = co0 + tx * 4;
  if (co >= cout) return;
  float bv[4] = {0.f, 0.f, 0.f, 0.f};
  if (epi.bias) {
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (co + c < cout) bv[c] = __ldg(epi.bias + co + c);
  }
#pragma unroll
  for (int p = 0; p < PXT; ++p) {
    const int xx = x0 + col0 + p;
    if (xx >= W) break;
    const size_t o = (img_off + (size_t)y * W + xx) * cout + co;
    float r[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float v = acc[p][c] + bv[c];
      if (epi.relu) v = fmaxf(v, 0.f);
      r[c] = v;
    }
    if (epi.sign_mask) {
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (co + c < cout && !(__ldg(epi.sign_mask + o + c) > 0.f)) r[c] = 0.f;
    }
    if (epi.add) {
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (co + c < cout) r[c] = fmaf(epi.add_scale, __ldg(epi.add + o + c), r[c]);
    }
    if (vec_w) {
      *reinterpret_cast<float4*>(out + o) = make_float4(r[0], r[1], r[2], r[3]);
    } else {
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (co + c < cout) out[o + c] = r[c];
    }
  }
}

}  // namespace

void launch_conv3x3_fp32(const float* x, const float* w9, int B, int H, int W, int cin, int cout,
                         const ConvEpilogue& epi, float* out, cudaStream_t s) {
  if (B <= 0 || H <= 0 || W <= 0) return;
  const int tiles_w = (W + TW - 1) / TW, tiles_h = (H + TH - 1) / TH;
  // gridDim.z is limited to 65535 items per launch
  for (int b0 = 0; b0 < B; b0 += 65535) {
    const int nb = (B - b0 < 65535) ? (B - b0) : 65535;
    ConvEpilogue e = epi;
    const size_t ooff = (size_t)b0 * H * W * cout;
    if (e.sign_mask) e.sign_mask += ooff;
    if (e.add) e.add += ooff;
    const size_t ctas64 = (size_t)tiles_w * tiles_h * ((cout + 63) / 64) * nb;
    if (ctas64 < 2 * (size_t)kNumSMs && cout > 32) {
      dim3 grid(tiles_w * tiles_h, (cout + 31) / 32, nb);
      conv3x3_fp32_kernel<32><<<grid, 256, 0, s>>>(x + (size_t)b0 * H * W * cin, w9, H, W, cin, cout, tiles_w, e, out + ooff);
    } else {
      dim3 grid(tiles_w * tiles_h, (cout + 63) / 64, nb);
      conv3x3_fp32_kernel<64><<<grid, 256, 0, s>>>(x + (size_t)b0 * H * W * cin, w9, H, W, cin, cout, tiles_w, e, out + ooff);
    }
  }
}

}  // namespace dcv
