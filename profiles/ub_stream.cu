// Micro-benchmark behind the choice of the roll / bilinear-resize kernel forms (not part of the library).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/ub_stream profiles/ub_stream.cu && /tmp/ub_stream
// Every variant moves the same algorithmic bytes; buffers (403 MB out) are far larger than the 126 MB L2.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#ifdef STCS
#define STORE(p, v) __stcs(p, v)
#else
#define STORE(p, v) (*(p) = (v))
#endif
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

// ---------------------------------------------------------------- roll ----------------------------------
// A: 4 scalar loads (lane stride 16 B) + one 16-byte store
__global__ void __launch_bounds__(256) roll_a(const float* __restrict__ x, int H, int wc, int oy, int shift, float* __restrict__ out) {
  const int y = blockIdx.y % H, b = blockIdx.y / H;
  int sy = (y - oy) % H; if (sy < 0) sy += H;
  const float* src = x + ((size_t)b * H + sy) * wc;
  float* dst = out + ((size_t)b * H + y) * wc;
  for (int f = (blockIdx.x * blockDim.x + threadIdx.x) * 4; f < wc; f += gridDim.x * blockDim.x * 4) {
    float v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { int g = f + k - shift; if (g < 0) g += wc; v[k] = __ldg(src + g); }
    *reinterpret_cast<float4*>(dst + f) = make_float4(v[0], v[1], v[2], v[3]);
  }
}
// B/C: U independent scalar loads, U scalar stores
template <int U>
__global__ void __launch_bounds__(256) roll_s(const float* __restrict__ x, int H, int wc, int oy, int shift, float* __restrict__ out) {
  const int y = blockIdx.y % H, b = blockIdx.y / H;
  int sy = (y - oy) % H; if (sy < 0) sy += H;
  const float* src = x + ((size_t)b * H + sy) * wc;
  float* dst = out + ((size_t)b * H + y) * wc;
  const int ts = gridDim.x * blockDim.x;
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  for (; f + (U - 1) * ts < wc; f += U * ts) {
    float v[U];
#pragma unroll
    for (int k = 0; k < U; ++k) { int g = f + k * ts - shift; if (g < 0) g += wc; v[k] = __ldg(src + g); }
#pragma unroll
    for (int k = 0; k < U; ++k) dst[f + k * ts] = v[k];
  }
  for (; f < wc; f += ts) { int g = f - shift; if (g < 0) g += wc; dst[f] = __ldg(src + g); }
}
// D: aligned 16-byte loads of a ROWS x row tile into shared memory, shifted scalar reads, 16-byte stores.
// One block = ROWS rows of wc floats (wc*ROWS*4 bytes of dynamic shared memory).
template <int ROWS>
__global__ void __launch_bounds__(512) roll_d(const float* __restrict__ x, int H, int wc, int oy, int shift, float* __restrict__ out) {
  extern __shared__ float4 sm4[];
  float* sm = reinterpret_cast<float*>(sm4);
  const int row0 = blockIdx.x * ROWS;                       // over B*H rows
  const int n4 = wc / 4;
  for (int i = threadIdx.x; i < ROWS * n4; i += blockDim.x) {
    const int r = i / n4, c = i - r * n4;
    const int gy = row0 + r, y = gy % H, b = gy / H;
    int sy = (y - oy) % H; if (sy < 0) sy += H;
    sm4[i] = __ldg(reinterpret_cast<const float4*>(x + ((size_t)b * H + sy) * wc) + c);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < ROWS * n4; i += blockDim.x) {
    const int r = i / n4, c = i - r * n4;
    const float* s = sm + r * wc;
    int g = c * 4 - shift; if (g < 0) g += wc;
    float v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { int gg = g + k; if (gg >= wc) gg -= wc; v[k] = s[gg]; }
    reinterpret_cast<float4*>(out + (size_t)(row0 + r) * wc)[c] = make_float4(v[0], v[1], v[2], v[3]);
  }
}
// E: plain 16-byte copy (upper bound)
__global__ void __launch_bounds__(256) copy_e(const float4* __restrict__ x, size_t n4, float4* __restrict__ out) {
  const size_t ts = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + 3 * ts < n4; i += 4 * ts) {
    const float4 a = __ldg(x + i), b = __ldg(x + i + ts), c = __ldg(x + i + 2 * ts), d = __ldg(x + i + 3 * ts);
    out[i] = a; out[i + ts] = b; out[i + 2 * ts] = c; out[i + 3 * ts] = d;
  }
  for (; i < n4; i += ts) out[i] = __ldg(x + i);
}
// F: per row, 16-byte loads from the aligned source, element funnel through warp shuffles, 16-byte stores.
// Thread t of a warp owns destination float4 q; it needs source floats [4q - shift, 4q - shift + 4), which straddle
// two aligned source float4s: A = floor((4q - shift)/4) and A+1.  Lane t loads source float4 (base + t) for its
// warp's span (33 float4 needed: lane 0 also loads the 33rd), then picks components with shuffles.
__global__ void __launch_bounds__(256) roll_f(const float* __restrict__ x, int H, int wc, int oy, int shift, float* __restrict__ out) {
  const int y = blockIdx.y % H, b = blockIdx.y / H;
  int sy = (y - oy) % H; if (sy < 0) sy += H;
  const float4* src4 = reinterpret_cast<const float4*>(x + ((size_t)b * H + sy) * wc);
  float4* dst4 = reinterpret_cast<float4*>(out + ((size_t)b * H + y) * wc);
  const int n4 = wc / 4;
  const int lane = threadIdx.x & 31;
  const int sh4 = shift >> 2, r = shift & 3;               // shift = 4*sh4 + r
  // destination float4 q takes floats 4q - shift .. +3 = 4(q - sh4) - r .. : source float4 (q - sh4 - 1) comps (4-r)..3
  // and source float4 (q - sh4) comps 0..(3-r)   (r == 0: only the latter)
  for (int q0 = (blockIdx.x * blockDim.x + threadIdx.x) - lane; q0 < n4; q0 += gridDim.x * blockDim.x) {
    const int q = q0 + lane;
    int a = q - sh4; a %= n4; if (a < 0) a += n4;
    const float4 cur = __ldg(src4 + a);                                         // source float4 (q - sh4)
    // previous source float4 comes from lane-1, lane 0 loads it itself
    float4 prev;
    prev.x = __shfl_up_sync(0xffffffffu, cur.x, 1); prev.y = __shfl_up_sync(0xffffffffu, cur.y, 1);
    prev.z = __shfl_up_sync(0xffffffffu, cur.z, 1); prev.w = __shfl_up_sync(0xffffffffu, cur.w, 1);
    if (lane == 0) { int p = a - 1; if (p < 0) p += n4; prev = __ldg(src4 + p); }
    float4 o;
    if (r == 0) o = cur;
    else if (r == 1) o = make_float4(prev.w, cur.x, cur.y, cur.z);
    else if (r == 2) o = make_float4(prev.z, prev.w, cur.x, cur.y);
    else o = make_float4(prev.y, prev.z, prev.w, cur.x);
    if (q < n4) dst4[q] = o;
  }
}

// ---------------------------------------------------------------- resize --------------------------------
struct RS { int H, W, C, OH, OW; };
__device__ __forceinline__ void coords(float scale, int o, int n, int& i0, int& i1, float& l) {
  float f = scale * ((float)o + 0.5f) - 0.5f; if (f < 0.f) f = 0.f;
  i0 = (int)f; if (i0 > n - 1) i0 = n - 1;
  i1 = i0 + (i0 < n - 1 ? 1 : 0);
  l = f - (float)i0;
}
// A: one thread per output pixel (3 channels)
__global__ void __launch_bounds__(256) rs_a(const float* __restrict__ x, RS p, float* __restrict__ out) {
  const int oy = blockIdx.y % p.OH, b = blockIdx.y / p.OH;
  int y0, y1; float ly; coords((float)p.H / p.OH, oy, p.H, y0, y1, ly); const float hy = 1.f - ly;
  const float* r0 = x + ((size_t)b * p.H + y0) * p.W * p.C;
  const float* r1 = x + ((size_t)b * p.H + y1) * p.W * p.C;
  float* o = out + ((size_t)b * p.OH + oy) * p.OW * p.C;
  const int ox = blockIdx.x * blockDim.x + threadIdx.x;
  if (ox >= p.OW) return;
  int x0, x1; float lx; coords((float)p.W / p.OW, ox, p.W, x0, x1, lx); const float hx = 1.f - lx;
  for (int c = 0; c < p.C; ++c)
    o[ox * p.C + c] = hy * (hx * __ldg(r0 + x0 * p.C + c) + lx * __ldg(r0 + x1 * p.C + c)) + ly * (hx * __ldg(r1 + x0 * p.C + c) + lx * __ldg(r1 + x1 * p.C + c));
}
// C/D: U output elements per thread (stride blockDim), 4U loads in flight
template <int U>
__global__ void __launch_bounds__(256) rs_e(const float* __restrict__ x, RS p, float* __restrict__ out) {
  const int oy = blockIdx.y % p.OH, b = blockIdx.y / p.OH;
  int y0, y1; float ly; coords((float)p.H / p.OH, oy, p.H, y0, y1, ly); const float hy = 1.f - ly;
  const float* r0 = x + ((size_t)b * p.H + y0) * p.W * p.C;
  const float* r1 = x + ((size_t)b * p.H + y1) * p.W * p.C;
  float* o = out + ((size_t)b * p.OH + oy) * p.OW * p.C;
  const int n = p.OW * p.C;
  const int f0 = blockIdx.x * blockDim.x * U + threadIdx.x;
  float v[U][4], lxs[U];
#pragma unroll
  for (int k = 0; k < U; ++k) {
    const int f = f0 + k * blockDim.x;
    if (f < n) {
      const int ox = f / p.C, c = f - ox * p.C;
      int x0, x1; coords((float)p.W / p.OW, ox, p.W, x0, x1, lxs[k]);
      v[k][0] = __ldg(r0 + x0 * p.C + c); v[k][1] = __ldg(r0 + x1 * p.C + c);
      v[k][2] = __ldg(r1 + x0 * p.C + c); v[k][3] = __ldg(r1 + x1 * p.C + c);
    }
  }
#pragma unroll
  for (int k = 0; k < U; ++k) {
    const int f = f0 + k * blockDim.x;
    if (f < n) { const float lx = lxs[k], hx = 1.f - lx; o[f] = hy * (hx * v[k][0] + lx * v[k][1]) + ly * (hx * v[k][2] + lx * v[k][3]); }
  }
}
// B: strip walk (STRIP rows per thread, horizontal interpolation cached)
template <int STRIP>
__global__ void __launch_bounds__(256) rs_b(const float* __restrict__ x, RS p, float* __restrict__ out) {
  const int strips = (p.OH + STRIP - 1) / STRIP;
  const int b = blockIdx.y / strips, oyb = (blockIdx.y % strips) * STRIP, oye = min(oyb + STRIP, p.OH);
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= p.OW * p.C) return;
  const int ox = f / p.C, c = f - ox * p.C;
  int x0, x1; float lx; coords((float)p.W / p.OW, ox, p.W, x0, x1, lx); const float hx = 1.f - lx;
  const int i0 = x0 * p.C + c, i1 = x1 * p.C + c;
  const float* img = x + (size_t)b * p.H * p.W * p.C;
  const size_t wc = (size_t)p.W * p.C;
  float* o = out + ((size_t)b * p.OH + oyb) * p.OW * p.C + f;
  int cy0 = -1, cy1 = -1; float h0 = 0.f, h1 = 0.f;
  for (int oy = oyb; oy < oye; ++oy, o += (size_t)p.OW * p.C) {
    int y0, y1; float ly; coords((float)p.H / p.OH, oy, p.H, y0, y1, ly); const float hy = 1.f - ly;
    if (y0 != cy0) { if (y0 == cy1) h0 = h1; else { const float* r = img + y0 * wc; h0 = hx * __ldg(r + i0) + lx * __ldg(r + i1); } cy0 = y0; }
    if (y1 != cy1) { if (y1 == cy0) h1 = h0; else { const float* r = img + y1 * wc; h1 = hx * __ldg(r + i0) + lx * __ldg(r + i1); } cy1 = y1; }
    *o = hy * h0 + ly * h1;
  }
}
// G: strip walk with the source-row span of the strip preloaded: all horizontal interpolations of the strip's
// source rows are issued first (independent loads), then the vertical blend.  STRIP output rows need at most
// SPAN = STRIP + 2 source rows when enlarging.
template <int STRIP, int SPAN>
__global__ void __launch_bounds__(256) rs_g(const float* __restrict__ x, RS p, float* __restrict__ out) {
  const int strips = (p.OH + STRIP - 1) / STRIP;
  const int b = blockIdx.y / strips, oyb = (blockIdx.y % strips) * STRIP, oye = min(oyb + STRIP, p.OH);
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= p.OW * p.C) return;
  const int ox = f / p.C, c = f - ox * p.C;
  int x0, x1; float lx; coords((float)p.W / p.OW, ox, p.W, x0, x1, lx); const float hx = 1.f - lx;
  const int i0 = x0 * p.C + c, i1 = x1 * p.C + c;
  const float* img = x + (size_t)b * p.H * p.W * p.C;
  const size_t wc = (size_t)p.W * p.C;
  int ya, yb_, yz, yz1; float t;
  coords((float)p.H / p.OH, oyb, p.H, ya, yb_, t);
  coords((float)p.H / p.OH, oye - 1, p.H, yz, yz1, t);
  float a0[SPAN], a1[SPAN];
#pragma unroll
  for (int k = 0; k < SPAN; ++k) {
    const int yy = min(ya + k, p.H - 1);
    if (ya + k <= yz1) { const float* r = img + yy * wc; a0[k] = __ldg(r + i0); a1[k] = __ldg(r + i1); }
  }
  float h[SPAN];
#pragma unroll
  for (int k = 0; k < SPAN; ++k) h[k] = hx * a0[k] + lx * a1[k];
  float* o = out + ((size_t)b * p.OH + oyb) * p.OW * p.C + f;
#pragma unroll
  for (int j = 0; j < STRIP; ++j) {
    const int oy = oyb + j;
    if (oy < oye) {
      int y0, y1; float ly; coords((float)p.H / p.OH, oy, p.H, y0, y1, ly); const float hy = 1.f - ly;
      float v0 = 0.f, v1 = 0.f;
#pragma unroll
      for (int k = 0; k < SPAN; ++k) { if (y0 - ya == k) v0 = h[k]; if (y1 - ya == k) v1 = h[k]; }
      o[(size_t)j * p.OW * p.C] = hy * v0 + ly * v1;
    }
  }
}

// H: U elements per thread (stride blockDim), R output rows per block; x coordinates computed once per thread,
// y coordinates once per block (shared memory).  CACHE: re-use the horizontally interpolated source rows between
// consecutive output rows (warp-uniform branches).
template <int U, int R, bool CACHE>
__global__ void __launch_bounds__(256) rs_h(const float* __restrict__ x, RS p, float* __restrict__ out) {
  __shared__ int sy0[R], sy1[R];
  __shared__ float sly[R];
  const int strips = (p.OH + R - 1) / R;
  const int b = blockIdx.y / strips, oyb = (blockIdx.y % strips) * R;
  const int rows = min(R, p.OH - oyb);
  if (threadIdx.x < R) {
    int y0, y1; float ly; coords((float)p.H / p.OH, min(oyb + (int)threadIdx.x, p.OH - 1), p.H, y0, y1, ly);
    sy0[threadIdx.x] = y0; sy1[threadIdx.x] = y1; sly[threadIdx.x] = ly;
  }
  const int n = p.OW * p.C;
  const int f0 = blockIdx.x * 256 * U + threadIdx.x;
  int i0[U], i1[U]; float lx[U];
  const float sw = (float)p.W / p.OW;
#pragma unroll
  for (int k = 0; k < U; ++k) {
    const int f = min(f0 + k * 256, n - 1);
    const int ox = f / p.C, c = f - ox * p.C;
    int x0, x1; coords(sw, ox, p.W, x0, x1, lx[k]);
    i0[k] = x0 * p.C + c; i1[k] = x1 * p.C + c;
  }
  __syncthreads();
  const float* img = x + (size_t)b * p.H * p.W * p.C;
  const size_t wc = (size_t)p.W * p.C;
  float* o = out + ((size_t)b * p.OH + oyb) * n;
  if (CACHE) {
    int cy0 = -1, cy1 = -1;
    float h0[U], h1[U];
    for (int r = 0; r < rows; ++r, o += n) {
      const int y0 = sy0[r], y1 = sy1[r]; const float ly = sly[r], hy = 1.f - ly;
      const bool need0 = (y0 != cy0) && (y0 != cy1), need1 = (y1 != cy1) && (y1 != y0 || true) && !(y1 == y0);
      // new h0
      float n0[U], n1[U];
      if (y0 == cy0) {
#pragma unroll
        for (int k = 0; k < U; ++k) n0[k] = h0[k];
      } else if (y0 == cy1) {
#pragma unroll
        for (int k = 0; k < U; ++k) n0[k] = h1[k];
      } else {
        const float* rp = img + y0 * wc;
        float a[U], bb[U];
#pragma unroll
        for (int k = 0; k < U; ++k) { a[k] = __ldg(rp + i0[k]); bb[k] = __ldg(rp + i1[k]); }
#pragma unroll
        for (int k = 0; k < U; ++k) n0[k] = (1.f - lx[k]) * a[k] + lx[k] * bb[k];
      }
      if (y1 == y0) {
#pragma unroll
        for (int k = 0; k < U; ++k) n1[k] = n0[k];
      } else if (y1 == cy1) {
#pragma unroll
        for (int k = 0; k < U; ++k) n1[k] = h1[k];
      } else {
        const float* rp = img + y1 * wc;
        float a[U], bb[U];
#pragma unroll
        for (int k = 0; k < U; ++k) { a[k] = __ldg(rp + i0[k]); bb[k] = __ldg(rp + i1[k]); }
#pragma unroll
        for (int k = 0; k < U; ++k) n1[k] = (1.f - lx[k]) * a[k] + lx[k] * bb[k];
      }
      (void)need0; (void)need1;
#pragma unroll
      for (int k = 0; k < U; ++k) { h0[k] = n0[k]; h1[k] = n1[k]; }
      cy0 = y0; cy1 = y1;
#pragma unroll
      for (int k = 0; k < U; ++k) if (f0 + k * 256 < n) STORE(&o[f0 + k * 256], hy * h0[k] + ly * h1[k]);
    }
  } else {
    for (int r = 0; r < rows; ++r, o += n) {
      const float* r0 = img + sy0[r] * wc; const float* r1 = img + sy1[r] * wc;
      const float ly = sly[r], hy = 1.f - ly;
      float v[U][4];
#pragma unroll
      for (int k = 0; k < U; ++k) { v[k][0] = __ldg(r0 + i0[k]); v[k][1] = __ldg(r0 + i1[k]); v[k][2] = __ldg(r1 + i0[k]); v[k][3] = __ldg(r1 + i1[k]); }
#pragma unroll
      for (int k = 0; k < U; ++k) {
        const float hx = 1.f - lx[k];
        if (f0 + k * 256 < n) o[f0 + k * 256] = hy * (hx * v[k][0] + lx[k] * v[k][1]) + ly * (hx * v[k][2] + lx[k] * v[k][3]);
      }
    }
  }
}

// S: two passes through shared memory used as a dynamically indexed, thread-private register file: first the
// horizontal interpolation of every distinct source row of the strip (independent loads), then the vertical
// blend of the strip's output rows.  A thread only ever reads the columns it wrote, so no barrier between passes.
template <int U, int R>
__global__ void __launch_bounds__(256) rs_s(const float* __restrict__ x, RS p, int rows_per_strip, float* __restrict__ out) {
  constexpr int S = R + 2;
  extern __shared__ float hs[];                    // [S][256 * U]
  __shared__ int srow[S], sa0[R], sa1[R], sns;
  __shared__ float sly[R];
  const int strips = (p.OH + rows_per_strip - 1) / rows_per_strip;
  const int b = blockIdx.y / strips, oyb = (blockIdx.y % strips) * rows_per_strip;
  const int rows = min(rows_per_strip, p.OH - oyb);
  if (threadIdx.x == 0) {
    int ns = 0;
    const float sh = (float)p.H / p.OH;
    for (int r = 0; r < rows; ++r) {
      int y0, y1; float ly; coords(sh, oyb + r, p.H, y0, y1, ly);
      int a0, a1;
      if (ns > 0 && srow[ns - 1] == y0) a0 = ns - 1;
      else if (ns > 1 && srow[ns - 2] == y0) a0 = ns - 2;
      else { srow[ns] = y0; a0 = ns++; }
      if (y1 == y0) a1 = a0;
      else if (ns > 0 && srow[ns - 1] == y1) a1 = ns - 1;
      else { srow[ns] = y1; a1 = ns++; }
      sa0[r] = a0; sa1[r] = a1; sly[r] = ly;
    }
    sns = ns;
  }
  const int n = p.OW * p.C;
  const int f0 = blockIdx.x * 256 * U + threadIdx.x;
  const float sw = (float)p.W / p.OW;
  int i0[U], i1[U]; float lx[U], hx[U];
#pragma unroll
  for (int k = 0; k < U; ++k) {
    const int f = min(f0 + k * 256, n - 1);
    const int ox = f / p.C, c = f - ox * p.C;
    int x0, x1; coords(sw, ox, p.W, x0, x1, lx[k]);
    hx[k] = 1.f - lx[k];
    i0[k] = x0 * p.C + c; i1[k] = x1 * p.C + c;
  }
  __syncthreads();
  const float* img = x + (size_t)b * p.H * p.W * p.C;
  const size_t wc = (size_t)p.W * p.C;
  const int ns = sns;
  float* col = hs + threadIdx.x;
#pragma unroll 4
  for (int j = 0; j < ns; ++j) {
    const float* rp = img + (size_t)srow[j] * wc;
#pragma unroll
    for (int k = 0; k < U; ++k) col[(j * U + k) * 256] = hx[k] * __ldg(rp + i0[k]) + lx[k] * __ldg(rp + i1[k]);
  }
  float* o = out + ((size_t)b * p.OH + oyb) * n + f0;
#pragma unroll 4
  for (int r = 0; r < rows; ++r, o += n) {
    const int a0 = sa0[r] * U * 256, a1 = sa1[r] * U * 256;
    const float ly = sly[r], hy = 1.f - ly;
#pragma unroll
    for (int k = 0; k < U; ++k)
      if (f0 + k * 256 < n) o[k * 256] = hy * col[a0 + k * 256] + ly * col[a1 + k * 256];
  }
}

// P: the H/cache form with the address arithmetic taken out of the row loop: per-thread source pointers that are
// advanced by a warp-uniform byte delta when a new source row is needed, one store pointer with immediate offsets,
// bounds predicates and (1 - l) weights computed once.
template <int U, int R>
__global__ void __launch_bounds__(256) rs_p(const float* __restrict__ x, RS p, float* __restrict__ out) {
  __shared__ int sy0[R], sy1[R];
  __shared__ float sly[R], shy[R];
  const int strips = (p.OH + R - 1) / R;
  const int b = blockIdx.y / strips, oyb = (blockIdx.y % strips) * R;
  const int rows = min(R, p.OH - oyb);
  if (threadIdx.x < R) {
    int y0, y1; float ly; coords((float)p.H / p.OH, min(oyb + (int)threadIdx.x, p.OH - 1), p.H, y0, y1, ly);
    sy0[threadIdx.x] = y0; sy1[threadIdx.x] = y1; sly[threadIdx.x] = ly; shy[threadIdx.x] = 1.f - ly;
  }
  const int n = p.OW * p.C;
  const int f0 = blockIdx.x * 256 * U + threadIdx.x;
  const float sw = (float)p.W / p.OW;
  const float* img = x + (size_t)b * p.H * p.W * p.C;
  const float* q0[U]; const float* q1[U];
  float lx[U], hx[U];
  bool ok[U];
#pragma unroll
  for (int k = 0; k < U; ++k) {
    ok[k] = f0 + k * 256 < n;
    const int f = min(f0 + k * 256, n - 1);
    const int ox = f / p.C, c = f - ox * p.C;
    int x0, x1; coords(sw, ox, p.W, x0, x1, lx[k]);
    hx[k] = 1.f - lx[k];
    q0[k] = img + x0 * p.C + c; q1[k] = img + x1 * p.C + c;
  }
  __syncthreads();
  const long long row_bytes = (long long)p.W * p.C;     // in floats
  int prow = 0;
  float* o = out + ((size_t)b * p.OH + oyb) * n + f0;
  int cy0 = -1, cy1 = -1;
  float h0[U], h1[U];
#pragma unroll
  for (int k = 0; k < U; ++k) { h0[k] = 0.f; h1[k] = 0.f; }
  for (int r = 0; r < rows; ++r, o += n) {
    const int y0 = sy0[r], y1 = sy1[r];
    if (y0 != cy0) {
      if (y0 == cy1) {
#pragma unroll
        for (int k = 0; k < U; ++k) h0[k] = h1[k];
      } else {
        const long long d = (long long)(y0 - prow) * row_bytes; prow = y0;
#pragma unroll
        for (int k = 0; k < U; ++k) { q0[k] += d; q1[k] += d; }
#pragma unroll
        for (int k = 0; k < U; ++k) h0[k] = hx[k] * __ldg(q0[k]) + lx[k] * __ldg(q1[k]);
      }
      cy0 = y0;
    }
    if (y1 != cy1) {
      if (y1 == cy0) {
#pragma unroll
        for (int k = 0; k < U; ++k) h1[k] = h0[k];
      } else {
        const long long d = (long long)(y1 - prow) * row_bytes; prow = y1;
#pragma unroll
        for (int k = 0; k < U; ++k) { q0[k] += d; q1[k] += d; }
#pragma unroll
        for (int k = 0; k < U; ++k) h1[k] = hx[k] * __ldg(q0[k]) + lx[k] * __ldg(q1[k]);
      }
      cy1 = y1;
    }
    const float ly = sly[r], hy = shy[r];
#pragma unroll
    for (int k = 0; k < U; ++k) if (ok[k]) o[k * 256] = hy * h0[k] + ly * h1[k];
  }
}

__global__ void fill_random(float* x, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    unsigned h = (unsigned)i * 2654435761u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    x[i] = (float)(int)(h & 0xFFFF) * (1.f / 65536.f) - 0.5f;
  }
}

template <class F> float time_ms(F f, int reps = 10) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); f();
  CK(cudaDeviceSynchronize());
  cudaEventRecord(a);
  for (int i = 0; i < reps; ++i) f();
  cudaEventRecord(b);
  CK(cudaEventSynchronize(b));
  float ms; cudaEventElapsedTime(&ms, a, b);
  CK(cudaGetLastError());
  return ms / reps;
}

int main() {
  const int B = 8, H = 2048, W = 2048, C = 3, wc = W * C;
  const size_t n = (size_t)B * H * W * C;
  float *x, *y;
  CK(cudaMalloc(&x, n * 4)); CK(cudaMalloc(&y, n * 4));
  CK(cudaMemset(x, 1, n * 4)); CK(cudaMemset(y, 0, n * 4));
#ifdef RANDOM_DATA
  fill_random<<<148 * 8, 256>>>(x, n);          // data-dependent effects (DRAM / datapath toggling) show up with this
  CK(cudaDeviceSynchronize());
#endif
  const int oy = 17, shift = ((-29 % W) + W) % W * C;
  const double gb = 2.0 * n * 4 / 1e9;
  auto rep = [&](const char* name, float ms, double g) { printf("%-44s %7.3f ms  %7.1f GB/s\n", name, ms, g / ms * 1e3); };
  rep("copy 16 B x4 (upper bound)", time_ms([&] { copy_e<<<148 * 16, 256>>>((const float4*)x, n / 4, (float4*)y); }), gb);
  rep("cudaMemcpyAsync D2D", time_ms([&] { cudaMemcpyAsync(y, x, n * 4, cudaMemcpyDeviceToDevice); }), gb);
  rep("roll A: 4 scalar ld + 16 B st", time_ms([&] { roll_a<<<dim3(7, B * H), 256>>>(x, H, wc, oy, shift, y); }), gb);
  rep("roll S4: 4 scalar ld/st", time_ms([&] { roll_s<4><<<dim3(6, B * H), 256>>>(x, H, wc, oy, shift, y); }), gb);
  rep("roll S8", time_ms([&] { roll_s<8><<<dim3(3, B * H), 256>>>(x, H, wc, oy, shift, y); }), gb);
  rep("roll S24 (one block per row)", time_ms([&] { roll_s<24><<<dim3(1, B * H), 256>>>(x, H, wc, oy, shift, y); }), gb);
  CK(cudaFuncSetAttribute(roll_d<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  CK(cudaFuncSetAttribute(roll_d<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  rep("roll D1: smem row, 512 thr", time_ms([&] { roll_d<1><<<B * H, 512, wc * 4>>>(x, H, wc, oy, shift, y); }), gb);
  rep("roll D2: smem 2 rows, 512 thr", time_ms([&] { roll_d<2><<<B * H / 2, 512, 2 * wc * 4>>>(x, H, wc, oy, shift, y); }), gb);
  rep("roll F: 16 B ld + shuffles + 16 B st", time_ms([&] { roll_f<<<dim3(6, B * H), 256>>>(x, H, wc, oy, shift, y); }), gb);
  rep("roll F: grid.x 2", time_ms([&] { roll_f<<<dim3(2, B * H), 256>>>(x, H, wc, oy, shift, y); }), gb);

  RS p{1462, 1462, 3, 2048, 2048};
  const double gr = 4.0 * ((double)B * 1462 * 1462 * 3 + (double)n) / 1e9;
  const int nrow = p.OW * p.C;
  rep("resize A: pixel/thread", time_ms([&] { rs_a<<<dim3((p.OW + 255) / 256, B * p.OH), 256>>>(x, p, y); }), gr);
  rep("resize E1: element/thread", time_ms([&] { rs_e<1><<<dim3((nrow + 255) / 256, B * p.OH), 256>>>(x, p, y); }), gr);
  rep("resize E4: 4 elements/thread", time_ms([&] { rs_e<4><<<dim3((nrow + 1023) / 1024, B * p.OH), 256>>>(x, p, y); }), gr);
  rep("resize E8", time_ms([&] { rs_e<8><<<dim3((nrow + 2047) / 2048, B * p.OH), 256>>>(x, p, y); }), gr);
  rep("resize B16: strip walk", time_ms([&] { rs_b<16><<<dim3((nrow + 255) / 256, B * ((p.OH + 15) / 16)), 256>>>(x, p, y); }), gr);
  rep("resize B64", time_ms([&] { rs_b<64><<<dim3((nrow + 255) / 256, B * ((p.OH + 63) / 64)), 256>>>(x, p, y); }), gr);
  rep("resize G8: strip, preloaded span", time_ms([&] { rs_g<8, 8><<<dim3((nrow + 255) / 256, B * ((p.OH + 7) / 8)), 256>>>(x, p, y); }), gr);
  rep("resize G16", time_ms([&] { rs_g<16, 14><<<dim3((nrow + 255) / 256, B * ((p.OH + 15) / 16)), 256>>>(x, p, y); }), gr);
#define RSH(U, R, CA) rep("resize H U=" #U " R=" #R " cache=" #CA, time_ms([&] { rs_h<U, R, CA><<<dim3((nrow + 256 * U - 1) / (256 * U), B * ((p.OH + R - 1) / R)), 256>>>(x, p, y); }), gr)
  RSH(4, 8, false); RSH(4, 16, false); RSH(2, 16, false); RSH(8, 8, false); RSH(1, 16, false); RSH(4, 32, false);
  RSH(4, 8, true); RSH(4, 16, true); RSH(2, 16, true); RSH(8, 16, true); RSH(1, 16, true); RSH(4, 32, true); RSH(2, 64, true); RSH(2, 32, true); RSH(3, 16, true); RSH(2, 8, true);
#define RSS(U, R) do { CK(cudaFuncSetAttribute(rs_s<U, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (R + 2) * 256 * U * 4)); \
  rep("resize S (smem) U=" #U " R=" #R, time_ms([&] { rs_s<U, R><<<dim3((nrow + 256 * U - 1) / (256 * U), B * ((p.OH + R - 1) / R)), 256, (R + 2) * 256 * U * 4>>>(x, p, R, y); }), gr); } while (0)
#define RSP(U, R) rep("resize P (pointers) U=" #U " R=" #R, time_ms([&] { rs_p<U, R><<<dim3((nrow + 256 * U - 1) / (256 * U), B * ((p.OH + R - 1) / R)), 256>>>(x, p, y); }), gr)
  RSP(2, 16); RSP(3, 16); RSP(4, 16); RSP(3, 32); RSP(6, 16); RSP(3, 8); RSP(2, 32); RSP(4, 32); RSP(6, 32);
  RSS(1, 16); RSS(2, 16); RSS(3, 16); RSS(1, 32); RSS(2, 32); RSS(1, 8); RSS(2, 8); RSS(4, 16);
  return 0;
}
