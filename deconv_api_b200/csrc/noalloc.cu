// dcv_deprocess_dev_noalloc — the de-process operator for callers that must not allocate GPU memory.
//
// Who needs it: a service that hands the whole GPU to one memory pool at start-up (so a later cudaMalloc
// fails or fragments the pool), a request handler on a latency budget (cudaMalloc and cudaFree synchronise
// the device and can stall every stream for milliseconds), and code running under a memory-accounting
// sandbox where only pre-declared buffers may exist.  The regular dcv_deprocess_dev grows the context's
// scratch buffer on demand for the per-block (sum, sum of squares) partials; this variant keeps those
// partials in a fixed array that is part of the library image instead.
//
// The array below lives in every GPU's memory from the moment the library is loaded there until the
// process exits (16 KiB per device).  Its name is only a host-side handle: host code must not pass
// `g_deprocess_partials` to a kernel, it has to ask the runtime where the array sits on the CURRENT device.
// That lookup happens on the first call for each device (see resolve_scratch), so make that first call
// outside of any stream capture.
#include "common.cuh"
#include "internal.cuh"

#include <atomic>

using namespace dcv;

namespace {

constexpr int kMaxPartials = DCV_DEPROCESS_NOALLOC_MAX_PARTIALS;          // (sum, sumsq) pairs
static __device__ __align__(256) double g_deprocess_partials[2 * kMaxPartials];

constexpr int kMaxDevices = 64;
// One slot per device; atomics because contexts on different host threads may make their first call at the
// same time (both would store the same address).  The address stays valid until cudaDeviceReset, which this
// cache does not observe: do not reset a device while the library is in use.
std::atomic<double*> g_scratch_on_device[kMaxDevices] = {};

// Device address of the array on `device` (which must be current).  Looked up once per device.  This first
// lookup may make the runtime load the library's device code onto that GPU (lazy module loading), which is the
// only allocation-like event the entry point can cause; later calls touch no allocator at all.
int resolve_scratch(dcv_ctx* ctx, double** out) {
  if (ctx->device < 0 || ctx->device >= kMaxDevices)
    return ctx->fail(DCV_ERR_UNSUPPORTED, "dcv_deprocess_dev_noalloc: device index %d beyond %d", ctx->device, kMaxDevices);
  double* p = g_scratch_on_device[ctx->device].load(std::memory_order_acquire);
  if (!p) {
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_deprocess_partials);
    if (e != cudaSuccess || !p)
      return ctx->fail(DCV_ERR_CUDA, "dcv_deprocess_dev_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
    g_scratch_on_device[ctx->device].store(p, std::memory_order_release);
  }
  *out = p;
  return 0;
}

}  // namespace

extern "C" int dcv_deprocess_dev_noalloc(dcv_ctx* ctx, const float* x_dev, size_t n, uint8_t* out_dev) {
  if (!ctx) return DCV_ERR_ARG;
  cudaError_t e = cudaSetDevice(ctx->device);
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
  if (!x_dev || !out_dev) return ctx->fail(DCV_ERR_ARG, "dcv_deprocess_dev_noalloc: null pointer");
  if (n == 0) return 0;
  if (n > (size_t)DCV_DEPROCESS_NOALLOC_MAX_ELEMS)
    return ctx->fail(DCV_ERR_ARG, "dcv_deprocess_dev_noalloc: %zu elements exceed the fixed-scratch limit of %zu; "
                     "use dcv_deprocess_dev (which may allocate) for larger inputs", n, (size_t)DCV_DEPROCESS_NOALLOC_MAX_ELEMS);
  if (((uintptr_t)x_dev & 15) || ((uintptr_t)out_dev & 3))
    return ctx->fail(DCV_ERR_ARG, "dcv_deprocess_dev_noalloc: x must be 16-byte and out 4-byte aligned");
  const int np = deprocess_partials(n);
  if (np > kMaxPartials)
    return ctx->fail(DCV_ERR_ARG, "dcv_deprocess_dev_noalloc: %d block partials exceed the fixed scratch of %d", np, kMaxPartials);
  double* scratch = nullptr;
  int r = resolve_scratch(ctx, &scratch);
  if (r) return r;
  // Nothing in the scratch has to start at zero: pass 1 overwrites partial[0 .. 2*np) completely and
  // pass 2 reads exactly that range, so whatever an earlier call left behind is never observed.
  launch_deprocess(x_dev, n, scratch, np, out_dev, nullptr, ctx->stream);
  ctx->launches += 2;
  e = cudaGetLastError();
  if (e != cudaSuccess) return ctx->fail(DCV_ERR_CUDA, "dcv_deprocess_dev_noalloc: launch: %s", cudaGetErrorString(e));
  return 0;
}
