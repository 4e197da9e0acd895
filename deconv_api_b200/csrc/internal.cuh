// Internal host-side types of the dcv library.
#pragma once
#include "common.cuh"

enum LayerKind { LK_CONV = 0, LK_POOL = 1, LK_FLATTEN = 2, LK_DENSE = 3 };

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(dcv_ctx* ctx, size_t bytes);   // grow-only; contents are NOT preserved on growth
  void release();
};

// tensor-core side of a conv layer (conv_tc.cu)
struct TcWeights {
  __nv_bfloat16* wd_bf16 = nullptr;   // backward weights packed [9][cin_pad(N)][cout_pad(K)] K-major bf16
  int n_pad = 0, k_pad = 0;
  void* tmap = nullptr;               // host copy of the CUtensorMap for the weights (128 B, 64 B aligned)
};

struct Layer {
  int kind = LK_CONV;
  // conv: cin/cout; dense: nin/nout
  int cin = 0, cout = 0, relu = 0, act_kind = 0;
  float* w = nullptr;        // conv: HWIO == [9][cin][cout]; dense: (nin, nout)
  float* bias = nullptr;
  float* wd = nullptr;       // conv backward kernel [9][cout][cin] (deepdream.py:80-81)
  TcWeights tc;
  // tape (valid for layers <= tape_upto)
  int B = 0, H = 0, W = 0, C = 0;      // output shape
  int in_H = 0, in_W = 0, in_C = 0;    // pool / flatten: input shape
  DevBuf act;                          // output, fp32
  DevBuf sw;                           // pool: switches, 1 byte per pooled element
};

struct ProfRec {
  cudaEvent_t a, b;
  int cat;
  double flops, bytes;
};

struct dcv_ctx {
  int device = 0;
  int num_sms = dcv::kNumSMs;
  cudaStream_t stream = nullptr;
  std::string err;
  long long launches = 0;
  size_t workspace_budget = 0;
  std::vector<Layer> net;
  int tape_upto = -1;
  int in_B = 0, in_H = 0, in_W = 0, in_C = 0;
  DevBuf input;
  DevBuf ws[6];        // projection workspace: [0],[1] ping-pong tensors, [2],[3] maps, [4],[5] bf16 ping-pong
  DevBuf feat_dev;
  DevBuf scratch;      // reduction partials
  DevBuf stage;        // staging for host-buffer entry points
  // optional per-kernel-class timing (dcv_profile): CUDA events around every launch of a class
  bool prof_on = false;
  std::vector<ProfRec> prof;
  std::vector<cudaEvent_t> ev_pool;
  void prof_begin(int cat, double flops, double bytes);
  void prof_end();
  int fail(int code, const char* fmt, ...);
};

struct ProfScope {
  dcv_ctx* c;
  ProfScope(dcv_ctx* ctx, int cat, double flops, double bytes) : c(ctx) { if (c->prof_on) c->prof_begin(cat, flops, bytes); }
  ~ProfScope() { if (c->prof_on) c->prof_end(); }
};

// conv_tc.cu
int tc_prepare_layer(dcv_ctx* ctx, Layer& l);
void tc_free_layer(Layer& l);
// backward conv of layer l on P items of (h, w, l.cout) fp32 -> (h, w, l.cin) fp32 with the tensor-core kernel
int tc_deconv(dcv_ctx* ctx, const Layer& l, const float* in, int P, int h, int w, float* out);
