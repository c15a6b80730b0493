// rnn_ops.cu — the gate of the DARTS recurrent cell, fused.
//
// Every node of the cell (reference: examples/hp_search_benchmarks/darts_penntreebank_pytorch/randomNAS_files/model.py:78-99)
// is a GEMM followed by a highway-style gate
//
//     [c, h] = (s_p ⊙ mask) W          s = s_p + σ(c) ⊙ (f(h) − s_p)          f ∈ {tanh, relu, sigmoid, identity}
//
// and through torch that gate is five elementwise launches forward and a dozen backward, per node, per time step: at
// batch 64 × width 850 each of them moves ~100 KB, so the recurrence (35 steps × 9 gates) is bound by launch and
// dependency latency, not by anything the hardware is good at. Here the gate of ALL nodes that share a predecessor (their
// GEMMs are already batched into one, see models/darts_rnn.py) is one launch forward and one backward:
//
//   fwd  reads ch [B, k·2·nh] and s_p [B, nh] once, writes the k node states and, when a dropout mask is given, their
//        masked copies (the next GEMM's operand), computing in fp32 and rounding once;
//   bwd  recomputes the gate from the same inputs (nothing but ch and s_p is saved), folds the gradients arriving through
//        a state and through its masked copy, and writes d(ch) for the whole group plus d(s_p) summed over the nodes in
//        registers.
//
// Width 850 is even but not a multiple of 8, so accesses are bf16 pairs (4 bytes).
#include "host_utils.h"
#include "sm100_ptx.cuh"

typedef __nv_bfloat16 bf16;

namespace det {

namespace {

constexpr int kMaxGate = 8;

struct GatePtrs {
  const bf16* ds[kMaxGate];    // bwd: gradient of each node state (nullable)
  const bf16* dsm[kMaxGate];   // bwd: gradient of each masked copy (nullable)
  bf16* s[kMaxGate];           // fwd: node states
  bf16* sm[kMaxGate];          // fwd: masked copies (nullable)
  int act[kMaxGate];
};

__device__ __forceinline__ float gate_sigmoid(float x) { return 1.0f / (1.0f + __expf(-x)); }

__device__ __forceinline__ float gate_act(float h, int act) {
  switch (act) {
    case 0: return tanhf(h);
    case 1: return fmaxf(h, 0.0f);
    case 2: return gate_sigmoid(h);
    default: return h;
  }
}

// derivative of the activation, from its input h and its output a
__device__ __forceinline__ float gate_act_grad(float h, float a, int act) {
  switch (act) {
    case 0: return 1.0f - a * a;
    case 1: return h > 0.0f ? 1.0f : 0.0f;
    case 2: return a * (1.0f - a);
    default: return 1.0f;
  }
}

__device__ __forceinline__ float2 ld2(const bf16* p) {
  const uint32_t u = *reinterpret_cast<const uint32_t*>(p);
  return make_float2(ptx::bf16_lo(u), ptx::bf16_hi(u));
}
__device__ __forceinline__ void st2(bf16* p, float a, float b) { *reinterpret_cast<uint32_t*>(p) = ptx::pack_bf16x2(a, b); }

__global__ void __launch_bounds__(256) highway_gate_fwd_kernel(const bf16* __restrict__ ch, const bf16* __restrict__ sp,
                                                               const bf16* __restrict__ mask, const GatePtrs p, int k, int B, int nh) {
  const int nh2 = nh >> 1;
  const long long total = static_cast<long long>(k) * B * nh2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % nh2) * 2;
    const long long r = i / nh2;
    const int b = static_cast<int>(r % B);
    const int j = static_cast<int>(r / B);
    const bf16* row = ch + (static_cast<long long>(b) * k + j) * 2 * nh;
    const float2 cc = ld2(row + c), hh = ld2(row + nh + c);
    const long long o = static_cast<long long>(b) * nh + c;
    const float2 s0 = ld2(sp + o);
    const int act = p.act[j];
    const float sx = s0.x + gate_sigmoid(cc.x) * (gate_act(hh.x, act) - s0.x);
    const float sy = s0.y + gate_sigmoid(cc.y) * (gate_act(hh.y, act) - s0.y);
    st2(p.s[j] + o, sx, sy);
    if (mask != nullptr) {
      const float2 m = ld2(mask + o);
      st2(p.sm[j] + o, sx * m.x, sy * m.y);
    }
  }
}

__global__ void __launch_bounds__(256) highway_gate_bwd_kernel(const bf16* __restrict__ ch, const bf16* __restrict__ sp,
                                                               const bf16* __restrict__ mask, const GatePtrs p,
                                                               bf16* __restrict__ dch, bf16* __restrict__ dsp, int k, int B, int nh) {
  const int nh2 = nh >> 1;
  const long long total = static_cast<long long>(B) * nh2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % nh2) * 2;
    const int b = static_cast<int>(i / nh2);
    const long long o = static_cast<long long>(b) * nh + c;
    const float2 s0 = ld2(sp + o);
    float2 m = make_float2(1.0f, 1.0f);
    if (mask != nullptr) m = ld2(mask + o);
    float ax = 0.0f, ay = 0.0f;      // d(s_p), summed over the nodes of the group
    for (int j = 0; j < k; ++j) {
      float dx = 0.0f, dy = 0.0f;
      if (p.ds[j] != nullptr) { const float2 d = ld2(p.ds[j] + o); dx = d.x; dy = d.y; }
      if (p.dsm[j] != nullptr) { const float2 d = ld2(p.dsm[j] + o); dx += d.x * m.x; dy += d.y * m.y; }
      const long long ro = (static_cast<long long>(b) * k + j) * 2 * nh;
      const float2 cc = ld2(ch + ro + c), hh = ld2(ch + ro + nh + c);
      const int act = p.act[j];
      const float gx = gate_sigmoid(cc.x), gy = gate_sigmoid(cc.y);
      const float fx = gate_act(hh.x, act), fy = gate_act(hh.y, act);
      ax += dx * (1.0f - gx);
      ay += dy * (1.0f - gy);
      st2(dch + ro + c, dx * (fx - s0.x) * gx * (1.0f - gx), dy * (fy - s0.y) * gy * (1.0f - gy));
      st2(dch + ro + nh + c, dx * gx * gate_act_grad(hh.x, fx, act), dy * gy * gate_act_grad(hh.y, fy, act));
    }
    st2(dsp + o, ax, ay);
  }
}

inline unsigned blocks_for(long long n) { return static_cast<unsigned>((n + 255) / 256); }

int fill_gate(GatePtrs& p, const int* acts, int k, int nh, int B) {
  DET_REQUIRE(k >= 1 && k <= kMaxGate, "a gate group holds 1 to 8 nodes");
  DET_REQUIRE(B > 0 && nh > 0 && nh % 2 == 0, "the gate kernels need an even width");
  for (int j = 0; j < kMaxGate; ++j) {
    p.ds[j] = p.dsm[j] = nullptr;
    p.s[j] = p.sm[j] = nullptr;
    p.act[j] = 3;
  }
  for (int j = 0; j < k; ++j) {
    DET_REQUIRE(acts[j] >= 0 && acts[j] <= 3, "activation must be 0 (tanh), 1 (relu), 2 (sigmoid) or 3 (identity)");
    p.act[j] = acts[j];
  }
  return 0;
}

}  // namespace

}  // namespace det

// ch [B, k·2·nh] (node j's [c | h] at columns j·2·nh), sp [B, nh], mask [B, nh] or null; s_out / sm_out: k pointers each.
extern "C" int det_highway_gate_fwd(const void* ch, const void* sp, const void* mask, void* const* s_out, void* const* sm_out,
                                    const int* acts, int k, int B, int nh, cudaStream_t stream) {
  det::GatePtrs p;
  if (int rc = det::fill_gate(p, acts, k, nh, B)) return rc;
  DET_REQUIRE(mask == nullptr || sm_out != nullptr, "a mask needs outputs for the masked copies");
  for (int j = 0; j < k; ++j) {
    p.s[j] = static_cast<bf16*>(s_out[j]);
    p.sm[j] = mask != nullptr ? static_cast<bf16*>(sm_out[j]) : nullptr;
  }
  const long long total = static_cast<long long>(k) * B * (nh / 2);
  det::highway_gate_fwd_kernel<<<det::blocks_for(total), 256, 0, stream>>>((const bf16*)ch, (const bf16*)sp, (const bf16*)mask, p, k, B, nh);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ds / dsm: k pointers each, entries may be null (no gradient arrived through that output). Writes dch [B, k·2·nh], dsp [B, nh].
extern "C" int det_highway_gate_bwd(const void* ch, const void* sp, const void* mask, const void* const* ds, const void* const* dsm,
                                    void* dch, void* dsp, const int* acts, int k, int B, int nh, cudaStream_t stream) {
  det::GatePtrs p;
  if (int rc = det::fill_gate(p, acts, k, nh, B)) return rc;
  for (int j = 0; j < k; ++j) {
    p.ds[j] = static_cast<const bf16*>(ds[j]);
    p.dsm[j] = (mask != nullptr && dsm != nullptr) ? static_cast<const bf16*>(dsm[j]) : nullptr;
  }
  const long long total = static_cast<long long>(B) * (nh / 2);
  det::highway_gate_bwd_kernel<<<det::blocks_for(total), 256, 0, stream>>>((const bf16*)ch, (const bf16*)sp, (const bf16*)mask, p,
                                                                          (bf16*)dch, (bf16*)dsp, k, B, nh);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}
