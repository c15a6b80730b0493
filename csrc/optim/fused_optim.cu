// fused_optim.cu — optimizer updates applied directly on flat gradient buckets: one launch per
// bucket instead of PyTorch-1.4-era per-parameter loops of 5–10 elementwise kernels
// (reference call site: harness/determined/pytorch/_pytorch_context.py:538-542 `optimizer.step()`),
// with the neighbouring steps the reference performs as separate passes folded in:
//   * gradient unscale / averaging (÷ loss_scale, ÷ world, ÷ aggregation_frequency — reference
//     `_average_gradients`, _pytorch_context.py:470-477) through a device-resident scale factor,
//   * global-norm clipping coefficient (reference `clip_grads`, :535-536) folded into that factor,
//   * skip-on-overflow (apex.amp dynamic loss scaling, :444-462) through a device-resident flag,
//   * `zero_grad()` (:545) by writing zeros back to the bucket while it is still in registers,
//   * the bf16 model copy ("shadow") that the tensor-core kernels read.
// Every hyper-parameter that changes between steps lives in device memory so a captured CUDA graph
// replays with new learning rates / bias corrections without re-capture.
#include <cuda_bf16.h>

#include <algorithm>

#include "host_utils.h"

namespace det {

// hyper[] layout (device fp32):
enum { HP_LR = 0, HP_B1 = 1, HP_B2 = 2, HP_EPS = 3, HP_WD = 4, HP_BC1 = 5, HP_BC2 = 6, HP_DAMP = 7, HP_FLAG = 8 };

struct OptArgs {
  float* p;
  float* g;
  float* s1;
  float* s2;
  __nv_bfloat16* shadow;
  long long n;
  const float* hyper;
  const float* grad_scale;
  const int* found_inf;
  int zero_grad;
  int variant;
};

__device__ __forceinline__ void store_shadow4(__nv_bfloat16* sh, long long i, const float4& p) {
  if (sh == nullptr) return;
  __nv_bfloat162 a = __floats2bfloat162_rn(p.x, p.y), b = __floats2bfloat162_rn(p.z, p.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(sh + i) = u;
}

enum { OPT_ADAM = 0, OPT_ADAMW = 1, OPT_SGD = 2, OPT_RMSPROP = 3, OPT_ADADELTA = 4 };

template <int OPT>
__device__ __forceinline__ void update1(float& p, float g, float& s1, float& s2, const float* hp) {
  const float lr = hp[HP_LR], wd = hp[HP_WD];
  if (OPT == OPT_ADAM || OPT == OPT_ADAMW) {
    const float b1 = hp[HP_B1], b2 = hp[HP_B2];
    if (OPT == OPT_ADAM) g += wd * p; else p -= lr * wd * p;
    s1 = b1 * s1 + (1.0f - b1) * g;
    s2 = b2 * s2 + (1.0f - b2) * g * g;
    const float denom = sqrtf(s2) * rsqrtf(hp[HP_BC2]) + hp[HP_EPS];
    p -= (lr / hp[HP_BC1]) * (s1 / denom);
  } else if (OPT == OPT_SGD) {
    const float mom = hp[HP_B1];
    g += wd * p;
    if (mom != 0.0f) {
      const bool first = hp[HP_FLAG] != 0.0f;  // first step: buf = g (torch semantics)
      s1 = first ? g : (mom * s1 + (1.0f - hp[HP_DAMP]) * g);
      g = (hp[HP_B2] != 0.0f) ? (g + mom * s1) : s1;  // HP_B2 doubles as the nesterov flag for SGD
    }
    p -= lr * g;
  } else if (OPT == OPT_RMSPROP) {
    const float alpha = hp[HP_B1], mom = hp[HP_B2];
    g += wd * p;
    s1 = alpha * s1 + (1.0f - alpha) * g * g;
    const float avg = sqrtf(s1) + hp[HP_EPS];
    if (mom > 0.0f) { s2 = mom * s2 + g / avg; p -= lr * s2; }
    else p -= lr * g / avg;
  } else {  // Adadelta
    const float rho = hp[HP_B1], eps = hp[HP_EPS];
    g += wd * p;
    s1 = rho * s1 + (1.0f - rho) * g * g;
    const float delta = sqrtf(s2 + eps) * rsqrtf(s1 + eps) * g;
    s2 = rho * s2 + (1.0f - rho) * delta * delta;
    p -= lr * delta;
  }
}

template <int OPT>
__global__ void __launch_bounds__(256)
fused_opt_kernel(const OptArgs a) {
  const bool skip = (a.found_inf != nullptr) && (*a.found_inf != 0);
  const float gs = a.grad_scale ? *a.grad_scale : 1.0f;
  float hp[12];  // registers: every thread reads the same 12 scalars once (L1-broadcast), none in the loop
#pragma unroll
  for (int i = 0; i < 12; ++i) hp[i] = __ldg(a.hyper + i);
  const long long n4 = a.n >> 2;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
    // All four loads are issued before ANY store: a store to the line `g` was just loaded from (the zero-grad write)
    // queues behind that outstanding miss in the LSU and would serialise the p / s1 / s2 loads behind it (measured:
    // 1.8 ms vs 0.5 ms for 109M parameters).
    float4 g = reinterpret_cast<const float4*>(a.g)[i];
    float4 p = reinterpret_cast<float4*>(a.p)[i];
    float4 s1 = reinterpret_cast<float4*>(a.s1)[i];
    float4 s2 = a.s2 ? reinterpret_cast<float4*>(a.s2)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    if (!skip) {
      update1<OPT>(p.x, g.x * gs, s1.x, s2.x, hp);
      update1<OPT>(p.y, g.y * gs, s1.y, s2.y, hp);
      update1<OPT>(p.z, g.z * gs, s1.z, s2.z, hp);
      update1<OPT>(p.w, g.w * gs, s1.w, s2.w, hp);
      reinterpret_cast<float4*>(a.p)[i] = p;
      reinterpret_cast<float4*>(a.s1)[i] = s1;
      if (a.s2) reinterpret_cast<float4*>(a.s2)[i] = s2;
      store_shadow4(a.shadow, i * 4, p);
    }
    if (a.zero_grad) reinterpret_cast<float4*>(a.g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  // tail (n not a multiple of 4)
  const long long tail0 = n4 << 2;
  const long long t = tail0 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t < a.n) {
    float g = a.g[t];
    if (a.zero_grad) a.g[t] = 0.0f;
    if (!skip) {
      float p = a.p[t], s1 = a.s1[t], s2 = a.s2 ? a.s2[t] : 0.0f;
      update1<OPT>(p, g * gs, s1, s2, hp);
      a.p[t] = p; a.s1[t] = s1;
      if (a.s2) a.s2[t] = s2;
      if (a.shadow) a.shadow[t] = __float2bfloat16_rn(p);
    }
  }
}

// sum of squares + non-finite detection over a bucket (one pass), pre-scaled by *grad_scale.
__global__ void __launch_bounds__(256)
sumsq_finite_kernel(const float* __restrict__ g, long long n, const float* grad_scale,
                    float* __restrict__ sumsq, int* __restrict__ found_inf) {
  const float gs = grad_scale ? *grad_scale : 1.0f;
  float acc = 0.0f;
  bool bad = false;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  // 16-byte loads only from a 16-byte aligned base (a view that starts mid-allocation takes the scalar loop below)
  const long long n4 = (reinterpret_cast<uintptr_t>(g) & 15) == 0 ? (n >> 2) : 0;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(g) + i);
    const float a = v.x * gs, b = v.y * gs, c = v.z * gs, d = v.w * gs;
    acc += a * a + b * b + c * c + d * d;
    bad |= !(isfinite(a) && isfinite(b) && isfinite(c) && isfinite(d));
  }
  for (long long i = (n4 << 2) + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float a = g[i] * gs;
    acc += a * a;
    bad |= !isfinite(a);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ float red[8];
  __shared__ int sbad;
  if (threadIdx.x == 0) sbad = 0;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  if (bad) sbad = 1;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.0f;
    for (int w = 0; w < 8; ++w) s += red[w];
    atomicAdd(sumsq, s);
    if (sbad && found_inf) atomicExch(found_inf, 1);
  }
}

// coef = min(1, max_norm / (sqrt(sumsq) + 1e-6)); out[0] = base_scale * coef; out[1] = grad norm.
__global__ void clip_coef_kernel(const float* sumsq, float max_norm, const float* base_scale, float* out) {
  const float base = base_scale ? *base_scale : 1.0f;
  const float norm = sqrtf(*sumsq);
  float coef = 1.0f;
  if (max_norm > 0.0f) coef = fminf(1.0f, max_norm / (norm + 1e-6f));
  out[0] = base * coef;
  out[1] = norm;
}

// Dynamic loss-scale state machine (apex/torch GradScaler semantics), entirely on device.
// state[0] = loss scale, state[1] = growth tracker, state[2] = 1/(scale*divisor) written for next step.
__global__ void loss_scale_update_kernel(float* state, int* found_inf, float growth, float backoff,
                                         int growth_interval, float divisor) {
  float scale = state[0];
  float tracker = state[1];
  if (*found_inf) {
    scale = fmaxf(scale * backoff, 1.0f);
    tracker = 0.0f;
  } else {
    tracker += 1.0f;
    if (tracker >= static_cast<float>(growth_interval)) { scale *= growth; tracker = 0.0f; }
  }
  state[0] = scale;
  state[1] = tracker;
  state[2] = 1.0f / (scale * divisor);
  *found_inf = 0;
}

}  // namespace det

extern "C" int det_fused_optimizer_step(int variant, float* param, float* grad, float* state1,
                                        float* state2, void* shadow_bf16, long long n,
                                        const float* hyper_dev, const float* grad_scale_dev,
                                        const int* found_inf_dev, int zero_grad, cudaStream_t stream) {
  using namespace det;
  if (n <= 0) return 0;
  DET_REQUIRE(param && grad && state1 && hyper_dev, "optimizer: null pointer");
  DET_REQUIRE(((reinterpret_cast<uintptr_t>(param) | reinterpret_cast<uintptr_t>(grad) |
                reinterpret_cast<uintptr_t>(state1) | reinterpret_cast<uintptr_t>(state2)) & 15) == 0,
              "optimizer: buckets must be 16-byte aligned");
  DET_REQUIRE((reinterpret_cast<uintptr_t>(shadow_bf16) & 7) == 0, "optimizer: shadow must be 8-byte aligned");
  OptArgs a{param, grad, state1, state2, reinterpret_cast<__nv_bfloat16*>(shadow_bf16), n, hyper_dev,
            grad_scale_dev, found_inf_dev, zero_grad, variant};
  const long long n4 = (n + 3) / 4;
  const int blocks = static_cast<int>(std::min<long long>((n4 + 255) / 256, static_cast<long long>(sm_count()) * 8));
  switch (variant) {
    case OPT_ADAM: fused_opt_kernel<OPT_ADAM><<<blocks, 256, 0, stream>>>(a); break;
    case OPT_ADAMW: fused_opt_kernel<OPT_ADAMW><<<blocks, 256, 0, stream>>>(a); break;
    case OPT_SGD: fused_opt_kernel<OPT_SGD><<<blocks, 256, 0, stream>>>(a); break;
    case OPT_RMSPROP: fused_opt_kernel<OPT_RMSPROP><<<blocks, 256, 0, stream>>>(a); break;
    case OPT_ADADELTA:
      DET_REQUIRE(state2 != nullptr, "adadelta needs two state buffers");
      fused_opt_kernel<OPT_ADADELTA><<<blocks, 256, 0, stream>>>(a);
      break;
    default: DET_REQUIRE(false, "unknown optimizer variant");
  }
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_grad_sumsq_finite(const float* grad, long long n, const float* grad_scale_dev,
                                     float* sumsq_dev, int* found_inf_dev, cudaStream_t stream) {
  using namespace det;
  if (n <= 0) return 0;
  const int blocks = static_cast<int>(std::min<long long>((n / 4 + 255) / 256 + 1, static_cast<long long>(sm_count()) * 8));
  sumsq_finite_kernel<<<blocks, 256, 0, stream>>>(grad, n, grad_scale_dev, sumsq_dev, found_inf_dev);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_clip_coef(const float* sumsq_dev, float max_norm, const float* base_scale_dev,
                             float* out2_dev, cudaStream_t stream) {
  det::clip_coef_kernel<<<1, 1, 0, stream>>>(sumsq_dev, max_norm, base_scale_dev, out2_dev);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_loss_scale_update(float* state3_dev, int* found_inf_dev, float growth, float backoff,
                                     int growth_interval, float divisor, cudaStream_t stream) {
  det::loss_scale_update_kernel<<<1, 1, 0, stream>>>(state3_dev, found_inf_dev, growth, backoff, growth_interval, divisor);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}
