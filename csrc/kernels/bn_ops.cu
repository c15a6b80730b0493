// bn_ops.cu — training-mode BatchNorm over the rows of an NHWC bf16 activation x[M, C] (C % 8 == 0), written for
// HBM bandwidth: every kernel uses the same "column strip" mapping — a thread owns 8 consecutive channels (one 16-byte
// vector) and walks down the rows with UNROLL independent 16-byte loads in flight, so per-channel constants (scale,
// shift, means) live in registers for the whole kernel instead of being re-fetched per element.
//
//   forward : bn_stats (column sums of x and x^2 -> mean, rstd, running stats)
//             bn_apply (y = act((x - mean) * rstd * gamma + beta + residual))
//   backward: bn_bwd_reduce (sum dyv, sum dyv*xhat with dyv = dy * act'(y))
//             bn_bwd_apply  (dx, and the residual-branch gradient dyv, in one pass), bn_param_grad (dgamma, dbeta +=)
//
// The reference reaches BatchNorm through torch.nn.BatchNorm2d → cuDNN in user model code (e.g. the torchvision
// backbones of examples/computer_vision/fasterrcnn_coco_pytorch); statistics are per process there too.
#include "host_utils.h"
#include "sm100_ptx.cuh"

namespace det {

namespace {

constexpr int BN_THREADS = 256;
constexpr int BN_UNROLL = 4;
constexpr int BN_CHUNK = 2048;  // channels per block column (256 threads x 8 channels)

__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  f[0] = ptx::bf16_lo(u.x); f[1] = ptx::bf16_hi(u.x); f[2] = ptx::bf16_lo(u.y); f[3] = ptx::bf16_hi(u.y);
  f[4] = ptx::bf16_lo(u.z); f[5] = ptx::bf16_hi(u.z); f[6] = ptx::bf16_lo(u.w); f[7] = ptx::bf16_hi(u.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 u;
  u.x = ptx::pack_bf16x2(f[0], f[1]); u.y = ptx::pack_bf16x2(f[2], f[3]);
  u.z = ptx::pack_bf16x2(f[4], f[5]); u.w = ptx::pack_bf16x2(f[6], f[7]);
  return u;
}
__device__ __forceinline__ uint4 ld_stream(const __nv_bfloat16* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

struct Strip {
  int vpr;       // vectors per row handled by this block column
  int rpi;       // rows covered by the block per pass
  int cv, ro;    // this thread's vector column and row offset inside a pass
  int c0;        // first channel of this thread
  bool active;
  long long row0, step;
};
__device__ __forceinline__ Strip make_strip(int C) {
  Strip s;
  const int cbase = blockIdx.y * BN_CHUNK;
  const int cc = min(BN_CHUNK, C - cbase);
  s.vpr = cc / 8;
  s.rpi = BN_THREADS / s.vpr;
  s.cv = threadIdx.x % s.vpr;
  s.ro = threadIdx.x / s.vpr;
  s.c0 = cbase + s.cv * 8;
  s.active = s.ro < s.rpi;
  s.row0 = static_cast<long long>(blockIdx.x) * s.rpi + s.ro;
  s.step = static_cast<long long>(gridDim.x) * s.rpi;
  return s;
}

__device__ __forceinline__ float act_mask(float a, int act, float slope) {
  return a > 0.0f ? 1.0f : (act == 2 ? slope : 0.0f);
}

// Fold per-thread column partials s[8], q[8] into global sums[0:C] / sums[C:2C]: lanes of a warp that share a
// column combine by shuffle (power-of-two strips), warps combine through shared-memory atomics, one coalesced global
// atomic per channel per block finishes.
__device__ __forceinline__ void fold_columns(const Strip& st, float* s, float* q, float* __restrict__ sums, int C,
                                             float (*sm)[BN_CHUNK]) {
  const int cc = st.vpr * 8;
  for (int i = threadIdx.x; i < cc; i += BN_THREADS) { sm[0][i] = 0.0f; sm[1][i] = 0.0f; }
  __syncthreads();
  const bool pow2 = (st.vpr & (st.vpr - 1)) == 0;
  bool writer = st.active;
  if (pow2 && st.vpr < 32) {
    for (int off = 16; off >= st.vpr; off >>= 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s[j] += __shfl_xor_sync(0xffffffffu, s[j], off);
        q[j] += __shfl_xor_sync(0xffffffffu, q[j], off);
      }
    }
    writer = (threadIdx.x & 31) < st.vpr;   // inactive lanes contributed zeros
  }
  if (writer) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(&sm[0][st.cv * 8 + j], s[j]);
      atomicAdd(&sm[1][st.cv * 8 + j], q[j]);
    }
  }
  __syncthreads();
  const int cbase = blockIdx.y * BN_CHUNK;
  for (int i = threadIdx.x; i < cc; i += BN_THREADS) {
    atomicAdd(sums + cbase + i, sm[0][i]);
    atomicAdd(sums + C + cbase + i, sm[1][i]);
  }
}

}  // namespace

// sums[0:C] += sum_rows x, sums[C:2C] += sum_rows x^2
__global__ void __launch_bounds__(BN_THREADS)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ sums, long long M, int C) {
  __shared__ float sm[2][BN_CHUNK];
  const Strip st = make_strip(C);
  float s[8] = {0, 0, 0, 0, 0, 0, 0, 0}, q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (st.active) {
    const __nv_bfloat16* base = x + st.c0;
    for (long long row = st.row0; row < M; row += BN_UNROLL * st.step) {
      uint4 v[BN_UNROLL];
#pragma unroll
      for (int u = 0; u < BN_UNROLL; ++u) {
        const long long r = row + u * st.step;
        v[u] = r < M ? ld_stream(base + r * C) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < BN_UNROLL; ++u) {
        float f[8];
        unpack8(v[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) { s[j] += f[j]; q[j] = fmaf(f[j], f[j], q[j]); }
      }
    }
  }
  fold_columns(st, s, q, sums, C, sm);
}

// mean / rstd from the sums; running statistics updated the way torch.nn.BatchNorm2d does (unbiased variance).
__global__ void bn_finalize_kernel(const float* __restrict__ sums, float* __restrict__ mean, float* __restrict__ rstd,
                                   float* running_mean, float* running_var, long long M, int C, float eps, float momentum) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float inv = 1.0f / static_cast<float>(M);
  const float mu = sums[c] * inv;
  const float var = fmaxf(sums[C + c] * inv - mu * mu, 0.0f);
  mean[c] = mu;
  rstd[c] = rsqrtf(var + eps);
  if (running_mean != nullptr) {
    const float unb = M > 1 ? var * static_cast<float>(M) / static_cast<float>(M - 1) : var;
    running_mean[c] = (1.0f - momentum) * running_mean[c] + momentum * mu;
    running_var[c] = (1.0f - momentum) * running_var[c] + momentum * unb;
  }
}

// y = act((x - mean) * rstd * gamma + beta + resid)
template <bool RESID>
__global__ void __launch_bounds__(BN_THREADS)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ resid, __nv_bfloat16* __restrict__ y,
                const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
                const float* __restrict__ beta, long long M, int C, int act, float slope) {
  const Strip st = make_strip(C);
  if (!st.active) return;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = __ldg(rstd + st.c0 + j) * (gamma ? __ldg(gamma + st.c0 + j) : 1.0f);
    sh[j] = (beta ? __ldg(beta + st.c0 + j) : 0.0f) - __ldg(mean + st.c0 + j) * sc[j];
  }
  for (long long row = st.row0; row < M; row += BN_UNROLL * st.step) {
    uint4 v[BN_UNROLL], rv[BN_UNROLL];
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u) {
      const long long r = row + u * st.step;
      if (r < M) {
        v[u] = ld_stream(x + r * C + st.c0);
        if (RESID) rv[u] = ld_stream(resid + r * C + st.c0);
      }
    }
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u) {
      const long long r = row + u * st.step;
      if (r >= M) break;
      float f[8], rr[8];
      unpack8(v[u], f);
      if (RESID) unpack8(rv[u], rr);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float t = fmaf(f[j], sc[j], sh[j]);
        if (RESID) t += rr[j];
        if (act == 1) t = fmaxf(t, 0.0f);
        else if (act == 2) t = t > 0.0f ? t : t * slope;
        f[j] = t;
      }
      *reinterpret_cast<uint4*>(y + r * C + st.c0) = pack8(f);
    }
  }
}

// sums[0:C] += sum dyv, sums[C:2C] += sum dyv * xhat   with dyv = dy * act'(a)
template <bool ACT>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const __nv_bfloat16* __restrict__ a, const float* __restrict__ mean, const float* __restrict__ rstd,
                     float* __restrict__ sums, long long M, int C, int act, float slope) {
  __shared__ float sm[2][BN_CHUNK];
  const Strip st = make_strip(C);
  float s[8] = {0, 0, 0, 0, 0, 0, 0, 0}, q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (st.active) {
    float mu[8], rs[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { mu[j] = __ldg(mean + st.c0 + j); rs[j] = __ldg(rstd + st.c0 + j); }
    for (long long row = st.row0; row < M; row += BN_UNROLL * st.step) {
      uint4 dv[BN_UNROLL], xv[BN_UNROLL], av[BN_UNROLL];
#pragma unroll
      for (int u = 0; u < BN_UNROLL; ++u) {
        const long long r = row + u * st.step;
        if (r < M) {
          const long long off = r * C + st.c0;
          dv[u] = ld_stream(dy + off);
          xv[u] = ld_stream(x + off);
          if (ACT) av[u] = ld_stream(a + off);
        } else {
          dv[u] = make_uint4(0, 0, 0, 0);
          xv[u] = make_uint4(0, 0, 0, 0);
          if (ACT) av[u] = make_uint4(0, 0, 0, 0);
        }
      }
#pragma unroll
      for (int u = 0; u < BN_UNROLL; ++u) {
        float d[8], xf[8], af[8];
        unpack8(dv[u], d);
        unpack8(xv[u], xf);
        if (ACT) unpack8(av[u], af);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float g = ACT ? d[j] * act_mask(af[j], act, slope) : d[j];
          s[j] += g;
          q[j] = fmaf(g, (xf[j] - mu[j]) * rs[j], q[j]);
        }
      }
    }
  }
  fold_columns(st, s, q, sums, C, sm);
}

// dx = gamma * rstd * (dyv - mean(dyv) - xhat * mean(dyv * xhat)); dresid = dyv (optional)
template <bool ACT>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                    const __nv_bfloat16* __restrict__ a, const float* __restrict__ mean, const float* __restrict__ rstd,
                    const float* __restrict__ gamma, const float* __restrict__ sums, __nv_bfloat16* __restrict__ dx,
                    __nv_bfloat16* __restrict__ dresid, long long M, int C, int act, float slope) {
  const Strip st = make_strip(C);
  if (!st.active) return;
  const float inv = 1.0f / static_cast<float>(M);
  // dx = k1 * dyv + k2 * x + k3 with xhat = (x - mu) * rs:  k1 = g*rs, k2 = -g*rs*rs*m2, k3 = -g*rs*(m1 - mu*rs*m2)
  float k1[8], k2[8], k3[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float rs = __ldg(rstd + st.c0 + j), mu = __ldg(mean + st.c0 + j);
    const float g = (gamma ? __ldg(gamma + st.c0 + j) : 1.0f) * rs;
    const float m1 = __ldg(sums + st.c0 + j) * inv, m2 = __ldg(sums + C + st.c0 + j) * inv;
    k1[j] = g;
    k2[j] = -g * rs * m2;
    k3[j] = -g * (m1 - mu * rs * m2);
  }
  for (long long row = st.row0; row < M; row += BN_UNROLL * st.step) {
    uint4 dv[BN_UNROLL], xv[BN_UNROLL], av[BN_UNROLL];
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u) {
      const long long r = row + u * st.step;
      if (r < M) {
        const long long off = r * C + st.c0;
        dv[u] = ld_stream(dy + off);
        xv[u] = ld_stream(x + off);
        if (ACT) av[u] = ld_stream(a + off);
      }
    }
#pragma unroll
    for (int u = 0; u < BN_UNROLL; ++u) {
      const long long r = row + u * st.step;
      if (r >= M) break;
      float d[8], xf[8], af[8], o[8];
      unpack8(dv[u], d);
      unpack8(xv[u], xf);
      if (ACT) unpack8(av[u], af);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (ACT) d[j] *= act_mask(af[j], act, slope);
        o[j] = fmaf(k1[j], d[j], fmaf(k2[j], xf[j], k3[j]));
      }
      const long long off = r * C + st.c0;
      *reinterpret_cast<uint4*>(dx + off) = pack8(o);
      if (dresid != nullptr) *reinterpret_cast<uint4*>(dresid + off) = pack8(d);
    }
  }
}

// dgamma += sums[C:2C], dbeta += sums[0:C]
__global__ void bn_param_grad_kernel(const float* __restrict__ sums, float* dgamma, float* dbeta, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  if (dbeta != nullptr) dbeta[c] += sums[c];
  if (dgamma != nullptr) dgamma[c] += sums[C + c];
}

// `reduce` kernels end with one global atomic per channel per block onto the SAME 2*C addresses, so their grid is kept
// at two blocks per SM (ncu: with eight per SM a 25 MB statistics pass spent most of its 25 us in those atomics);
// the elementwise kernels take the full eight.
static dim3 strip_grid(long long M, int C, bool reduce = false) {
  const int vpr = (C < BN_CHUNK ? C : BN_CHUNK) / 8;
  const int rpi = BN_THREADS / vpr;
  long long bx = (M + static_cast<long long>(rpi) * BN_UNROLL - 1) / (static_cast<long long>(rpi) * BN_UNROLL);
  const long long cap = static_cast<long long>(sm_count()) * (reduce ? 2 : 8);
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  return dim3(static_cast<unsigned>(bx), static_cast<unsigned>((C + BN_CHUNK - 1) / BN_CHUNK), 1);
}

}  // namespace det

typedef __nv_bfloat16 bf16;

// Training-mode statistics: sums_scratch [2*C] fp32 is cleared here; mean/rstd [C] are written; running stats updated
// in place when given.
extern "C" int det_bn_stats(const void* x, float* sums_scratch, float* mean, float* rstd, float* running_mean,
                            float* running_var, long long M, int C, float eps, float momentum, cudaStream_t stream) {
  DET_REQUIRE(M > 0 && C > 0 && C % 8 == 0, "BatchNorm needs C % 8 == 0");
  DET_REQUIRE(C <= det::BN_CHUNK || C % det::BN_CHUNK == 0, "BatchNorm: C above 2048 must be a multiple of 2048");
  DET_CHECK_CUDA(cudaMemsetAsync(sums_scratch, 0, sizeof(float) * 2 * C, stream));
  det::bn_stats_kernel<<<det::strip_grid(M, C, true), det::BN_THREADS, 0, stream>>>((const bf16*)x, sums_scratch, M, C);
  det::bn_finalize_kernel<<<(C + 127) / 128, 128, 0, stream>>>(sums_scratch, mean, rstd, running_mean, running_var, M, C, eps, momentum);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_bn_apply(const void* x, const void* resid, void* y, const float* mean, const float* rstd,
                            const float* gamma, const float* beta, long long M, int C, int act, float slope,
                            cudaStream_t stream) {
  DET_REQUIRE(M > 0 && C > 0 && C % 8 == 0, "BatchNorm needs C % 8 == 0");
  DET_REQUIRE(C <= det::BN_CHUNK || C % det::BN_CHUNK == 0, "BatchNorm: C above 2048 must be a multiple of 2048");
  const dim3 grid = det::strip_grid(M, C);
  if (resid != nullptr)
    det::bn_apply_kernel<true><<<grid, det::BN_THREADS, 0, stream>>>((const bf16*)x, (const bf16*)resid, (bf16*)y, mean, rstd, gamma, beta, M, C, act, slope);
  else
    det::bn_apply_kernel<false><<<grid, det::BN_THREADS, 0, stream>>>((const bf16*)x, nullptr, (bf16*)y, mean, rstd, gamma, beta, M, C, act, slope);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Backward of y = act(bn(x) + resid). `a` is the saved forward output (needed when act != 0). dgamma/dbeta are
// accumulated (+=). dresid (optional) receives the gradient of the residual branch.
extern "C" int det_bn_bwd(const void* dy, const void* x, const void* a, const float* mean, const float* rstd,
                          const float* gamma, float* sums_scratch, void* dx, void* dresid, float* dgamma, float* dbeta,
                          long long M, int C, int act, float slope, cudaStream_t stream) {
  DET_REQUIRE(M > 0 && C > 0 && C % 8 == 0, "BatchNorm needs C % 8 == 0");
  DET_REQUIRE(C <= det::BN_CHUNK || C % det::BN_CHUNK == 0, "BatchNorm: C above 2048 must be a multiple of 2048");
  DET_REQUIRE(act == 0 || a != nullptr, "BatchNorm backward through an activation needs the saved output");
  DET_CHECK_CUDA(cudaMemsetAsync(sums_scratch, 0, sizeof(float) * 2 * C, stream));
  const dim3 grid = det::strip_grid(M, C);
  const dim3 rgrid = det::strip_grid(M, C, true);
  if (act != 0) {
    det::bn_bwd_reduce_kernel<true><<<rgrid, det::BN_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)x, (const bf16*)a, mean, rstd, sums_scratch, M, C, act, slope);
    det::bn_bwd_apply_kernel<true><<<grid, det::BN_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)x, (const bf16*)a, mean, rstd, gamma, sums_scratch, (bf16*)dx, (bf16*)dresid, M, C, act, slope);
  } else {
    det::bn_bwd_reduce_kernel<false><<<rgrid, det::BN_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)x, nullptr, mean, rstd, sums_scratch, M, C, act, slope);
    det::bn_bwd_apply_kernel<false><<<grid, det::BN_THREADS, 0, stream>>>((const bf16*)dy, (const bf16*)x, nullptr, mean, rstd, gamma, sums_scratch, (bf16*)dx, (bf16*)dresid, M, C, act, slope);
  }
  if (dgamma != nullptr || dbeta != nullptr)
    det::bn_param_grad_kernel<<<(C + 127) / 128, 128, 0, stream>>>(sums_scratch, dgamma, dbeta, C);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// mean / rstd (and the running statistics) from column sums produced elsewhere — by det_bn_stats' reduction or by the
// fused epilogues det_gemm_bf16_stats / det_conv2d_fprop_stats. sums = [sum | sum of squares], 2*C floats.
extern "C" int det_bn_finalize(const float* sums, float* mean, float* rstd, float* running_mean, float* running_var, long long M,
                               int C, float eps, float momentum, cudaStream_t stream) {
  DET_REQUIRE(M > 0 && C > 0, "empty BatchNorm");
  det::bn_finalize_kernel<<<(C + 127) / 128, 128, 0, stream>>>(sums, mean, rstd, running_mean, running_var, M, C, eps, momentum);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}
