// norm_ops.cu — memory-bound companions of the tensor-core kernels: LayerNorm forward/backward
// (with the dropout masks of the neighbouring GEMM epilogues fused in), column sums for bias
// gradients, embedding gather / scatter-add, softmax cross-entropy (fwd+bwd in one pass), a
// strided CUDA-core GEMM for tiny heads (pooler / classifier), and small elementwise helpers.
//
// One warp owns one row and keeps it in registers, so each activation is read exactly once per
// pass; per-column reductions (dgamma/dbeta/bias grads) accumulate in registers across the rows a
// warp visits and hit global memory once per block through fp32 atomics into the gradient bucket.
//
// Replaces torch.nn.LayerNorm / Embedding / CrossEntropyLoss / small nn.Linear kernels used by
// the reference's examples (reference: examples/nlp/bert_glue_pytorch/model_def.py:60-68).
#include <algorithm>

#include "host_utils.h"
#include "sm100_ptx.cuh"

namespace det {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
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

struct DropCfg {
  uint32_t threshold;  // 0 = off
  float scale;
  uint32_t seed;
  const uint32_t* seed_dev;
};
__device__ __forceinline__ uint32_t eff_seed(const DropCfg& d) {
  return dropout_effective_seed(d.seed, d.seed_dev);
}
static DropCfg make_drop(float p, unsigned seed, const unsigned* seed_dev) {
  DropCfg d;
  if (p > 0.0f) {
    d.threshold = static_cast<uint32_t>(static_cast<double>(p) * 4294967296.0);
    d.scale = 1.0f / (1.0f - p);
  } else {
    d.threshold = 0; d.scale = 1.0f;
  }
  d.seed = seed;
  d.seed_dev = seed_dev;
  return d;
}

// ------------------------------------------------------------------------------------------------
// LayerNorm forward: y = dropout_out( (x - mean) * rstd * gamma + beta ).  H = CH * 256.
// ------------------------------------------------------------------------------------------------
template <int CH>
__global__ void __launch_bounds__(128)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
                     const float* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                     float* __restrict__ mean, float* __restrict__ rstd, int T, float eps,
                     DropCfg drop) {
  constexpr int H = CH * 256;
  const int lane = threadIdx.x & 31;
  const int warps_total = gridDim.x * 4;
  const uint32_t dseed = eff_seed(drop);
  for (int row = blockIdx.x * 4 + (threadIdx.x >> 5); row < T; row += warps_total) {
    float v[CH * 8];
    const uint4* xr = reinterpret_cast<const uint4*>(x + static_cast<long long>(row) * H);
#pragma unroll
    for (int c = 0; c < CH; ++c) unpack8(__ldg(xr + c * 32 + lane), v + c * 8);
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < CH * 8; ++i) s += v[i];
    const float mu = warp_sum(s) * (1.0f / H);
    float ss = 0.0f;
#pragma unroll
    for (int i = 0; i < CH * 8; ++i) { const float d = v[i] - mu; ss += d * d; }
    const float rs = rsqrtf(warp_sum(ss) * (1.0f / H) + eps);
    if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
    uint4* yr = reinterpret_cast<uint4*>(y + static_cast<long long>(row) * H);
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      const int col = c * 256 + lane * 8;
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + col));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + col + 4));
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + col));
      const float4 b1 = __ldg(reinterpret_cast<const float4*>(beta + col + 4));
      const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        float t = (v[c * 8 + e] - mu) * rs * g[e] + b[e];
        if (drop.threshold != 0u)
          t = dropout_keep(dseed, static_cast<uint32_t>(row) * H + col + e, drop.threshold) ? t * drop.scale : 0.0f;
        o[e] = t;
      }
      yr[c * 32 + lane] = pack8(o);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// LayerNorm backward.
//   dy_eff = drop_in(dy)                      (mask of a dropout applied to the LN *output*)
//   dx     = rstd * (dy_eff*gamma - mean(dy_eff*gamma) - xhat * mean(dy_eff*gamma*xhat))
//   dx_masked = drop_out(dx)                  (mask of a dropout applied to the producer of x,
//                                              i.e. the GEMM epilogue in front of the residual add)
//   dgamma += sum_rows dy_eff * xhat ; dbeta += sum_rows dy_eff      (fp32 atomics)
// ------------------------------------------------------------------------------------------------
// Twelve rows in flight per block: every block ends with one atomic per column onto the same 2*H addresses, so fewer,
// fatter blocks (one per SM) keep that serialised tail short (the BatchNorm reductions taught the same lesson).
constexpr int LN_BWD_WARPS = 12;   // 160 registers per thread: 12 warps fill the register file with ONE block per SM

template <int CH>
__global__ void __launch_bounds__(LN_BWD_WARPS * 32)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const float* __restrict__ gamma, const float* __restrict__ mean,
                     const float* __restrict__ rstd, __nv_bfloat16* __restrict__ dx,
                     __nv_bfloat16* __restrict__ dx_masked, float* __restrict__ dgamma,
                     float* __restrict__ dbeta, float* __restrict__ dxsum, int T, DropCfg drop_in, DropCfg drop_out) {
  constexpr int H = CH * 256;
  extern __shared__ float ln_bwd_smem[];           // [LN_BWD_WARPS + 1][H]: 52 KB at H = 1024
  float (*red)[CH * 256] = reinterpret_cast<float (*)[CH * 256]>(ln_bwd_smem);
  float* sacc = ln_bwd_smem + LN_BWD_WARPS * H;    // column sums of the gradient this kernel hands to the producer GEMM
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (dxsum != nullptr) {
    for (int i = threadIdx.x; i < H; i += LN_BWD_WARPS * 32) sacc[i] = 0.0f;
    __syncthreads();
  }
  const int warps_total = gridDim.x * LN_BWD_WARPS;
  const uint32_t seed_in = eff_seed(drop_in), seed_out = eff_seed(drop_out);
  float g[CH * 8], dg[CH * 8], db[CH * 8];
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const int col = c * 256 + lane * 8;
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + col));
    const float4 g1 = __ldg(reinterpret_cast<const float4*>(gamma + col + 4));
    g[c * 8 + 0] = g0.x; g[c * 8 + 1] = g0.y; g[c * 8 + 2] = g0.z; g[c * 8 + 3] = g0.w;
    g[c * 8 + 4] = g1.x; g[c * 8 + 5] = g1.y; g[c * 8 + 6] = g1.z; g[c * 8 + 7] = g1.w;
  }
#pragma unroll
  for (int i = 0; i < CH * 8; ++i) { dg[i] = 0.0f; db[i] = 0.0f; }

  for (int row = blockIdx.x * LN_BWD_WARPS + warp; row < T; row += warps_total) {
    const long long off = static_cast<long long>(row) * H;
    const uint4* xr = reinterpret_cast<const uint4*>(x + off);
    const uint4* dr = reinterpret_cast<const uint4*>(dy + off);
    float xh[CH * 8], d[CH * 8];
    const float mu = __ldg(mean + row), rs = __ldg(rstd + row);
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      unpack8(__ldg(xr + c * 32 + lane), xh + c * 8);
      unpack8(__ldg(dr + c * 32 + lane), d + c * 8);
    }
    float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
    for (int c = 0; c < CH; ++c) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int i = c * 8 + e;
        if (drop_in.threshold != 0u)
          d[i] = dropout_keep(seed_in, static_cast<uint32_t>(row) * H + c * 256 + lane * 8 + e, drop_in.threshold)
                     ? d[i] * drop_in.scale : 0.0f;
        xh[i] = (xh[i] - mu) * rs;
        dg[i] += d[i] * xh[i];
        db[i] += d[i];
        d[i] *= g[i];
        s1 += d[i];
        s2 += d[i] * xh[i];
      }
    }
    s1 = warp_sum(s1) * (1.0f / H);
    s2 = warp_sum(s2) * (1.0f / H);
    uint4* oxr = reinterpret_cast<uint4*>(dx + off);
    uint4* omr = dx_masked ? reinterpret_cast<uint4*>(dx_masked + off) : nullptr;
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = rs * (d[c * 8 + e] - s1 - xh[c * 8 + e] * s2);
      oxr[c * 32 + lane] = pack8(o);
      if (omr != nullptr) {
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (drop_out.threshold != 0u)
            o[e] = dropout_keep(seed_out, static_cast<uint32_t>(row) * H + c * 256 + lane * 8 + e, drop_out.threshold)
                       ? o[e] * drop_out.scale : 0.0f;
        omr[c * 32 + lane] = pack8(o);
      }
      if (dxsum != nullptr) {   // bias gradient of the GEMM in front: column sums of what that GEMM's backward reads
#pragma unroll
        for (int e = 0; e < 8; ++e) atomicAdd(sacc + c * 256 + lane * 8 + e, o[e]);
      }
    }
  }
  if (dxsum != nullptr) {
    __syncthreads();
    for (int col = threadIdx.x; col < H; col += LN_BWD_WARPS * 32) atomicAdd(dxsum + col, sacc[col]);
  }
  // block reduce of dgamma / dbeta, then one atomic per column per block
#pragma unroll 1
  for (int pass = 0; pass < 2; ++pass) {
    const float* src = pass == 0 ? dg : db;
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int e = 0; e < 8; ++e) red[warp][c * 256 + lane * 8 + e] = src[c * 8 + e];
    __syncthreads();
    float* dst = pass == 0 ? dgamma : dbeta;
    if (dst != nullptr)
      for (int col = threadIdx.x; col < H; col += LN_BWD_WARPS * 32) {
        float acc = 0.0f;
#pragma unroll
        for (int w = 0; w < LN_BWD_WARPS; ++w) acc += red[w][col];
        atomicAdd(dst + col, acc);
      }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------
// Column sum: out[n] += sum_t x[t, n]   (bias gradients).  N % 8 == 0.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
colsum_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, int T, int N, long long ld,
              const long long* __restrict__ sel, long long sel_value) {
  __shared__ float red[8][256 + 4];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = blockIdx.x * 256 + lane * 8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (col < N) {
    // four independent 16-byte loads in flight per thread (the column-strip lesson of the BatchNorm kernels)
    const int step = gridDim.y * 8;
    for (int row = blockIdx.y * 8 + warp; row < T; row += 4 * step) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int r = row + u * step;
        const bool take = r < T && (sel == nullptr || __ldg(sel + r) == sel_value);
        v[u] = take ? __ldg(reinterpret_cast<const uint4*>(x + static_cast<long long>(r) * ld + col)) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float f[8];
        unpack8(v[u], f);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += f[e];
      }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) red[warp][lane * 8 + e] = acc[e];
  __syncthreads();
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < N) {
    float s = 0.0f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w][threadIdx.x];
    atomicAdd(out + c, s);
  }
}

// ------------------------------------------------------------------------------------------------
// BERT embedding gather: out[t] = word[ids[t]] + pos[t % S] + type[tt[t]]   (bf16 out, fp32 tables)
// and its scatter-add backward.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
embed_gather_kernel(const long long* __restrict__ ids, const long long* __restrict__ tts,
                    const float* __restrict__ word, const float* __restrict__ pos,
                    const float* __restrict__ type, __nv_bfloat16* __restrict__ out, int T, int S, int H) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (row >= T) return;
  const float* w = word + ids[row] * H;
  const float* p = pos + static_cast<long long>(row % S) * H;
  const float* ty = type + (tts ? tts[row] : 0) * H;
  for (int c = lane * 8; c < H; c += 256) {
    float o[8];
#pragma unroll
    for (int e = 0; e < 8; e += 4) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(w + c + e));
      const float4 b = __ldg(reinterpret_cast<const float4*>(p + c + e));
      const float4 d = __ldg(reinterpret_cast<const float4*>(ty + c + e));
      o[e] = a.x + b.x + d.x; o[e + 1] = a.y + b.y + d.y; o[e + 2] = a.z + b.z + d.z; o[e + 3] = a.w + b.w + d.w;
    }
    *reinterpret_cast<uint4*>(out + static_cast<long long>(row) * H + c) = pack8(o);
  }
}

__device__ __forceinline__ void red_add_f32x4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(128)
embed_scatter_kernel(const long long* __restrict__ ids, const long long* __restrict__ tts,
                     const __nv_bfloat16* __restrict__ de, float* __restrict__ dword,
                     float* __restrict__ dpos, float* __restrict__ dtype, int T, int S, int H) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (row >= T) return;
  float* w = dword + ids[row] * H;
  float* p = dpos ? dpos + static_cast<long long>(row % S) * H : nullptr;   // null: the caller reduces positions over the batch
  float* ty = dtype ? dtype + (tts ? tts[row] : 0) * H : nullptr;
  for (int c = lane * 8; c < H; c += 256) {
    float f[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(de + static_cast<long long>(row) * H + c)), f);
    // 16-byte vector reductions: two per lane cover its 8 columns (scalar atomics here were 4.7 M requests per BERT
    // micro-batch, each warp instruction touching 32 sectors — 210 us for a 19 MB update)
    red_add_f32x4(w + c, f[0], f[1], f[2], f[3]);
    red_add_f32x4(w + c + 4, f[4], f[5], f[6], f[7]);
    if (dpos != nullptr) {
      red_add_f32x4(p + c, f[0], f[1], f[2], f[3]);
      red_add_f32x4(p + c + 4, f[4], f[5], f[6], f[7]);
    }
    if (dtype != nullptr) {
      red_add_f32x4(ty + c, f[0], f[1], f[2], f[3]);
      red_add_f32x4(ty + c + 4, f[4], f[5], f[6], f[7]);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Softmax cross-entropy over fp32 logits [B, C]: per-row loss and dlogits = (softmax - onehot) * gscale
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
softmax_ce_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                  float* __restrict__ loss, float* __restrict__ dlogits, int B, int C, float gscale) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (row >= B) return;
  const float* l = logits + static_cast<long long>(row) * C;
  float mx = -INFINITY;
  for (int c = lane; c < C; c += 32) mx = fmaxf(mx, l[c]);
  mx = warp_max(mx);
  float s = 0.0f;
  for (int c = lane; c < C; c += 32) s += __expf(l[c] - mx);
  s = warp_sum(s);
  const float lse = mx + __logf(s);
  const long long lab = labels[row];
  if (lane == 0) loss[row] = lse - l[lab];
  if (dlogits != nullptr) {
    float* d = dlogits + static_cast<long long>(row) * C;
    for (int c = lane; c < C; c += 32) d[c] = (__expf(l[c] - lse) - (c == lab ? 1.0f : 0.0f)) * gscale;
  }
}

// ------------------------------------------------------------------------------------------------
// Tiny strided GEMM on CUDA cores: C[m,n] (+)= act(alpha * sum_k A[m,k] B[n,k] + bias[n]).
// One warp per output element. For heads where M*N is a few thousand (pooler, classifier).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float load_any(const void* p, long long i, int is_f32) {
  return is_f32 ? reinterpret_cast<const float*>(p)[i] : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
}
__global__ void __launch_bounds__(256)
gemm_small_kernel(const void* __restrict__ A, const void* __restrict__ B, void* __restrict__ C, int M,
                  int N, int K, long long sam, long long sak, long long sbn, long long sbk,
                  long long scm, long long scn, int a_f32, int b_f32, int c_f32, const float* bias,
                  int act, float alpha, int accumulate) {
  const int lane = threadIdx.x & 31;
  const long long o = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (o >= static_cast<long long>(M) * N) return;
  const int m = static_cast<int>(o / N), n = static_cast<int>(o % N);
  float acc = 0.0f;
  for (int k = lane; k < K; k += 32) acc += load_any(A, m * sam + k * sak, a_f32) * load_any(B, n * sbn + k * sbk, b_f32);
  acc = warp_sum(acc);
  if (lane == 0) {
    acc *= alpha;
    if (bias != nullptr) acc += bias[n];
    if (act == 1) acc = tanhf(acc);
    const long long ci = m * scm + n * scn;
    if (c_f32) {
      float* c = reinterpret_cast<float*>(C);
      c[ci] = accumulate ? c[ci] + acc : acc;
    } else {
      __nv_bfloat16* c = reinterpret_cast<__nv_bfloat16*>(C);
      c[ci] = __float2bfloat16_rn(accumulate ? __bfloat162float(c[ci]) + acc : acc);
    }
  }
}

// Same contract for very short contractions (K <= 8, e.g. the input gradient of a two-logit head): a thread per output.
__global__ void __launch_bounds__(256)
gemm_smallk_kernel(const void* __restrict__ A, const void* __restrict__ B, void* __restrict__ C, int M, int N, int K,
                   long long sam, long long sak, long long sbn, long long sbk, long long scm, long long scn, int a_f32,
                   int b_f32, int c_f32, const float* bias, int act, float alpha, int accumulate) {
  const long long o = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (o >= static_cast<long long>(M) * N) return;
  const int m = static_cast<int>(o / N), n = static_cast<int>(o % N);
  float acc = 0.0f;
  for (int k = 0; k < K; ++k) acc += load_any(A, m * sam + k * sak, a_f32) * load_any(B, n * sbn + k * sbk, b_f32);
  acc *= alpha;
  if (bias != nullptr) acc += bias[n];
  if (act == 1) acc = tanhf(acc);
  const long long ci = m * scm + n * scn;
  if (c_f32) {
    float* c = reinterpret_cast<float*>(C);
    c[ci] = accumulate ? c[ci] + acc : acc;
  } else {
    __nv_bfloat16* c = reinterpret_cast<__nv_bfloat16*>(C);
    c[ci] = __float2bfloat16_rn(accumulate ? __bfloat162float(c[ci]) + acc : acc);
  }
}

// ------------------------------------------------------------------------------------------------
// fp32 -> bf16 cast (shadow weights) and bf16 transposed copy helpers
// ------------------------------------------------------------------------------------------------
__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n) {
  const long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 8;
  if (i + 8 <= n) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(in + i));
    const float4 b = __ldg(reinterpret_cast<const float4*>(in + i + 4));
    const float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    *reinterpret_cast<uint4*>(out + i) = pack8(f);
  } else {
    for (long long j = i; j < n; ++j) out[j] = __float2bfloat16_rn(in[j]);
  }
}

template <int CH>
static int launch_ln_fwd(const void* x, const float* g, const float* b, void* y, float* mean, float* rstd,
                         int T, float eps, DropCfg d, cudaStream_t st) {
  const int blocks = min((T + 3) / 4, sm_count() * 8);
  layernorm_fwd_kernel<CH><<<blocks, 128, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(x), g, b,
                                                  reinterpret_cast<__nv_bfloat16*>(y), mean, rstd, T, eps, d);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}
template <int CH>
static int launch_ln_bwd(const void* dy, const void* x, const float* g, const float* mean, const float* rstd,
                         void* dx, void* dxm, float* dg, float* db, float* dxsum, int T, DropCfg din, DropCfg dout, cudaStream_t st) {
  const int blocks = min((T + LN_BWD_WARPS - 1) / LN_BWD_WARPS, sm_count());
  constexpr int smem = (LN_BWD_WARPS + 1) * CH * 256 * static_cast<int>(sizeof(float));
  static bool attr_set = false;
  if (smem > 48 * 1024 && !attr_set) {
    DET_CHECK_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel<CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  layernorm_bwd_kernel<CH><<<blocks, LN_BWD_WARPS * 32, smem, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const __nv_bfloat16*>(x), g, mean, rstd,
      reinterpret_cast<__nv_bfloat16*>(dx), reinterpret_cast<__nv_bfloat16*>(dxm), dg, db, dxsum, T, din, dout);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------
// LayerNorm for any width (H % 8 == 0; ALBERT's 128-wide embeddings, 4096-wide xxlarge layers): a warp per row,
// lanes striding over 16-byte vectors, the row re-read from L1 for each pass. Same dropout semantics as above.
// ------------------------------------------------------------------------------------------------
constexpr int LNG_WARPS = 8;

__global__ void __launch_bounds__(LNG_WARPS * 32)
layernorm_fwd_generic_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
                             const float* __restrict__ beta, __nv_bfloat16* __restrict__ y, float* __restrict__ mean,
                             float* __restrict__ rstd, int T, int H, float eps, DropCfg drop) {
  const int lane = threadIdx.x & 31;
  const int nvec = H / 8;
  const float inv_h = 1.0f / static_cast<float>(H);
  const uint32_t dseed = eff_seed(drop);
  for (int row = blockIdx.x * LNG_WARPS + (threadIdx.x >> 5); row < T; row += gridDim.x * LNG_WARPS) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + static_cast<long long>(row) * H);
    float s = 0.0f;
    for (int v = lane; v < nvec; v += 32) {
      float f[8];
      unpack8(xr[v], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) s += f[e];
    }
    const float mu = warp_sum(s) * inv_h;
    float ss = 0.0f;
    for (int v = lane; v < nvec; v += 32) {
      float f[8];
      unpack8(xr[v], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) { const float d = f[e] - mu; ss += d * d; }
    }
    const float rs = rsqrtf(warp_sum(ss) * inv_h + eps);
    if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
    uint4* yr = reinterpret_cast<uint4*>(y + static_cast<long long>(row) * H);
    for (int v = lane; v < nvec; v += 32) {
      float f[8], o[8];
      unpack8(xr[v], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int col = v * 8 + e;
        float t = (f[e] - mu) * rs * __ldg(gamma + col) + __ldg(beta + col);
        if (drop.threshold != 0u)
          t = dropout_keep(dseed, static_cast<uint32_t>(row) * H + col, drop.threshold) ? t * drop.scale : 0.0f;
        o[e] = t;
      }
      yr[v] = pack8(o);
    }
  }
}

__global__ void __launch_bounds__(LNG_WARPS * 32)
layernorm_bwd_generic_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                             const float* __restrict__ gamma, const float* __restrict__ mean,
                             const float* __restrict__ rstd, __nv_bfloat16* __restrict__ dx,
                             __nv_bfloat16* __restrict__ dx_masked, float* __restrict__ dgamma, float* __restrict__ dbeta,
                             float* __restrict__ dxsum, int T, int H, DropCfg drop_in, DropCfg drop_out) {
  extern __shared__ float lng_red[];   // [3][H]: per-block dgamma / dbeta / column sums of the outgoing gradient
  float* sg = lng_red;
  float* sb = lng_red + H;
  float* sx = lng_red + 2 * H;
  for (int i = threadIdx.x; i < 3 * H; i += blockDim.x) lng_red[i] = 0.0f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int nvec = H / 8;
  const float inv_h = 1.0f / static_cast<float>(H);
  const uint32_t seed_in = eff_seed(drop_in), seed_out = eff_seed(drop_out);
  for (int row = blockIdx.x * LNG_WARPS + (threadIdx.x >> 5); row < T; row += gridDim.x * LNG_WARPS) {
    const long long off = static_cast<long long>(row) * H;
    const uint4* xr = reinterpret_cast<const uint4*>(x + off);
    const uint4* dr = reinterpret_cast<const uint4*>(dy + off);
    const float mu = __ldg(mean + row), rs = __ldg(rstd + row);
    float s1 = 0.0f, s2 = 0.0f;
    for (int v = lane; v < nvec; v += 32) {
      float xf[8], d[8];
      unpack8(xr[v], xf);
      unpack8(dr[v], d);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int col = v * 8 + e;
        if (drop_in.threshold != 0u)
          d[e] = dropout_keep(seed_in, static_cast<uint32_t>(row) * H + col, drop_in.threshold) ? d[e] * drop_in.scale : 0.0f;
        const float xh = (xf[e] - mu) * rs;
        atomicAdd(sg + col, d[e] * xh);
        atomicAdd(sb + col, d[e]);
        const float dgm = d[e] * __ldg(gamma + col);
        s1 += dgm;
        s2 += dgm * xh;
      }
    }
    s1 = warp_sum(s1) * inv_h;
    s2 = warp_sum(s2) * inv_h;
    uint4* oxr = reinterpret_cast<uint4*>(dx + off);
    uint4* omr = dx_masked ? reinterpret_cast<uint4*>(dx_masked + off) : nullptr;
    for (int v = lane; v < nvec; v += 32) {
      float xf[8], d[8], o[8];
      unpack8(xr[v], xf);
      unpack8(dr[v], d);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int col = v * 8 + e;
        if (drop_in.threshold != 0u)
          d[e] = dropout_keep(seed_in, static_cast<uint32_t>(row) * H + col, drop_in.threshold) ? d[e] * drop_in.scale : 0.0f;
        const float xh = (xf[e] - mu) * rs;
        o[e] = rs * (d[e] * __ldg(gamma + col) - s1 - xh * s2);
      }
      oxr[v] = pack8(o);
      if (omr != nullptr) {
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (drop_out.threshold != 0u)
            o[e] = dropout_keep(seed_out, static_cast<uint32_t>(row) * H + v * 8 + e, drop_out.threshold) ? o[e] * drop_out.scale : 0.0f;
        omr[v] = pack8(o);
      }
      if (dxsum != nullptr) {
#pragma unroll
        for (int e = 0; e < 8; ++e) atomicAdd(sx + v * 8 + e, o[e]);
      }
    }
  }
  __syncthreads();
  for (int col = threadIdx.x; col < H; col += blockDim.x) {
    if (dgamma != nullptr) atomicAdd(dgamma + col, sg[col]);
    if (dbeta != nullptr) atomicAdd(dbeta + col, sb[col]);
    if (dxsum != nullptr) atomicAdd(dxsum + col, sx[col]);
  }
}

static int launch_ln_fwd_generic(const void* x, const float* g, const float* b, void* y, float* mean, float* rstd, int T,
                                 int H, float eps, DropCfg d, cudaStream_t st) {
  const int blocks = min((T + LNG_WARPS - 1) / LNG_WARPS, sm_count() * 8);
  layernorm_fwd_generic_kernel<<<blocks, LNG_WARPS * 32, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(x), g, b,
                                                                 reinterpret_cast<__nv_bfloat16*>(y), mean, rstd, T, H, eps, d);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}
static int launch_ln_bwd_generic(const void* dy, const void* x, const float* g, const float* mean, const float* rstd,
                                 void* dx, void* dxm, float* dg, float* db, float* dxsum, int T, int H, DropCfg din, DropCfg dout,
                                 cudaStream_t st) {
  const size_t smem = sizeof(float) * 3 * H;
  DET_REQUIRE(smem <= 200 * 1024, "layernorm: hidden size too large");
  if (smem > 48 * 1024)
    DET_CHECK_CUDA(cudaFuncSetAttribute(layernorm_bwd_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  const int blocks = min((T + LNG_WARPS - 1) / LNG_WARPS, sm_count() * 2);
  layernorm_bwd_generic_kernel<<<blocks, LNG_WARPS * 32, smem, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<const __nv_bfloat16*>(x), g, mean, rstd,
      reinterpret_cast<__nv_bfloat16*>(dx), reinterpret_cast<__nv_bfloat16*>(dxm), dg, db, dxsum, T, H, din, dout);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace det

#define DET_LN_DISPATCH(FN, ...)                                           \
  switch (H / 256) {                                                       \
    case 1: return det::FN<1>(__VA_ARGS__);                                \
    case 2: return det::FN<2>(__VA_ARGS__);                                \
    case 3: return det::FN<3>(__VA_ARGS__);                                \
    case 4: return det::FN<4>(__VA_ARGS__);                                \
    default: break;                                                        \
  }                                                                        \
  det::set_last_error("layernorm: unexpected hidden size");               \
  return 2;

extern "C" int det_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y,
                                 float* mean, float* rstd, int T, int H, float eps, float dropout_p,
                                 unsigned int seed, const unsigned int* seed_dev, cudaStream_t stream) {
  DET_REQUIRE(H >= 8 && H % 8 == 0, "layernorm hidden size must be a multiple of 8");
  det::DropCfg d = det::make_drop(dropout_p, seed, seed_dev);
  if (H % 256 != 0 || H > 1024) return det::launch_ln_fwd_generic(x, gamma, beta, y, mean, rstd, T, H, eps, d, stream);
  DET_LN_DISPATCH(launch_ln_fwd, x, gamma, beta, y, mean, rstd, T, eps, d, stream)
}

extern "C" int det_layernorm_bwd(const void* dy, const void* x, const float* gamma, const float* mean,
                                 const float* rstd, void* dx, void* dx_masked, float* dgamma,
                                 float* dbeta, float* dxsum, int T, int H, float in_dropout_p, unsigned int in_seed,
                                 float out_dropout_p, unsigned int out_seed,
                                 const unsigned int* seed_dev, cudaStream_t stream) {
  DET_REQUIRE(H >= 8 && H % 8 == 0, "layernorm hidden size must be a multiple of 8");
  det::DropCfg din = det::make_drop(in_dropout_p, in_seed, seed_dev);
  det::DropCfg dout = det::make_drop(out_dropout_p, out_seed, seed_dev);
  if (H % 256 != 0 || H > 1024)
    return det::launch_ln_bwd_generic(dy, x, gamma, mean, rstd, dx, dx_masked, dgamma, dbeta, dxsum, T, H, din, dout, stream);
  DET_LN_DISPATCH(launch_ln_bwd, dy, x, gamma, mean, rstd, dx, dx_masked, dgamma, dbeta, dxsum, T, din, dout, stream)
}

extern "C" int det_colsum_bf16(const void* x, float* out, int T, int N, long long ld, cudaStream_t stream) {
  DET_REQUIRE(N % 8 == 0 && ld % 8 == 0, "colsum needs N and ld multiples of 8");
  const int gx = (N + 255) / 256;
  int gy = std::max(1, std::min((T + 31) / 32, (det::sm_count() * 2 + gx - 1) / gx));   // few fat blocks: one atomic per column per block at the end
  det::colsum_kernel<<<dim3(gx, gy), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), out, T, N, ld,
                                                       nullptr, 0);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// out[n] += sum over rows t with sel[t] == sel_value of x[t, n]  (token-type embedding gradients:
// a handful of destination rows, so a reduction instead of millions of same-address atomics).
extern "C" int det_colsum_select_bf16(const void* x, const long long* sel, long long sel_value, float* out,
                                      int T, int N, long long ld, cudaStream_t stream) {
  DET_REQUIRE(N % 8 == 0 && ld % 8 == 0, "colsum needs N and ld multiples of 8");
  const int gx = (N + 255) / 256;
  int gy = std::max(1, std::min((T + 31) / 32, (det::sm_count() * 2 + gx - 1) / gx));   // few fat blocks: one atomic per column per block at the end
  det::colsum_kernel<<<dim3(gx, gy), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), out, T, N, ld,
                                                       sel, sel_value);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_embed_gather(const long long* ids, const long long* type_ids, const float* word,
                                const float* pos, const float* type, void* out, int T, int S, int H,
                                cudaStream_t stream) {
  DET_REQUIRE(H % 8 == 0, "embedding width must be a multiple of 8");
  det::embed_gather_kernel<<<(T + 3) / 4, 128, 0, stream>>>(ids, type_ids, word, pos, type,
                                                           reinterpret_cast<__nv_bfloat16*>(out), T, S, H);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_embed_scatter(const long long* ids, const long long* type_ids, const void* de,
                                 float* dword, float* dpos, float* dtype, int T, int S, int H,
                                 cudaStream_t stream) {
  DET_REQUIRE(H % 8 == 0, "embedding width must be a multiple of 8");
  det::embed_scatter_kernel<<<(T + 3) / 4, 128, 0, stream>>>(ids, type_ids, reinterpret_cast<const __nv_bfloat16*>(de),
                                                            dword, dpos, dtype, T, S, H);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_softmax_ce(const float* logits, const long long* labels, float* loss, float* dlogits,
                              int B, int C, float gscale, cudaStream_t stream) {
  det::softmax_ce_kernel<<<(B + 3) / 4, 128, 0, stream>>>(logits, labels, loss, dlogits, B, C, gscale);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_gemm_small(const void* A, const void* B, void* C, int M, int N, int K, long long sam,
                              long long sak, long long sbn, long long sbk, long long scm, long long scn,
                              int a_f32, int b_f32, int c_f32, const float* bias, int act, float alpha,
                              int accumulate, cudaStream_t stream) {
  const long long outs = static_cast<long long>(M) * N;
  DET_REQUIRE(outs > 0 && outs < (1LL << 31), "gemm_small: bad output size");
  if (K <= 8) {
    det::gemm_smallk_kernel<<<static_cast<unsigned>((outs + 255) / 256), 256, 0, stream>>>(
        A, B, C, M, N, K, sam, sak, sbn, sbk, scm, scn, a_f32, b_f32, c_f32, bias, act, alpha, accumulate);
    DET_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  det::gemm_small_kernel<<<static_cast<unsigned>((outs + 7) / 8), 256, 0, stream>>>(
      A, B, C, M, N, K, sam, sak, sbn, sbk, scm, scn, a_f32, b_f32, c_f32, bias, act, alpha, accumulate);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_cast_f32_bf16(const float* in, void* out, long long n, cudaStream_t stream) {
  if (n <= 0) return 0;
  DET_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
              "cast: pointers must be 16-byte aligned");
  const long long threads = (n + 7) / 8;
  det::cast_f32_bf16_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, stream>>>(
      in, reinterpret_cast<__nv_bfloat16*>(out), n);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}
