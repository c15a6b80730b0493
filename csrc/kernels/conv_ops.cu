// conv_ops.cu — the bandwidth-bound half of the convolutional models (NHWC bf16 activations):
//   * gather / scatter-free col2im for the convolutions the implicit-GEMM kernel does not take
//     (stride > 1, or channel counts that are not a multiple of 8 such as the RGB stem),
//   * (training-mode BatchNorm lives in bn_ops.cu),
//   * max / global-average pooling, elementwise activations, stateless-hash dropout, NCHW fp32 -> NHWC bf16.
// Every kernel moves 16 bytes per thread per access where the channel count allows it.
//
// Reference counterparts are library calls made by user model code (cuDNN BatchNorm / pooling through
// torch.nn in e.g. examples/computer_vision/cifar10_pytorch/model_def.py:49-74 and
// examples/gan/gan_mnist_pytorch/model_def.py:30-72); SURVEY.md §2.9 K14 lists them as kernels to author.
#include "host_utils.h"
#include "sm100_ptx.cuh"

namespace det {

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

struct ConvGeom {
  int N, H, W, C, R, S, stride, pad_h, pad_w, Ho, Wo, Kp;
};

// ------------------------------------------------------------------------------------------------
// im2col: col[(n,ho,wo), (r,s,c)] = x[n, ho*stride + r - pad, wo*stride + s - pad, c]; columns >= R*S*C are zero.
// ------------------------------------------------------------------------------------------------
template <bool VEC>
__global__ void __launch_bounds__(256)
im2col_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col, const ConvGeom g, long long total) {
  const long long t = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;
  if (t >= total) return;
  const int kg = g.Kp / 8;
  const int k8 = static_cast<int>(t % kg);
  long long row = t / kg;
  const int wo = static_cast<int>(row % g.Wo);
  const long long r2 = row / g.Wo;
  const int ho = static_cast<int>(r2 % g.Ho);
  const int n = static_cast<int>(r2 / g.Ho);
  const int K = g.R * g.S * g.C;
  uint4 outv = make_uint4(0, 0, 0, 0);
  if (VEC) {
    const int k = k8 * 8;
    if (k < K) {
      const int tap = k / g.C, c = k - tap * g.C;
      const int fr = tap / g.S, fs = tap - fr * g.S;
      const int h = ho * g.stride + fr - g.pad_h, w = wo * g.stride + fs - g.pad_w;
      if (h >= 0 && h < g.H && w >= 0 && w < g.W)
        outv = __ldg(reinterpret_cast<const uint4*>(x + ((static_cast<long long>(n) * g.H + h) * g.W + w) * g.C + c));
    }
  } else {
    // few-channel layers (RGB stem): decode (tap row, tap column, channel) once, then step — no division per element
    uint32_t bits[8];
    int k = k8 * 8;
    const int tap0 = k / g.C;
    int c = k - tap0 * g.C;
    int fr = tap0 / g.S, fs = tap0 - fr * g.S;
    const long long img = static_cast<long long>(n) * g.H;
#pragma unroll
    for (int j = 0; j < 8; ++j, ++k) {
      bits[j] = 0;
      if (k < K) {
        const int h = ho * g.stride + fr - g.pad_h, w = wo * g.stride + fs - g.pad_w;
        if (h >= 0 && h < g.H && w >= 0 && w < g.W)
          bits[j] = __ldg(reinterpret_cast<const unsigned short*>(x) + ((img + h) * g.W + w) * g.C + c);
      }
      if (++c == g.C) { c = 0; if (++fs == g.S) { fs = 0; ++fr; } }
    }
    outv.x = bits[0] | (bits[1] << 16); outv.y = bits[2] | (bits[3] << 16);
    outv.z = bits[4] | (bits[5] << 16); outv.w = bits[6] | (bits[7] << 16);
  }
  *reinterpret_cast<uint4*>(col + row * g.Kp + k8 * 8) = outv;
}

// ------------------------------------------------------------------------------------------------
// col2im in gather form (deterministic, no atomics): dx[n,h,w,c] = sum over taps whose window covers (h,w).
// ------------------------------------------------------------------------------------------------
template <bool VEC>
__global__ void __launch_bounds__(256)
col2im_kernel(const __nv_bfloat16* __restrict__ col, __nv_bfloat16* __restrict__ dx, const ConvGeom g, long long total) {
  const long long t = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;
  if (t >= total) return;
  const int cg = VEC ? g.C / 8 : g.C;
  const int c = static_cast<int>(t % cg) * (VEC ? 8 : 1);
  long long pix = t / cg;
  const int w = static_cast<int>(pix % g.W);
  const long long r2 = pix / g.W;
  const int h = static_cast<int>(r2 % g.H);
  const int n = static_cast<int>(r2 / g.H);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int fr = 0; fr < g.R; ++fr) {
    const int hn = h + g.pad_h - fr;
    if (hn < 0 || hn % g.stride != 0) continue;
    const int ho = hn / g.stride;
    if (ho >= g.Ho) continue;
    for (int fs = 0; fs < g.S; ++fs) {
      const int wn = w + g.pad_w - fs;
      if (wn < 0 || wn % g.stride != 0) continue;
      const int wo = wn / g.stride;
      if (wo >= g.Wo) continue;
      const __nv_bfloat16* src = col + ((static_cast<long long>(n) * g.Ho + ho) * g.Wo + wo) * g.Kp + (fr * g.S + fs) * g.C + c;
      if (VEC) {
        float f[8];
        unpack8(__ldg(reinterpret_cast<const uint4*>(src)), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      } else {
        acc[0] += __bfloat162float(*src);
      }
    }
  }
  __nv_bfloat16* dst = dx + pix * g.C + c;
  if (VEC) *reinterpret_cast<uint4*>(dst) = pack8(acc);
  else *dst = __float2bfloat16_rn(acc[0]);
}

// ------------------------------------------------------------------------------------------------
// Pooling (NHWC, C % 8 == 0)
// ------------------------------------------------------------------------------------------------
struct PoolGeom {
  int N, H, W, C, k, stride, pad, Ho, Wo;
};

// y = max over the window; idx = position (r*k + s) of the first maximum, which is what backward routes to.
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, uint8_t* __restrict__ idx,
                   const PoolGeom g, long long total) {
  const long long t = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;
  if (t >= total) return;
  const int cg = g.C / 8;
  const int c = static_cast<int>(t % cg) * 8;
  const long long pix = t / cg;
  const int wo = static_cast<int>(pix % g.Wo);
  const long long r2 = pix / g.Wo;
  const int ho = static_cast<int>(r2 % g.Ho);
  const int n = static_cast<int>(r2 / g.Ho);
  float best[8];
  int bi[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; bi[j] = 0; }
  for (int fr = 0; fr < g.k; ++fr) {
    const int h = ho * g.stride + fr - g.pad;
    if (h < 0 || h >= g.H) continue;
    for (int fs = 0; fs < g.k; ++fs) {
      const int w = wo * g.stride + fs - g.pad;
      if (w < 0 || w >= g.W) continue;
      float f[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(x + ((static_cast<long long>(n) * g.H + h) * g.W + w) * g.C + c)), f);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (f[j] > best[j]) { best[j] = f[j]; bi[j] = fr * g.k + fs; }
    }
  }
  *reinterpret_cast<uint4*>(y + pix * g.C + c) = pack8(best);
  uint2 pk;
  pk.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
  pk.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
  *reinterpret_cast<uint2*>(idx + pix * g.C + c) = pk;
}

__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ idx, __nv_bfloat16* __restrict__ dx,
                   const PoolGeom g, long long total) {
  const long long t = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;
  if (t >= total) return;
  const int cg = g.C / 8;
  const int c = static_cast<int>(t % cg) * 8;
  const long long pix = t / cg;
  const int w = static_cast<int>(pix % g.W);
  const long long r2 = pix / g.W;
  const int h = static_cast<int>(r2 % g.H);
  const int n = static_cast<int>(r2 / g.H);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int fr = 0; fr < g.k; ++fr) {
    const int hn = h + g.pad - fr;
    if (hn < 0 || hn % g.stride != 0) continue;
    const int ho = hn / g.stride;
    if (ho >= g.Ho) continue;
    for (int fs = 0; fs < g.k; ++fs) {
      const int wn = w + g.pad - fs;
      if (wn < 0 || wn % g.stride != 0) continue;
      const int wo = wn / g.stride;
      if (wo >= g.Wo) continue;
      const long long o = ((static_cast<long long>(n) * g.Ho + ho) * g.Wo + wo) * g.C + c;
      const uint2 pk = __ldg(reinterpret_cast<const uint2*>(idx + o));
      float f[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(dy + o)), f);
      const int me = fr * g.k + fs;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int sel = ((j < 4 ? pk.x : pk.y) >> (8 * (j & 3))) & 0xFF;
        if (sel == me) acc[j] += f[j];
      }
    }
  }
  *reinterpret_cast<uint4*>(dx + pix * g.C + c) = pack8(acc);
}

// y[n, c] = mean over HW of x[n, hw, c]; one thread per (n, 8 channels)
__global__ void __launch_bounds__(256)
avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int N, int HW, int C) {
  const int t = blockIdx.x * 256 + threadIdx.x;
  const int cg = C / 8;
  if (t >= N * cg) return;
  const int n = t / cg, c = (t % cg) * 8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int i = 0; i < HW; ++i) {
    float f[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(x + (static_cast<long long>(n) * HW + i) * C + c)), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] += f[j];
  }
  const float inv = 1.0f / HW;
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] *= inv;
  *reinterpret_cast<uint4*>(y + static_cast<long long>(n) * C + c) = pack8(acc);
}

__global__ void __launch_bounds__(256)
avgpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, long long total_vec, int HW, int C) {
  const long long t = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;
  if (t >= total_vec) return;
  const int cg = C / 8;
  const int c = static_cast<int>(t % cg) * 8;
  const long long n = t / cg / HW;
  float f[8];
  unpack8(__ldg(reinterpret_cast<const uint4*>(dy + n * C + c)), f);
  const float inv = 1.0f / HW;
#pragma unroll
  for (int j = 0; j < 8; ++j) f[j] *= inv;
  reinterpret_cast<uint4*>(dx)[t] = pack8(f);
}

// ------------------------------------------------------------------------------------------------
// Elementwise: activations (fwd from x, bwd from the saved output), dropout, layout conversion
// ------------------------------------------------------------------------------------------------
enum Act : int { ACT_RELU = 1, ACT_LEAKY = 2, ACT_TANH = 3, ACT_SIGMOID = 4 };

__global__ void __launch_bounds__(256)
act_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n, int act, float slope) {
  const long long i = (static_cast<long long>(blockIdx.x) * 256 + threadIdx.x) * 8;
  if (i >= n) return;
  float f[8];
  if (i + 8 <= n) unpack8(__ldg(reinterpret_cast<const uint4*>(x + i)), f);
  else for (int j = 0; j < 8; ++j) f[j] = i + j < n ? __bfloat162float(x[i + j]) : 0.0f;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float v = f[j];
    if (act == ACT_RELU) v = fmaxf(v, 0.0f);
    else if (act == ACT_LEAKY) v = v > 0.0f ? v : v * slope;
    else if (act == ACT_TANH) v = tanhf(v);
    else if (act == ACT_SIGMOID) v = 1.0f / (1.0f + __expf(-v));
    f[j] = v;
  }
  if (i + 8 <= n) *reinterpret_cast<uint4*>(y + i) = pack8(f);
  else for (int j = 0; j < 8; ++j) if (i + j < n) y[i + j] = __float2bfloat16_rn(f[j]);
}

// dx = dy * act'(.) expressed through the saved OUTPUT y
__global__ void __launch_bounds__(256)
act_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ dx,
               long long n, int act, float slope) {
  const long long i = (static_cast<long long>(blockIdx.x) * 256 + threadIdx.x) * 8;
  if (i >= n) return;
  float d[8], o[8];
  if (i + 8 <= n) {
    unpack8(__ldg(reinterpret_cast<const uint4*>(dy + i)), d);
    unpack8(__ldg(reinterpret_cast<const uint4*>(y + i)), o);
  } else {
    for (int j = 0; j < 8; ++j) {
      d[j] = i + j < n ? __bfloat162float(dy[i + j]) : 0.0f;
      o[j] = i + j < n ? __bfloat162float(y[i + j]) : 0.0f;
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float g = 1.0f;
    if (act == ACT_RELU) g = o[j] > 0.0f ? 1.0f : 0.0f;
    else if (act == ACT_LEAKY) g = o[j] > 0.0f ? 1.0f : slope;
    else if (act == ACT_TANH) g = 1.0f - o[j] * o[j];
    else if (act == ACT_SIGMOID) g = o[j] * (1.0f - o[j]);
    d[j] *= g;
  }
  if (i + 8 <= n) *reinterpret_cast<uint4*>(dx + i) = pack8(d);
  else for (int j = 0; j < 8; ++j) if (i + j < n) dx[i + j] = __float2bfloat16_rn(d[j]);
}

// y = keep(seed, salt + i) ? x / (1 - p) : 0. Backward is the same kernel applied to dy.
__global__ void __launch_bounds__(256)
dropout_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n, uint32_t threshold,
               float scale, uint32_t seed, const uint32_t* __restrict__ seed_dev) {
  const long long i = (static_cast<long long>(blockIdx.x) * 256 + threadIdx.x) * 8;
  if (i >= n) return;
  const uint32_t s = dropout_effective_seed(seed, seed_dev);
  float f[8];
  if (i + 8 <= n) unpack8(__ldg(reinterpret_cast<const uint4*>(x + i)), f);
  else for (int j = 0; j < 8; ++j) f[j] = i + j < n ? __bfloat162float(x[i + j]) : 0.0f;
#pragma unroll
  for (int j = 0; j < 8; ++j) f[j] = dropout_keep(s, static_cast<uint32_t>(i + j), threshold) ? f[j] * scale : 0.0f;
  if (i + 8 <= n) *reinterpret_cast<uint4*>(y + i) = pack8(f);
  else for (int j = 0; j < 8; ++j) if (i + j < n) y[i + j] = __float2bfloat16_rn(f[j]);
}

// NCHW fp32 -> NHWC bf16 through a 32x32 shared-memory transpose of the [C, HW] plane of each image.
__global__ void __launch_bounds__(256)
nchw_to_nhwc_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, int C, int HW) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8) {
    const int c = c0 + j, p = p0 + tx;
    tile[j][tx] = (c < C && p < HW) ? in[(static_cast<long long>(n) * C + c) * HW + p] : 0.0f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int p = p0 + j, c = c0 + tx;
    if (c < C && p < HW) out[(static_cast<long long>(n) * HW + p) * C + c] = __float2bfloat16_rn(tile[tx][j]);
  }
}

// uint8 NHWC (decoded image bytes) -> bf16 NHWC with per-channel (x/255 - mean) / std folded into scale/shift.
// 16 input bytes per thread; C <= 8.
struct U8Norm { float scale[8], shift[8]; int C; };
__global__ void __launch_bounds__(256)
u8_to_bf16_norm_kernel(const uint8_t* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n, const U8Norm p) {
  const long long i = (static_cast<long long>(blockIdx.x) * 256 + threadIdx.x) * 16;
  if (i >= n) return;
  int c = static_cast<int>(i % p.C);
  if (i + 16 <= n) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(in + i));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    float f[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float b = static_cast<float>((w[j >> 2] >> (8 * (j & 3))) & 0xFFu);
      f[j] = fmaf(b, p.scale[c], p.shift[c]);
      c = (c + 1 == p.C) ? 0 : c + 1;
    }
    uint4* o = reinterpret_cast<uint4*>(out + i);
    o[0] = pack8(f);
    o[1] = pack8(f + 8);
  } else {
    for (long long k = i; k < n; ++k) {
      out[k] = __float2bfloat16_rn(fmaf(static_cast<float>(in[k]), p.scale[c], p.shift[c]));
      c = (c + 1 == p.C) ? 0 : c + 1;
    }
  }
}

static inline unsigned blocks_for(long long n) { return static_cast<unsigned>((n + 255) / 256); }

}  // namespace det

using det::ConvGeom;
using det::PoolGeom;
typedef __nv_bfloat16 bf16;

static int fill_geom(ConvGeom& g, int N, int H, int W, int C, int R, int S, int stride, int pad_h, int pad_w, int Kp) {
  g.N = N; g.H = H; g.W = W; g.C = C; g.R = R; g.S = S; g.stride = stride; g.pad_h = pad_h; g.pad_w = pad_w;
  g.Ho = (H + 2 * pad_h - R) / stride + 1;
  g.Wo = (W + 2 * pad_w - S) / stride + 1;
  g.Kp = Kp;
  DET_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0 && R > 0 && S > 0 && stride > 0, "bad convolution geometry");
  DET_REQUIRE(g.Ho > 0 && g.Wo > 0, "convolution output is empty");
  DET_REQUIRE(Kp % 8 == 0 && Kp >= R * S * C, "Kp must be a multiple of 8 that covers R*S*C");
  return 0;
}

extern "C" int det_im2col(const void* x, void* col, int N, int H, int W, int C, int R, int S, int stride, int pad_h,
                          int pad_w, int Kp, cudaStream_t stream) {
  ConvGeom g;
  if (int rc = fill_geom(g, N, H, W, C, R, S, stride, pad_h, pad_w, Kp)) return rc;
  const long long total = static_cast<long long>(N) * g.Ho * g.Wo * (Kp / 8);
  if (C % 8 == 0)
    det::im2col_kernel<true><<<det::blocks_for(total), 256, 0, stream>>>((const bf16*)x, (bf16*)col, g, total);
  else
    det::im2col_kernel<false><<<det::blocks_for(total), 256, 0, stream>>>((const bf16*)x, (bf16*)col, g, total);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_col2im(const void* col, void* dx, int N, int H, int W, int C, int R, int S, int stride, int pad_h,
                          int pad_w, int Kp, cudaStream_t stream) {
  ConvGeom g;
  if (int rc = fill_geom(g, N, H, W, C, R, S, stride, pad_h, pad_w, Kp)) return rc;
  if (C % 8 == 0) {
    const long long total = static_cast<long long>(N) * H * W * (C / 8);
    det::col2im_kernel<true><<<det::blocks_for(total), 256, 0, stream>>>((const bf16*)col, (bf16*)dx, g, total);
  } else {
    const long long total = static_cast<long long>(N) * H * W * C;
    det::col2im_kernel<false><<<det::blocks_for(total), 256, 0, stream>>>((const bf16*)col, (bf16*)dx, g, total);
  }
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static int fill_pool(PoolGeom& g, int N, int H, int W, int C, int k, int stride, int pad) {
  g.N = N; g.H = H; g.W = W; g.C = C; g.k = k; g.stride = stride; g.pad = pad;
  g.Ho = (H + 2 * pad - k) / stride + 1;
  g.Wo = (W + 2 * pad - k) / stride + 1;
  DET_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0, "pooling needs C % 8 == 0");
  DET_REQUIRE(k > 0 && k <= 15 && stride > 0 && g.Ho > 0 && g.Wo > 0, "bad pooling window");
  return 0;
}

extern "C" int det_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride, int pad,
                               cudaStream_t stream) {
  PoolGeom g;
  if (int rc = fill_pool(g, N, H, W, C, k, stride, pad)) return rc;
  const long long total = static_cast<long long>(N) * g.Ho * g.Wo * (C / 8);
  det::maxpool_fwd_kernel<<<det::blocks_for(total), 256, 0, stream>>>((const bf16*)x, (bf16*)y, (uint8_t*)idx, g, total);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_maxpool_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int k, int stride,
                               int pad, cudaStream_t stream) {
  PoolGeom g;
  if (int rc = fill_pool(g, N, H, W, C, k, stride, pad)) return rc;
  const long long total = static_cast<long long>(N) * H * W * (C / 8);
  det::maxpool_bwd_kernel<<<det::blocks_for(total), 256, 0, stream>>>((const bf16*)dy, (const uint8_t*)idx, (bf16*)dx, g, total);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_avgpool_fwd(const void* x, void* y, int N, int HW, int C, cudaStream_t stream) {
  DET_REQUIRE(N > 0 && HW > 0 && C > 0 && C % 8 == 0, "avgpool needs C % 8 == 0");
  det::avgpool_fwd_kernel<<<(N * (C / 8) + 255) / 256, 256, 0, stream>>>((const bf16*)x, (bf16*)y, N, HW, C);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, cudaStream_t stream) {
  DET_REQUIRE(N > 0 && HW > 0 && C > 0 && C % 8 == 0, "avgpool needs C % 8 == 0");
  const long long tv = static_cast<long long>(N) * HW * (C / 8);
  det::avgpool_bwd_kernel<<<det::blocks_for(tv), 256, 0, stream>>>((const bf16*)dy, (bf16*)dx, tv, HW, C);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_act_fwd(const void* x, void* y, long long n, int act, float slope, cudaStream_t stream) {
  DET_REQUIRE(act >= 1 && act <= 4, "activation must be 1 (relu), 2 (leaky relu), 3 (tanh) or 4 (sigmoid)");
  if (n <= 0) return 0;
  det::act_fwd_kernel<<<det::blocks_for((n + 7) / 8), 256, 0, stream>>>((const bf16*)x, (bf16*)y, n, act, slope);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_act_bwd(const void* dy, const void* y, void* dx, long long n, int act, float slope, cudaStream_t stream) {
  DET_REQUIRE(act >= 1 && act <= 4, "activation must be 1 (relu), 2 (leaky relu), 3 (tanh) or 4 (sigmoid)");
  if (n <= 0) return 0;
  det::act_bwd_kernel<<<det::blocks_for((n + 7) / 8), 256, 0, stream>>>((const bf16*)dy, (const bf16*)y, (bf16*)dx, n, act, slope);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_dropout(const void* x, void* y, long long n, float p, unsigned int seed, const unsigned int* seed_dev,
                           cudaStream_t stream) {
  DET_REQUIRE(p >= 0.0f && p < 1.0f, "dropout p must be in [0, 1)");
  DET_REQUIRE(n < (1LL << 32), "dropout tensors are limited to 2^32 elements");
  if (n <= 0) return 0;
  const uint32_t thr = static_cast<uint32_t>(static_cast<double>(p) * 4294967296.0);
  det::dropout_kernel<<<det::blocks_for((n + 7) / 8), 256, 0, stream>>>((const bf16*)x, (bf16*)y, n, thr, 1.0f / (1.0f - p), seed, seed_dev);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_nchw_to_nhwc_bf16(const float* in, void* out, int N, int C, int HW, cudaStream_t stream) {
  DET_REQUIRE(N > 0 && C > 0 && HW > 0 && N <= 65535, "bad tensor shape for the layout conversion");
  dim3 grid((HW + 31) / 32, (C + 31) / 32, N);
  det::nchw_to_nhwc_kernel<<<grid, 256, 0, stream>>>(in, (bf16*)out, C, HW);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// in: uint8 [n] laid out ...HWC with C channels (C <= 8); out: bf16 [n]; y = (x / 255 - mean[c]) / std[c].
extern "C" int det_u8_to_bf16_norm(const void* in, void* out, long long n, int C, const float* mean_host,
                                   const float* std_host, cudaStream_t stream) {
  DET_REQUIRE(n > 0 && C >= 1 && C <= 8, "u8 normalise: 1..8 channels");
  DET_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
              "u8 normalise: pointers must be 16-byte aligned");
  det::U8Norm p;
  p.C = C;
  for (int c = 0; c < 8; ++c) {
    const float sd = (c < C && std_host) ? std_host[c] : 1.0f;
    const float mu = (c < C && mean_host) ? mean_host[c] : 0.0f;
    p.scale[c] = 1.0f / (255.0f * sd);
    p.shift[c] = -mu / sd;
  }
  det::u8_to_bf16_norm_kernel<<<det::blocks_for((n + 15) / 16), 256, 0, stream>>>((const uint8_t*)in, (bf16*)out, n, p);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}
