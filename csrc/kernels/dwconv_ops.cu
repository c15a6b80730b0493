// dwconv_ops.cu — depthwise 2-D convolution (groups == channels) on NHWC bf16, with stride, padding and dilation.
//
// A depthwise convolution has no reduction over channels, so there is nothing for the tensor cores to do: per output
// element it is R*S multiply-adds, and the kernels are written for HBM bandwidth instead — a thread owns 8 consecutive
// channels (one 16-byte vector) of one pixel, the R*S taps of those channels come from a tap-major fp32 weight array
// [R*S, C] (two float4 loads per tap, L1 resident), activations are read once per tap from L1/L2.
//
//   fwd    y[n,ho,wo,c]  = bias[c] + sum_{r,s} x[n, ho*st - p + r*d, wo*st - p + s*d, c] * w[r,s,c]
//   dgrad  dx[n,hi,wi,c] = sum_{r,s : (hi + p - r*d) % st == 0 ...} dy[n,ho,wo,c] * w[r,s,c]
//   wgrad  dw[r,s,c]    += sum_{n,ho,wo} x[n,hi,wi,c] * dy[n,ho,wo,c]        (fp32 atomics into the gradient bucket)
//
// Together with the 1x1 GEMM this is the separable / dilated convolution of the reference's NAS examples
// (examples/hp_search_benchmarks/darts_cifar10_pytorch/operations.py: SepConv, DilConv; examples/nas/gaea_pytorch), which
// reach it through torch.nn.Conv2d(groups=C) → cuDNN.
#include "host_utils.h"
#include "sm100_ptx.cuh"

namespace det {

namespace {

struct DwParams {
  int N, H, W, C, R, S, stride, pad, dil, Ho, Wo;
};

__device__ __forceinline__ void dw_unpack8(const uint4& u, float* f) {
  f[0] = ptx::bf16_lo(u.x); f[1] = ptx::bf16_hi(u.x); f[2] = ptx::bf16_lo(u.y); f[3] = ptx::bf16_hi(u.y);
  f[4] = ptx::bf16_lo(u.z); f[5] = ptx::bf16_hi(u.z); f[6] = ptx::bf16_lo(u.w); f[7] = ptx::bf16_hi(u.w);
}
__device__ __forceinline__ uint4 dw_pack8(const float* f) {
  uint4 u;
  u.x = ptx::pack_bf16x2(f[0], f[1]); u.y = ptx::pack_bf16x2(f[2], f[3]);
  u.z = ptx::pack_bf16x2(f[4], f[5]); u.w = ptx::pack_bf16x2(f[6], f[7]);
  return u;
}
__device__ __forceinline__ void dw_load_w(const float* __restrict__ w, int tap, int C, int c0, float* out) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(w + static_cast<long long>(tap) * C + c0));
  const float4 b = __ldg(reinterpret_cast<const float4*>(w + static_cast<long long>(tap) * C + c0 + 4));
  out[0] = a.x; out[1] = a.y; out[2] = a.z; out[3] = a.w; out[4] = b.x; out[5] = b.y; out[6] = b.z; out[7] = b.w;
}

}  // namespace

__global__ void __launch_bounds__(256)
dwconv_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                  __nv_bfloat16* __restrict__ y, DwParams p, int relu) {
  const int cv = p.C / 8;
  const long long total = static_cast<long long>(p.N) * p.Ho * p.Wo * cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(i % cv) * 8;
    long long pix = i / cv;
    const int wo = static_cast<int>(pix % p.Wo); pix /= p.Wo;
    const int ho = static_cast<int>(pix % p.Ho);
    const int n = static_cast<int>(pix / p.Ho);
    float acc[8];
    if (bias != nullptr) dw_load_w(bias, 0, p.C, c0, acc);
    else {
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = 0.0f;
    }
    for (int r = 0; r < p.R; ++r) {
      const int hi = ho * p.stride - p.pad + r * p.dil;
      if (hi < 0 || hi >= p.H) continue;
      for (int s = 0; s < p.S; ++s) {
        const int wi = wo * p.stride - p.pad + s * p.dil;
        if (wi < 0 || wi >= p.W) continue;
        float xv[8], wv[8];
        dw_unpack8(__ldg(reinterpret_cast<const uint4*>(x + ((static_cast<long long>(n) * p.H + hi) * p.W + wi) * p.C + c0)), xv);
        dw_load_w(w, r * p.S + s, p.C, c0, wv);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(xv[e], wv[e], acc[e]);
      }
    }
    if (relu) {
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = fmaxf(acc[e], 0.0f);
    }
    *reinterpret_cast<uint4*>(y + ((static_cast<long long>(n) * p.Ho + ho) * p.Wo + wo) * p.C + c0) = dw_pack8(acc);
  }
}

__global__ void __launch_bounds__(256)
dwconv_dgrad_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ w, __nv_bfloat16* __restrict__ dx, DwParams p) {
  const int cv = p.C / 8;
  const long long total = static_cast<long long>(p.N) * p.H * p.W * cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(i % cv) * 8;
    long long pix = i / cv;
    const int wi = static_cast<int>(pix % p.W); pix /= p.W;
    const int hi = static_cast<int>(pix % p.H);
    const int n = static_cast<int>(pix / p.H);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int r = 0; r < p.R; ++r) {
      const int hn = hi + p.pad - r * p.dil;
      if (hn < 0 || hn % p.stride != 0) continue;
      const int ho = hn / p.stride;
      if (ho >= p.Ho) continue;
      for (int s = 0; s < p.S; ++s) {
        const int wn = wi + p.pad - s * p.dil;
        if (wn < 0 || wn % p.stride != 0) continue;
        const int wo = wn / p.stride;
        if (wo >= p.Wo) continue;
        float gv[8], wv[8];
        dw_unpack8(__ldg(reinterpret_cast<const uint4*>(dy + ((static_cast<long long>(n) * p.Ho + ho) * p.Wo + wo) * p.C + c0)), gv);
        dw_load_w(w, r * p.S + s, p.C, c0, wv);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(gv[e], wv[e], acc[e]);
      }
    }
    *reinterpret_cast<uint4*>(dx + ((static_cast<long long>(n) * p.H + hi) * p.W + wi) * p.C + c0) = dw_pack8(acc);
  }
}

// Block = DW_LANES pixel lanes x up to 32 channel vectors; every tap ends with one atomic per channel per block, so the
// grid is kept to a couple of blocks per SM (the lesson of the BatchNorm reductions).
constexpr int DW_WG_THREADS = 256;

__global__ void __launch_bounds__(DW_WG_THREADS)
dwconv_wgrad_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy, float* __restrict__ dw, DwParams p) {
  __shared__ float red[DW_WG_THREADS][8];
  const int cv_total = p.C / 8;
  const int vpr = min(cv_total - static_cast<int>(blockIdx.y) * 32, 32);      // channel vectors of this block column
  const int lanes = DW_WG_THREADS / vpr;
  const int v = threadIdx.x % vpr, lane = threadIdx.x / vpr;
  const bool active = lane < lanes;
  const int c0 = (blockIdx.y * 32 + v) * 8;
  const long long P = static_cast<long long>(p.N) * p.Ho * p.Wo;
  for (int r = 0; r < p.R; ++r) {
    for (int s = 0; s < p.S; ++s) {
      float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      if (active) {
        for (long long pix = static_cast<long long>(blockIdx.x) * lanes + lane; pix < P; pix += static_cast<long long>(gridDim.x) * lanes) {
          const int wo = static_cast<int>(pix % p.Wo);
          const int ho = static_cast<int>((pix / p.Wo) % p.Ho);
          const int n = static_cast<int>(pix / (static_cast<long long>(p.Wo) * p.Ho));
          const int hi = ho * p.stride - p.pad + r * p.dil, wi = wo * p.stride - p.pad + s * p.dil;
          if (hi < 0 || hi >= p.H || wi < 0 || wi >= p.W) continue;
          float xv[8], gv[8];
          dw_unpack8(__ldg(reinterpret_cast<const uint4*>(x + ((static_cast<long long>(n) * p.H + hi) * p.W + wi) * p.C + c0)), xv);
          dw_unpack8(__ldg(reinterpret_cast<const uint4*>(dy + pix * p.C + c0)), gv);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[e] = fmaf(xv[e], gv[e], acc[e]);
        }
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) red[threadIdx.x][e] = acc[e];
      __syncthreads();
      if (threadIdx.x < vpr * 8) {                    // thread t finishes channel (t / 8, t % 8) of this block column
        const int vv = threadIdx.x / 8, e = threadIdx.x % 8;
        float sum = 0.0f;
        for (int l = 0; l < lanes; ++l) sum += red[l * vpr + vv][e];
        atomicAdd(dw + static_cast<long long>(r * p.S + s) * p.C + (blockIdx.y * 32 + vv) * 8 + e, sum);
      }
      __syncthreads();
    }
  }
}

// ---- windowed average pooling: the same traversal with weights 1 / (window size); count_include_pad as in torch ----
__device__ __forceinline__ float avg_window_inv(const DwParams& p, int ho, int wo, int include_pad) {
  if (include_pad) return 1.0f / static_cast<float>(p.R * p.S);
  const int h0 = max(ho * p.stride - p.pad, 0), h1 = min(ho * p.stride - p.pad + p.R, p.H);
  const int w0 = max(wo * p.stride - p.pad, 0), w1 = min(wo * p.stride - p.pad + p.S, p.W);
  return 1.0f / static_cast<float>(max((h1 - h0) * (w1 - w0), 1));
}

__global__ void __launch_bounds__(256)
avgpool2d_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, DwParams p, int include_pad) {
  const int cv = p.C / 8;
  const long long total = static_cast<long long>(p.N) * p.Ho * p.Wo * cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(i % cv) * 8;
    long long pix = i / cv;
    const int wo = static_cast<int>(pix % p.Wo); pix /= p.Wo;
    const int ho = static_cast<int>(pix % p.Ho);
    const int n = static_cast<int>(pix / p.Ho);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int r = 0; r < p.R; ++r) {
      const int hi = ho * p.stride - p.pad + r;
      if (hi < 0 || hi >= p.H) continue;
      for (int s = 0; s < p.S; ++s) {
        const int wi = wo * p.stride - p.pad + s;
        if (wi < 0 || wi >= p.W) continue;
        float xv[8];
        dw_unpack8(__ldg(reinterpret_cast<const uint4*>(x + ((static_cast<long long>(n) * p.H + hi) * p.W + wi) * p.C + c0)), xv);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += xv[e];
      }
    }
    const float inv = avg_window_inv(p, ho, wo, include_pad);
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] *= inv;
    *reinterpret_cast<uint4*>(y + ((static_cast<long long>(n) * p.Ho + ho) * p.Wo + wo) * p.C + c0) = dw_pack8(acc);
  }
}

__global__ void __launch_bounds__(256)
avgpool2d_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, DwParams p, int include_pad) {
  const int cv = p.C / 8;
  const long long total = static_cast<long long>(p.N) * p.H * p.W * cv;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(i % cv) * 8;
    long long pix = i / cv;
    const int wi = static_cast<int>(pix % p.W); pix /= p.W;
    const int hi = static_cast<int>(pix % p.H);
    const int n = static_cast<int>(pix / p.H);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int r = 0; r < p.R; ++r) {
      const int hn = hi + p.pad - r;
      if (hn < 0 || hn % p.stride != 0) continue;
      const int ho = hn / p.stride;
      if (ho >= p.Ho) continue;
      for (int s = 0; s < p.S; ++s) {
        const int wn = wi + p.pad - s;
        if (wn < 0 || wn % p.stride != 0) continue;
        const int wo = wn / p.stride;
        if (wo >= p.Wo) continue;
        float gv[8];
        dw_unpack8(__ldg(reinterpret_cast<const uint4*>(dy + ((static_cast<long long>(n) * p.Ho + ho) * p.Wo + wo) * p.C + c0)), gv);
        const float inv = avg_window_inv(p, ho, wo, include_pad);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(gv[e], inv, acc[e]);
      }
    }
    *reinterpret_cast<uint4*>(dx + ((static_cast<long long>(n) * p.H + hi) * p.W + wi) * p.C + c0) = dw_pack8(acc);
  }
}

static int dw_fill(DwParams& p, int N, int H, int W, int C, int R, int S, int stride, int pad, int dil) {
  DET_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0, "depthwise conv: channels must be a positive multiple of 8");
  DET_REQUIRE(R > 0 && S > 0 && stride > 0 && dil > 0 && pad >= 0, "depthwise conv: bad filter geometry");
  p.N = N; p.H = H; p.W = W; p.C = C; p.R = R; p.S = S; p.stride = stride; p.pad = pad; p.dil = dil;
  p.Ho = (H + 2 * pad - dil * (R - 1) - 1) / stride + 1;
  p.Wo = (W + 2 * pad - dil * (S - 1) - 1) / stride + 1;
  DET_REQUIRE(p.Ho > 0 && p.Wo > 0, "depthwise conv: empty output");
  return 0;
}

}  // namespace det

// x [N,H,W,C] bf16; w fp32 tap-major [R*S, C]; bias fp32 [C] or null; y [N,Ho,Wo,C] bf16.
extern "C" int det_dwconv2d_fwd(const void* x, const float* w, const float* bias, void* y, int N, int H, int W, int C, int R, int S,
                                int stride, int pad, int dil, int relu, cudaStream_t stream) {
  using namespace det;
  DwParams p;
  if (int rc = dw_fill(p, N, H, W, C, R, S, stride, pad, dil)) return rc;
  const long long total = static_cast<long long>(N) * p.Ho * p.Wo * (C / 8);
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, static_cast<long long>(sm_count()) * 16));
  dwconv_fwd_kernel<<<blocks, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), w, bias, reinterpret_cast<__nv_bfloat16*>(y), p, relu);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// dy [N,Ho,Wo,C] bf16 -> dx [N,H,W,C] bf16 (H, W are the INPUT sizes).
extern "C" int det_dwconv2d_dgrad(const void* dy, const float* w, void* dx, int N, int H, int W, int C, int R, int S, int stride,
                                  int pad, int dil, cudaStream_t stream) {
  using namespace det;
  DwParams p;
  if (int rc = dw_fill(p, N, H, W, C, R, S, stride, pad, dil)) return rc;
  const long long total = static_cast<long long>(N) * H * W * (C / 8);
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, static_cast<long long>(sm_count()) * 16));
  dwconv_dgrad_kernel<<<blocks, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(dy), w, reinterpret_cast<__nv_bfloat16*>(dx), p);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// dw fp32 tap-major [R*S, C], accumulated into (+=).
extern "C" int det_dwconv2d_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int R, int S, int stride,
                                  int pad, int dil, cudaStream_t stream) {
  using namespace det;
  DwParams p;
  if (int rc = dw_fill(p, N, H, W, C, R, S, stride, pad, dil)) return rc;
  const int cols = (C / 8 + 31) / 32;
  const long long P = static_cast<long long>(N) * p.Ho * p.Wo;
  const int lanes = DW_WG_THREADS / std::min(C / 8, 32);
  const int rows = static_cast<int>(std::max<long long>(1, std::min<long long>((P + lanes * 8 - 1) / (lanes * 8), std::max(1, 2 * sm_count() / cols))));
  dwconv_wgrad_kernel<<<dim3(rows, cols), DW_WG_THREADS, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x),
                                                                      reinterpret_cast<const __nv_bfloat16*>(dy), dw, p);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Windowed average pooling (k x k window); bwd takes the INPUT sizes H, W.
extern "C" int det_avgpool2d_fwd(const void* x, void* y, int N, int H, int W, int C, int k, int stride, int pad, int count_include_pad,
                                 cudaStream_t stream) {
  using namespace det;
  DwParams p;
  if (int rc = dw_fill(p, N, H, W, C, k, k, stride, pad, 1)) return rc;
  const long long total = static_cast<long long>(N) * p.Ho * p.Wo * (C / 8);
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, static_cast<long long>(sm_count()) * 16));
  avgpool2d_fwd_kernel<<<blocks, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y), p, count_include_pad);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_avgpool2d_bwd(const void* dy, void* dx, int N, int H, int W, int C, int k, int stride, int pad, int count_include_pad,
                                 cudaStream_t stream) {
  using namespace det;
  DwParams p;
  if (int rc = dw_fill(p, N, H, W, C, k, k, stride, pad, 1)) return rc;
  const long long total = static_cast<long long>(N) * H * W * (C / 8);
  const int blocks = static_cast<int>(std::min<long long>((total + 255) / 256, static_cast<long long>(sm_count()) * 16));
  avgpool2d_bwd_kernel<<<blocks, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(dy), reinterpret_cast<__nv_bfloat16*>(dx), p, count_include_pad);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}
