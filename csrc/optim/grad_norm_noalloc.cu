// grad_norm_noalloc.cu — gradient L2 norm + non-finite check for callers that must not allocate.
//
// The regular path (det_grad_sumsq_finite in fused_optim.cu) accumulates into a sum-of-squares cell and
// an overflow flag that the CALLER allocates and clears. Some users of this library have no way to do
// that: a monitoring or health-check hook that is attached to a training process whose device memory is
// entirely owned by the framework's caching allocator (a cudaMalloc there can fail outright, or succeed
// only by forcing the allocator to give memory back in the middle of a step), and plain-C plugins that
// are handed gradient pointers inside the step loop, where cudaMalloc/cudaFree would also synchronise
// the whole device. For them this file offers ONE extra entry point, det_grad_norm_noalloc, whose
// scratch (per-block partial sums and the overflow flag) lives in a fixed array inside the library image.
//
// Properties to know before using it:
//   * Size limit: at most DET_NOALLOC_MAX_ELEMS (2^30) fp32 elements per call; larger inputs are rejected
//     with an error, nothing is launched.
//   * Memory cost: every GPU that loads the library carries the array (about 256 KB) until the process
//     exits, whether or not this entry point is ever called.
//   * Set-up: the first call on each device resolves the array's device address (with that device
//     current) and caches it. Make that first call before any stream capture begins.
//   * All calls on a device share the one array. It still holds the previous call's partial sums; the
//     entry point clears the flag cell it needs at zero and overwrites exactly the partials it reads.
//     Two calls that overlap on the GPU (different streams, no ordering between them) race on it.
//   * It calls no cudaMalloc/cudaFree and no allocating library routine.
//
// There is no counterpart of this entry point in the reference.
#include <cuda_runtime.h>

#include <mutex>

#include "host_utils.h"

namespace det {

constexpr int NA_THREADS = 256;
constexpr int NA_VEC_PER_THREAD = 16;                                   // float4 loads per thread
constexpr long long NA_BLOCK_ELEMS = 1LL * NA_THREADS * NA_VEC_PER_THREAD * 4;  // 16384 fp32 per block
constexpr int NA_MAX_BLOCKS = 65536;
constexpr long long NA_MAX_ELEMS = NA_BLOCK_ELEMS * NA_MAX_BLOCKS;      // 2^30
constexpr int NA_FLAG_SLOT = NA_MAX_BLOCKS;                             // one int cell after the partials

// Scratch for det_grad_norm_noalloc: [0, NA_MAX_BLOCKS) per-block partial sums, [NA_FLAG_SLOT] overflow flag.
// In host code this name is only a handle for cudaGetSymbolAddress, never a pointer to pass to a kernel.
static __device__ __align__(256) float g_noalloc_scratch[NA_MAX_BLOCKS + 64];

__global__ void __launch_bounds__(NA_THREADS)
noalloc_partial_sumsq_kernel(const float* __restrict__ g, long long n, const float* grad_scale,
                             float* __restrict__ scratch) {
  const float gs = grad_scale ? *grad_scale : 1.0f;
  const long long base = static_cast<long long>(blockIdx.x) * NA_BLOCK_ELEMS;
  const long long end = (base + NA_BLOCK_ELEMS < n) ? base + NA_BLOCK_ELEMS : n;
  float acc = 0.0f;
  bool bad = false;
  const bool aligned = (reinterpret_cast<uintptr_t>(g) & 15) == 0;
  if (aligned && end - base == NA_BLOCK_ELEMS) {
    const float4* p = reinterpret_cast<const float4*>(g + base);
#pragma unroll 4
    for (int i = 0; i < NA_VEC_PER_THREAD; ++i) {
      const float4 v = __ldg(p + i * NA_THREADS + threadIdx.x);
      const float a = v.x * gs, b = v.y * gs, c = v.z * gs, d = v.w * gs;
      acc += a * a + b * b + c * c + d * d;
      bad |= !(isfinite(a) && isfinite(b) && isfinite(c) && isfinite(d));
    }
  } else {
    for (long long i = base + threadIdx.x; i < end; i += NA_THREADS) {
      const float a = g[i] * gs;
      acc += a * a;
      bad |= !isfinite(a);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ float red[NA_THREADS / 32];
  __shared__ int sbad;
  if (threadIdx.x == 0) sbad = 0;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  if (bad) sbad = 1;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.0f;
#pragma unroll
    for (int w = 0; w < NA_THREADS / 32; ++w) s += red[w];
    scratch[blockIdx.x] = s;  // overwrite: stale partials of an earlier call never survive in [0, gridDim.x)
    if (sbad) atomicExch(reinterpret_cast<int*>(scratch + NA_FLAG_SLOT), 1);
  }
}

__global__ void __launch_bounds__(1024)
noalloc_final_kernel(const float* __restrict__ scratch, int nblocks, float* __restrict__ out2,
                     int* __restrict__ found_inf_out) {
  double acc = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += 1024) acc += static_cast<double>(scratch[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ double red[32];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < 32; ++w) s += red[w];
    out2[0] = static_cast<float>(sqrt(s));
    out2[1] = static_cast<float>(s);
    if (found_inf_out) *found_inf_out = *reinterpret_cast<const int*>(scratch + NA_FLAG_SLOT);
  }
}

// Device address of g_noalloc_scratch per device, resolved on first use with that device current.
static int noalloc_scratch_ptr(float** out) {
  static std::mutex mu;
  static float* ptrs[64] = {nullptr};
  int dev = 0;
  DET_CHECK_CUDA(cudaGetDevice(&dev));
  DET_REQUIRE(dev >= 0 && dev < 64, "device ordinal out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    float* p = nullptr;
    DET_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch));
    DET_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return 0;
}

}  // namespace det

extern "C" long long det_grad_norm_noalloc_max_elems() { return det::NA_MAX_ELEMS; }

// out2_dev[0] = || grad * (*grad_scale_dev) ||_2, out2_dev[1] = its square; *found_inf_out_dev = 1 when a
// non-finite value was seen, else 0 (may be null). grad_scale_dev may be null (scale 1). n <= 2^30.
// See the header of this file for the sharing, set-up and size rules.
extern "C" int det_grad_norm_noalloc(const float* grad, long long n, const float* grad_scale_dev,
                                     float* out2_dev, int* found_inf_out_dev, cudaStream_t stream) {
  using namespace det;
  DET_REQUIRE(grad != nullptr && out2_dev != nullptr, "det_grad_norm_noalloc: null pointer");
  DET_REQUIRE(n > 0, "det_grad_norm_noalloc: n must be positive");
  if (n > NA_MAX_ELEMS) {
    set_last_error("det_grad_norm_noalloc: n = " + std::to_string(n) + " exceeds the built-in scratch limit of " +
                   std::to_string(NA_MAX_ELEMS) + " elements; use det_grad_sumsq_finite with caller-owned scratch");
    return 4;
  }
  float* scratch = nullptr;
  int rc = noalloc_scratch_ptr(&scratch);
  if (rc) return rc;
  const int nblocks = static_cast<int>((n + NA_BLOCK_ELEMS - 1) / NA_BLOCK_ELEMS);
  DET_CHECK_CUDA(cudaMemsetAsync(scratch + NA_FLAG_SLOT, 0, sizeof(int), stream));  // the one cell that must start at zero
  noalloc_partial_sumsq_kernel<<<nblocks, NA_THREADS, 0, stream>>>(grad, n, grad_scale_dev, scratch);
  DET_CHECK_CUDA(cudaGetLastError());
  noalloc_final_kernel<<<1, 1024, 0, stream>>>(scratch, nblocks, out2_dev, found_inf_out_dev);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}
