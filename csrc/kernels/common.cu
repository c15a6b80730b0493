// common.cu — error string, device properties cache and TMA tensor-map cache for the kernel library.
#include <mutex>
#include <unordered_map>

#include "host_utils.h"

namespace det {

static thread_local std::string g_last_error;

void set_last_error(const std::string& s) { g_last_error = s; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

struct TmaKey {
  const void* p;
  uint64_t inner, outer, ld;
  uint32_t bi, bo;
  bool operator==(const TmaKey& o) const {
    return p == o.p && inner == o.inner && outer == o.outer && ld == o.ld && bi == o.bi &&
           bo == o.bo;
  }
};
struct TmaKeyHash {
  size_t operator()(const TmaKey& k) const {
    uint64_t h = reinterpret_cast<uint64_t>(k.p) * 0x9E3779B97F4A7C15ull;
    h ^= (k.inner + 0x9E37ull) * 0xBF58476D1CE4E5B9ull;
    h ^= (k.outer + 0x79B9ull) * 0x94D049BB133111EBull;
    h ^= k.ld * 31 + k.bi * 131 + k.bo * 1031;
    return static_cast<size_t>(h ^ (h >> 29));
  }
};

int make_tma_2d_bf16(CUtensorMap* out, const void* gptr, uint64_t inner, uint64_t outer,
                     uint64_t ld_elems, uint32_t box_inner, uint32_t box_outer) {
  static std::mutex mu;
  static std::unordered_map<TmaKey, CUtensorMap, TmaKeyHash> cache;
  TmaKey key{gptr, inner, outer, ld_elems, box_inner, box_outer};
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  EncodeTiledFn fn = get_encode_fn();
  DET_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  DET_REQUIRE((reinterpret_cast<uintptr_t>(gptr) & 15) == 0, "TMA base must be 16-byte aligned");
  DET_REQUIRE((ld_elems * 2) % 16 == 0, "TMA row pitch must be a multiple of 16 bytes");
  DET_REQUIRE(box_inner == 64, "SWIZZLE_128B box needs 64 bf16 in the inner dimension");
  cuuint64_t gdim[2] = {inner, outer};
  cuuint64_t gstr[1] = {ld_elems * 2};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(gptr), gdim, gstr,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    return 3;
  }
  std::lock_guard<std::mutex> g(mu);
  if (cache.size() > 65536) cache.clear();
  cache.emplace(key, *out);
  return 0;
}

// 2-D fp32 tensor map, SWIZZLE_128B (box_inner = 32 floats = 128 B): the fp32 store / reduce-add epilogues of the GEMM.
int make_tma_2d_f32(CUtensorMap* out, const void* gptr, uint64_t inner, uint64_t outer, uint64_t ld_elems,
                    uint32_t box_inner, uint32_t box_outer) {
  static std::mutex mu;
  static std::unordered_map<TmaKey, CUtensorMap, TmaKeyHash> cache;
  TmaKey key{gptr, inner, outer, ld_elems, box_inner, box_outer};
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return 0; }
  }
  EncodeTiledFn fn = get_encode_fn();
  DET_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  DET_REQUIRE((reinterpret_cast<uintptr_t>(gptr) & 15) == 0, "TMA base must be 16-byte aligned");
  DET_REQUIRE((ld_elems * 4) % 16 == 0, "TMA row pitch must be a multiple of 16 bytes");
  DET_REQUIRE(box_inner == 32, "SWIZZLE_128B box needs 32 fp32 in the inner dimension");
  cuuint64_t gdim[2] = {inner, outer};
  cuuint64_t gstr[1] = {ld_elems * 4};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(gptr), gdim, gstr, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled (fp32) failed with CUresult " + std::to_string((int)r));
    return 3;
  }
  std::lock_guard<std::mutex> g(mu);
  if (cache.size() > 65536) cache.clear();
  cache.emplace(key, *out);
  return 0;
}

// N-D (rank 2..4) bf16 tensor map, SWIZZLE_128B, innermost box = 64 elements. dims/box are innermost first;
// strides_bytes has rank-1 entries (dims 1..rank-1). Cached on every argument.
int make_tma_nd_bf16(CUtensorMap* out, const void* gptr, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box) {
  struct NdKey {
    const void* p; int rank; uint64_t d[4]; uint64_t s[3]; uint32_t b[4];
    bool operator==(const NdKey& o) const { return std::memcmp(this, &o, sizeof(NdKey)) == 0; }
  };
  struct NdHash {
    size_t operator()(const NdKey& k) const {
      const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
      uint64_t h = 0xcbf29ce484222325ull;
      for (size_t i = 0; i < sizeof(NdKey) / 8; ++i) { h ^= w[i]; h *= 0x100000001b3ull; h ^= h >> 31; }
      return static_cast<size_t>(h);
    }
  };
  static_assert(sizeof(NdKey) % 8 == 0, "key must be a whole number of words");
  static std::mutex mu;
  static std::unordered_map<NdKey, CUtensorMap, NdHash> cache;
  DET_REQUIRE(rank >= 2 && rank <= 4, "tensor map rank must be 2..4");
  NdKey key;
  std::memset(&key, 0, sizeof(key));
  key.p = gptr; key.rank = rank;
  for (int i = 0; i < rank; ++i) { key.d[i] = dims[i]; key.b[i] = box[i]; }
  for (int i = 0; i < rank - 1; ++i) key.s[i] = strides_bytes[i];
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return 0; }
  }
  EncodeTiledFn fn = get_encode_fn();
  DET_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  DET_REQUIRE((reinterpret_cast<uintptr_t>(gptr) & 15) == 0, "TMA base must be 16-byte aligned");
  DET_REQUIRE(box[0] == 64, "SWIZZLE_128B box needs 64 bf16 in the inner dimension");
  cuuint64_t gdim[4]; cuuint64_t gstr[3]; cuuint32_t bx[4]; cuuint32_t estr[4] = {1, 1, 1, 1};
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i]; bx[i] = box[i];
    DET_REQUIRE(box[i] >= 1 && box[i] <= 256, "TMA box extents must be 1..256");
  }
  for (int i = 0; i < rank - 1; ++i) {
    gstr[i] = strides_bytes[i];
    DET_REQUIRE(strides_bytes[i] % 16 == 0, "TMA strides must be multiples of 16 bytes");
  }
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(gptr), gdim, gstr,
                  bx, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled (rank " + std::to_string(rank) + ") failed with CUresult " +
                   std::to_string((int)r));
    return 3;
  }
  std::lock_guard<std::mutex> g(mu);
  if (cache.size() > 65536) cache.clear();
  cache.emplace(key, *out);
  return 0;
}

// SMs the persistent kernels may occupy (0 = all). Lowered by the host while a gradient all-reduce runs on a side
// stream, so that the collective's CTAs find free SMs instead of queueing behind full-width persistent grids.
static int g_sm_limit = 0;

int sm_count() {
  static int counts[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (counts[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    counts[dev] = n;
  }
  if (g_sm_limit > 0 && g_sm_limit < counts[dev]) return g_sm_limit;
  return counts[dev];
}

}  // namespace det

extern "C" const char* det_last_error() { return det::g_last_error.c_str(); }

extern "C" int det_kernels_abi_version() { return 1; }

// Limit the SM count every launcher in this library sizes its grid from (0 restores the device's count).
extern "C" int det_set_sm_limit(int n) {
  det::g_sm_limit = n < 0 ? 0 : n;
  return 0;
}
