// host_utils.h — host-side helpers shared by the kernel translation units: error reporting through
// the C ABI (every entry point returns 0 on success, non-zero otherwise, det_last_error() gives
// the text) and TMA tensor-map construction through the driver entry point (no -lcuda link).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstring>
#include <string>

namespace det {

void set_last_error(const std::string& s);  // defined in common.cu

#define DET_CHECK_CUDA(expr)                                                                   \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      ::det::set_last_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " @ " +      \
                            __FILE__ + ":" + std::to_string(__LINE__));                        \
      return 1;                                                                                \
    }                                                                                          \
  } while (0)

#define DET_REQUIRE(cond, msg)                                                                 \
  do {                                                                                         \
    if (!(cond)) {                                                                             \
      ::det::set_last_error(std::string("requirement failed: ") + #cond + " — " + (msg));     \
      return 2;                                                                                \
    }                                                                                          \
  } while (0)

// Encode a 2-D bf16 tensor map: global tensor [outer][inner] with row pitch `ld_elems`, box
// [box_outer][box_inner], SWIZZLE_128B (box_inner must be 64 bf16 = 128 B). Cached per key.
// Returns 0 on success.
int make_tma_2d_bf16(CUtensorMap* out, const void* gptr, uint64_t inner, uint64_t outer,
                     uint64_t ld_elems, uint32_t box_inner, uint32_t box_outer);

// Same for fp32 tensors (box_inner must be 32 floats = 128 B).
int make_tma_2d_f32(CUtensorMap* out, const void* gptr, uint64_t inner, uint64_t outer,
                    uint64_t ld_elems, uint32_t box_inner, uint32_t box_outer);

// Rank 2..4 variant (dims/box innermost first, strides in bytes for dims 1..rank-1); box[0] must be 64.
int make_tma_nd_bf16(CUtensorMap* out, const void* gptr, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box);

int sm_count();  // multiprocessor count of the current device (cached per device)

}  // namespace det
