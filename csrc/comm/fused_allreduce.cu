// fused_allreduce.cu — in-kernel collectives over NVLink peer memory for one 8xB200 NVSwitch box.
//
// Every rank owns a "symmetric arena": the same allocation on each GPU, mapped into every peer's
// address space (CUDA IPC or torch symmetric memory; see determined_b200/parallel/arena.py). The
// kernels below take the table of peer pointers and do the whole collective themselves:
//
//   det_fused_allreduce  — bucketed gradient all-reduce fused with averaging (÷world, ÷agg_freq),
//                          loss-scale unscale, non-finite detection and the grad-norm partial sum.
//                          One pass: rank r pulls slice r of every peer's bucket with 16-byte P2P
//                          loads, reduces in fp32 registers, and pushes the finished slice into every
//                          peer's bucket with P2P stores (reduce-scatter and all-gather in one sweep).
//                          Optional NVLS path: multimem.ld_reduce / multimem.st through a multicast
//                          address so the switch does the arithmetic.
//                          Replaces Horovod's NCCL ring allreduce + the per-parameter
//                          `_average_gradients` kernels + apex unscale
//                          (reference: harness/determined/pytorch/_pytorch_context.py:192-198, :462,
//                          :470-477, :522).
//   det_p2p_broadcast    — parameter / optimizer-state broadcast from rank 0 (reference:
//                          hvd.broadcast_parameters, _pytorch_trial.py:71-73): every rank pulls the
//                          root's bucket.
//   det_metric_allreduce — AVG/SUM/MIN/MAX of a small metric vector across ranks (reference: ZMQ +
//                          pickle gather, _pytorch_trial.py:292-335, :542-616): one-shot, every rank
//                          reads all peers.
//
// Cross-rank synchronisation is a flag barrier in peer memory: block b of rank r writes an epoch
// number into slot [b][r] of every peer's flag table and spins until its own row holds that epoch
// from everyone. Grids are identical on all ranks and never larger than the SM count, so every
// block is resident while it waits. Spins are bounded (trap after ~10 s) so a lost rank cannot
// wedge the GPU.
#include <cuda_bf16.h>

#include <algorithm>

#include "host_utils.h"
#include "sm100_ptx.cuh"

namespace det {

constexpr int COMM_MAX_RANKS = 8;
constexpr int COMM_MAX_BLOCKS = 256;

struct PeerTable {
  void* buf[COMM_MAX_RANKS];        // bucket base on each rank (peer-mapped addresses)
  uint32_t* flags[COMM_MAX_RANKS];  // flag table on each rank: [COMM_MAX_BLOCKS][COMM_MAX_RANKS]
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// All threads of the block call this. `epoch` must be the same on every rank for this barrier.
__device__ __forceinline__ void rank_barrier(const PeerTable& t, int rank, int world, uint32_t epoch) {
  __syncthreads();
  if (threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(t.flags[threadIdx.x] + blockIdx.x * COMM_MAX_RANKS + rank, epoch);
    const uint32_t* mine = t.flags[rank] + blockIdx.x * COMM_MAX_RANKS + threadIdx.x;
    const long long t0 = clock64();
    while (static_cast<int32_t>(ld_acquire_sys(mine) - epoch) < 0) {
      if (clock64() - t0 > 20000000000LL) __trap();
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float4 ld_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_f4(float* p, const float4& v) {
  asm volatile("st.global.relaxed.sys.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w) : "memory");
}

// Layout of the symmetric flag region (uint32 words): [COMM_MAX_BLOCKS][COMM_MAX_RANKS] barrier flags, then the
// gradient-statistics mailbox: SS_VALUE + r = rank r's sum of squares of its slice (+inf when it saw a non-finite
// value), SS_TAG + r = the launch number that value belongs to. The local epoch array carries, after its
// COMM_MAX_BLOCKS per-block counters: [SS_EPOCH] launch number, [SS_DONE] blocks finished, [SS_PART] running partial.
constexpr int SS_VALUE = COMM_MAX_BLOCKS * COMM_MAX_RANKS;
constexpr int SS_TAG = SS_VALUE + COMM_MAX_RANKS;
// TURN + c * COMM_MAX_RANKS + r: 1 once rank r's share of chunk c of the current step's pipelined all-gather has landed
// in THIS rank's bucket
constexpr int TURN = SS_TAG + COMM_MAX_RANKS;
constexpr int FLAG_WORDS = TURN + 8 * COMM_MAX_RANKS;   // 8 = PIPE_MAX_CHUNKS
constexpr int SS_EPOCH = COMM_MAX_BLOCKS, SS_DONE = COMM_MAX_BLOCKS + 1, SS_PART = COMM_MAX_BLOCKS + 2, AG_DONE = COMM_MAX_BLOCKS + 3;
constexpr int EPOCH_WORDS = COMM_MAX_BLOCKS + 4;

struct AllreduceArgs {
  PeerTable t;
  float* mc_ptr;          // multicast (NVLS) address of the bucket, or null
  uint32_t* epoch;        // local device counter [COMM_MAX_BLOCKS]
  long long n;            // fp32 elements in the bucket (multiple of 4)
  int rank, world;
  const float* scale_dev; // optional device scalar multiplied into `scale`
  float scale;
  float* sumsq;           // optional: += sum of squares of the WHOLE reduced, scaled bucket (all ranks' slices)
  int* found_inf;         // optional: set to 1 when ANY rank saw a non-finite value in its slice
};

// Gradient statistics across ranks without a second launch: every rank reduces only its own slice, so the sum of squares
// (grad-norm clipping) and the non-finite flag (fp16 loss scaling) of the whole bucket exist only as `world` partials. The
// last block of each rank to finish publishes the rank's partial into every peer's mailbox, waits for the other ranks'
// partials of the same launch and adds them up in rank order — every rank computes bit-identical totals, so every rank
// derives the same clip coefficient and takes the same skip decision.
__device__ __forceinline__ void exchange_grad_stats(const AllreduceArgs& a, float block_sum, bool block_bad) {
  // called by thread 0 of every block
  float* part = reinterpret_cast<float*>(a.epoch + SS_PART);
  atomicAdd(part, block_bad ? INFINITY : block_sum);
  __threadfence();
  const uint32_t done = atomicAdd(a.epoch + SS_DONE, 1u);
  if (done != gridDim.x - 1) return;
  __threadfence();
  const float mine = atomicExch(part, 0.0f);
  a.epoch[SS_DONE] = 0;
  const uint32_t tag = a.epoch[SS_EPOCH] + 1;
  a.epoch[SS_EPOCH] = tag;
  for (int r = 0; r < a.world; ++r) {
    uint32_t* f = a.t.flags[r];
    asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(reinterpret_cast<float*>(f + SS_VALUE + a.rank)), "f"(mine) : "memory");
    __threadfence_system();
    st_release_sys(f + SS_TAG + a.rank, tag);
  }
  float total = 0.0f;
  const long long t0 = clock64();
  for (int r = 0; r < a.world; ++r) {
    const uint32_t* f = a.t.flags[a.rank];
    while (static_cast<int32_t>(ld_acquire_sys(f + SS_TAG + r) - tag) < 0) {
      if (clock64() - t0 > 20000000000LL) __trap();
    }
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(reinterpret_cast<const float*>(f + SS_VALUE + r)) : "memory");
    total += v;
  }
  if (a.sumsq != nullptr && isfinite(total)) atomicAdd(a.sumsq, total);
  if (a.found_inf != nullptr && !isfinite(total)) atomicExch(a.found_inf, 1);
}

// Memory-level parallelism is what sets the bandwidth of one CTA here (a remote load / switch reduction takes a few
// microseconds): every thread keeps AR_UNROLL independent 16-byte transactions in flight, so a handful of CTAs already
// saturates the links — which is what lets the collective run next to backward on a few reserved SMs.
constexpr int AR_THREADS = 1024;
constexpr int AR_UNROLL_MC = 8;
constexpr int AR_UNROLL_P2P = 1;   // the peer loads of one element are already `world` independent transactions

template <bool USE_MC>
__global__ void __launch_bounds__(AR_THREADS)
fused_allreduce_kernel(const AllreduceArgs a) {
  constexpr int U = USE_MC ? AR_UNROLL_MC : AR_UNROLL_P2P;
  const int rank = a.rank, world = a.world;
  uint32_t ep = a.epoch[blockIdx.x];
  rank_barrier(a.t, rank, world, ep + 1);  // every rank's gradients for this bucket are complete

  const float scale = a.scale * (a.scale_dev ? *a.scale_dev : 1.0f);
  // slice owned by this rank, in float4 units
  const long long n4 = a.n >> 2;
  const long long per = (n4 + world - 1) / world;
  const long long lo = std::min<long long>(per * rank, n4), hi = std::min<long long>(lo + per, n4);
  float acc = 0.0f;
  bool bad = false;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * U;
  for (long long i0 = lo + static_cast<long long>(blockIdx.x) * blockDim.x * U + threadIdx.x; i0 < hi; i0 += stride) {
    float4 s[U];
    if (USE_MC) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + static_cast<long long>(u) * blockDim.x;
        if (i < hi)
          asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(s[u].x), "=f"(s[u].y), "=f"(s[u].z), "=f"(s[u].w) : "l"(a.mc_ptr + i * 4) : "memory");
      }
    } else {
      float4 v[U][COMM_MAX_RANKS];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + static_cast<long long>(u) * blockDim.x;
        if (i < hi) {
#pragma unroll
          for (int r = 0; r < COMM_MAX_RANKS; ++r)
            if (r < world) v[u][r] = ld_f4(reinterpret_cast<const float*>(a.t.buf[(rank + r) % world]) + i * 4);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        s[u] = v[u][0];
#pragma unroll
        for (int r = 1; r < COMM_MAX_RANKS; ++r)
          if (r < world) { s[u].x += v[u][r].x; s[u].y += v[u][r].y; s[u].z += v[u][r].z; s[u].w += v[u][r].w; }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + static_cast<long long>(u) * blockDim.x;
      if (i >= hi) break;
      float4 t = s[u];
      t.x *= scale; t.y *= scale; t.z *= scale; t.w *= scale;
      acc += t.x * t.x + t.y * t.y + t.z * t.z + t.w * t.w;
      bad |= !(isfinite(t.x) && isfinite(t.y) && isfinite(t.z) && isfinite(t.w));
      if (USE_MC) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a.mc_ptr + i * 4),
                     "f"(t.x), "f"(t.y), "f"(t.z), "f"(t.w) : "memory");
      } else {
#pragma unroll
        for (int r = 0; r < COMM_MAX_RANKS; ++r)
          if (r < world) st_f4(reinterpret_cast<float*>(a.t.buf[(rank + r) % world]) + i * 4, t);
      }
    }
  }
  if (a.sumsq != nullptr || a.found_inf != nullptr) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ float red[AR_THREADS / 32];
    __shared__ int sbad;
    if (threadIdx.x == 0) sbad = 0;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    if (bad) sbad = 1;
    __syncthreads();
    if (threadIdx.x == 0) {
      float s = 0.0f;
      for (int w = 0; w < (blockDim.x >> 5); ++w) s += red[w];
      exchange_grad_stats(a, s, sbad != 0);
    }
  }
  rank_barrier(a.t, rank, world, ep + 2);  // every slice has landed everywhere
  if (threadIdx.x == 0) a.epoch[blockIdx.x] = ep + 2;
}

// ------------------------------------------------------------------------------------------------------------------
// Pipelined form: reduce-scatter now, all-gather chunk by chunk UNDER the optimizer.
//
// A fused all-reduce ends before the optimizer starts; with grad-norm clipping nothing can be updated until the whole
// bucket has been reduced (the norm needs every element). But only the reduce-scatter half feeds the norm. So the bucket
// is cut into `chunks` contiguous chunks and every chunk into `world` sub-slices, rank r owning sub-slice r of each chunk:
//   1. reduce_scatter_stats_kernel   every rank reduces ITS sub-slices (switch reduction or peer loads), applies the scale,
//                                    accumulates Σg² / non-finite, keeps the result in its OWN bucket, and the per-rank
//                                    statistics are exchanged in-kernel — after it every rank knows the clip coefficient;
//   2. allgather_send_kernel(c)      every rank multicasts its sub-slice of chunk c into every rank's bucket (all ranks at
//                                    once: every link busy, as in a plain all-gather) and then raises TURN[c][rank] on every
//                                    rank; the launches for c = 0, 1, ... follow each other on a side stream;
//   3. the optimizer kernel for chunk c (one contiguous range) starts on every rank as soon as all of TURN[c][*] are up
//      there — while chunk c+1 is on the wire. The wire (NVLink in, little HBM) and the optimizer (HBM-bound) use
//      different resources.
// Waiting is done by wait_chunk_kernel (one warp) placed in the stream BEFORE the kernel that needs the data, never by the
// working kernels themselves: the only kernels that ever spin hold one warp and wait for OTHER GPUs' send kernels, which
// wait for nothing — no two kernels of one GPU wait for each other and nothing can deadlock for SM space.
// ------------------------------------------------------------------------------------------------------------------
constexpr int PIPE_MAX_CHUNKS = 8;

struct PipeGeom {
  long long n4, chunk4, sub4;   // float4 units: bucket, one chunk (= world * sub4), one rank's share of a chunk
  int chunks;
};
__host__ __device__ inline PipeGeom pipe_geom(long long n_f32, int world, int chunks) {
  PipeGeom g;
  g.n4 = n_f32 >> 2;
  g.chunks = chunks;
  const long long c4 = (g.n4 + chunks - 1) / chunks;
  g.sub4 = (c4 + world - 1) / world;
  g.chunk4 = g.sub4 * world;
  return g;
}

template <bool USE_MC>
__global__ void __launch_bounds__(AR_THREADS)
reduce_scatter_stats_kernel(const AllreduceArgs a, const int chunks) {
  constexpr int U = USE_MC ? AR_UNROLL_MC : AR_UNROLL_P2P;
  const int rank = a.rank, world = a.world;
  uint32_t ep = a.epoch[blockIdx.x];
  if (blockIdx.x == 0 && threadIdx.x < PIPE_MAX_CHUNKS * COMM_MAX_RANKS)   // last step's turns were all consumed
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(a.t.flags[rank] + TURN + threadIdx.x), "r"(0u) : "memory");
  rank_barrier(a.t, rank, world, ep + 1);   // gradients complete everywhere; nobody is still gathering the last step
  const float scale = a.scale * (a.scale_dev ? *a.scale_dev : 1.0f);
  const PipeGeom g = pipe_geom(a.n, world, chunks);
  float* mine = reinterpret_cast<float*>(a.t.buf[rank]);
  float acc = 0.0f;
  bool bad = false;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * U;
  for (int c = 0; c < chunks; ++c) {
    const long long lo = std::min<long long>(c * g.chunk4 + rank * g.sub4, g.n4), hi = std::min<long long>(lo + g.sub4, g.n4);
    for (long long i0 = lo + static_cast<long long>(blockIdx.x) * blockDim.x * U + threadIdx.x; i0 < hi; i0 += stride) {
      float4 s[U];
      if (USE_MC) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long i = i0 + static_cast<long long>(u) * blockDim.x;
          if (i < hi)
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(s[u].x), "=f"(s[u].y), "=f"(s[u].z), "=f"(s[u].w) : "l"(a.mc_ptr + i * 4) : "memory");
        }
      } else {
        float4 v[U][COMM_MAX_RANKS];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long i = i0 + static_cast<long long>(u) * blockDim.x;
          if (i < hi) {
#pragma unroll
            for (int r = 0; r < COMM_MAX_RANKS; ++r)
              if (r < world) v[u][r] = ld_f4(reinterpret_cast<const float*>(a.t.buf[(rank + r) % world]) + i * 4);
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          s[u] = v[u][0];
#pragma unroll
          for (int r = 1; r < COMM_MAX_RANKS; ++r)
            if (r < world) { s[u].x += v[u][r].x; s[u].y += v[u][r].y; s[u].z += v[u][r].z; s[u].w += v[u][r].w; }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + static_cast<long long>(u) * blockDim.x;
        if (i >= hi) break;
        float4 t = s[u];
        t.x *= scale; t.y *= scale; t.z *= scale; t.w *= scale;
        acc += t.x * t.x + t.y * t.y + t.z * t.z + t.w * t.w;
        bad |= !(isfinite(t.x) && isfinite(t.y) && isfinite(t.z) && isfinite(t.w));
        reinterpret_cast<float4*>(mine)[i] = t;
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ float red[AR_THREADS / 32];
  __shared__ int sbad;
  if (threadIdx.x == 0) sbad = 0;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  if (bad) sbad = 1;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.0f;
    for (int w = 0; w < (blockDim.x >> 5); ++w) t += red[w];
    exchange_grad_stats(a, t, sbad != 0);
    a.epoch[blockIdx.x] = ep + 1;
  }
}

template <bool USE_MC>
__global__ void __launch_bounds__(AR_THREADS)
allgather_send_kernel(const AllreduceArgs a, const int chunks, const int c) {
  constexpr int U = 8;
  const int rank = a.rank, world = a.world;
  const PipeGeom g = pipe_geom(a.n, world, chunks);
  const long long lo = std::min<long long>(c * g.chunk4 + rank * g.sub4, g.n4), hi = std::min<long long>(lo + g.sub4, g.n4);
  const float4* mine = reinterpret_cast<const float4*>(a.t.buf[rank]);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * U;
  for (long long i0 = lo + static_cast<long long>(blockIdx.x) * blockDim.x * U + threadIdx.x; i0 < hi; i0 += stride) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + static_cast<long long>(u) * blockDim.x;
      if (i < hi) v[u] = mine[i];
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + static_cast<long long>(u) * blockDim.x;
      if (i >= hi) break;
      if (USE_MC) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a.mc_ptr + i * 4),
                     "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w) : "memory");
      } else {
#pragma unroll
        for (int r = 1; r < COMM_MAX_RANKS; ++r)
          if (r < world) st_f4(reinterpret_cast<float*>(a.t.buf[(rank + r) % world]) + i * 4, v[u]);
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t done = atomicAdd(a.epoch + AG_DONE, 1u);
    if (done == gridDim.x - 1) {            // last block: everything this rank sent of chunk c is ordered before the flags
      __threadfence();
      a.epoch[AG_DONE] = 0;
      __threadfence_system();
      for (int r = 0; r < world; ++r)
        if (r != rank) st_release_sys(a.t.flags[r] + TURN + c * COMM_MAX_RANKS + rank, 1u);
    }
  }
}

// lanes 0 .. world-1 (except this rank's own) each wait for one sender's flag of chunk c
__global__ void __launch_bounds__(32)
wait_chunk_kernel(const uint32_t* flags, const int c, const int rank, const int world) {
  if (static_cast<int>(threadIdx.x) < world && static_cast<int>(threadIdx.x) != rank) {
    const uint32_t* f = flags + TURN + c * COMM_MAX_RANKS + threadIdx.x;
    const long long t0 = clock64();
    while (ld_acquire_sys(f) == 0u) {
      if (clock64() - t0 > 20000000000LL) __trap();
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------
// bf16 on the wire (`optimizations.gradient_compression`; the reference's hvd.Compression.fp16,
// _pytorch_context.py:191-198). The fp32 bucket is packed into a bf16 staging buffer of the symmetric arena (local pass),
// the collective runs on the staging buffers — the switch adds bf16 pairs with fp32 accumulation
// (multimem.ld_reduce ... .acc::f32 ... .bf16x2), or peers are read with 16-byte loads and summed in fp32 registers —
// scale / unscale / non-finite test / grad-norm partial happen on the fp32 values exactly as in the fp32 kernel, the
// result goes back out as bf16 (multimem.st / peer stores) and is unpacked into the fp32 bucket (local pass). Half the
// bytes cross NVLink; bf16 keeps fp32's exponent range, so no overflow handling is needed.
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack2_bf16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float bf16lo_f32(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16hi_f32(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

__global__ void __launch_bounds__(256)
wire_pack_kernel(const float* __restrict__ g, uint4* __restrict__ wire, long long n8) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 a = reinterpret_cast<const float4*>(g)[2 * i], b = reinterpret_cast<const float4*>(g)[2 * i + 1];
    wire[i] = make_uint4(pack2_bf16(a.x, a.y), pack2_bf16(a.z, a.w), pack2_bf16(b.x, b.y), pack2_bf16(b.z, b.w));
  }
}
__global__ void __launch_bounds__(256)
wire_unpack_kernel(const uint4* __restrict__ wire, float* __restrict__ g, long long n8) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const uint4 w = wire[i];
    reinterpret_cast<float4*>(g)[2 * i] = make_float4(bf16lo_f32(w.x), bf16hi_f32(w.x), bf16lo_f32(w.y), bf16hi_f32(w.y));
    reinterpret_cast<float4*>(g)[2 * i + 1] = make_float4(bf16lo_f32(w.z), bf16hi_f32(w.z), bf16lo_f32(w.w), bf16hi_f32(w.w));
  }
}

// a.t.buf / a.mc_ptr address the bf16 staging buffers here; a.n counts bf16 elements (multiple of 8).
template <bool USE_MC>
__global__ void __launch_bounds__(AR_THREADS)
fused_allreduce_bf16wire_kernel(const AllreduceArgs a) {
  constexpr int U = USE_MC ? AR_UNROLL_MC : 1;
  const int rank = a.rank, world = a.world;
  uint32_t ep = a.epoch[blockIdx.x];
  rank_barrier(a.t, rank, world, ep + 1);   // every rank has packed its bucket
  const float scale = a.scale * (a.scale_dev ? *a.scale_dev : 1.0f);
  const long long n8 = a.n >> 3;            // 16-byte units of eight bf16
  const long long per = (n8 + world - 1) / world;
  const long long lo = std::min<long long>(per * rank, n8), hi = std::min<long long>(lo + per, n8);
  const char* mc = reinterpret_cast<const char*>(a.mc_ptr);
  float acc = 0.0f;
  bool bad = false;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * U;
  for (long long i0 = lo + static_cast<long long>(blockIdx.x) * blockDim.x * U + threadIdx.x; i0 < hi; i0 += stride) {
    float f[U][8];
    if (USE_MC) {
      uint4 w[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + static_cast<long long>(u) * blockDim.x;
        if (i < hi)
          asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                       : "=r"(w[u].x), "=r"(w[u].y), "=r"(w[u].z), "=r"(w[u].w) : "l"(mc + i * 16) : "memory");
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        f[u][0] = bf16lo_f32(w[u].x); f[u][1] = bf16hi_f32(w[u].x); f[u][2] = bf16lo_f32(w[u].y); f[u][3] = bf16hi_f32(w[u].y);
        f[u][4] = bf16lo_f32(w[u].z); f[u][5] = bf16hi_f32(w[u].z); f[u][6] = bf16lo_f32(w[u].w); f[u][7] = bf16hi_f32(w[u].w);
      }
    } else {
      uint4 v[COMM_MAX_RANKS];
#pragma unroll
      for (int r = 0; r < COMM_MAX_RANKS; ++r)
        if (r < world) {
          const char* src = reinterpret_cast<const char*>(a.t.buf[(rank + r) % world]) + i0 * 16;
          asm volatile("ld.global.relaxed.sys.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v[r].x), "=r"(v[r].y), "=r"(v[r].z), "=r"(v[r].w) : "l"(src) : "memory");
        }
#pragma unroll
      for (int e = 0; e < 8; ++e) f[0][e] = 0.0f;
      // fixed order (rank 0 .. world-1) so that every rank would compute the same sum for the same element
#pragma unroll
      for (int r0 = 0; r0 < COMM_MAX_RANKS; ++r0)
        if (r0 < world) {
          const uint4 x = v[(r0 - rank + world) % world];
          f[0][0] += bf16lo_f32(x.x); f[0][1] += bf16hi_f32(x.x); f[0][2] += bf16lo_f32(x.y); f[0][3] += bf16hi_f32(x.y);
          f[0][4] += bf16lo_f32(x.z); f[0][5] += bf16hi_f32(x.z); f[0][6] += bf16lo_f32(x.w); f[0][7] += bf16hi_f32(x.w);
        }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + static_cast<long long>(u) * blockDim.x;
      if (i >= hi) break;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        f[u][e] *= scale;
        acc += f[u][e] * f[u][e];
        bad |= !isfinite(f[u][e]);
      }
      const uint4 o = make_uint4(pack2_bf16(f[u][0], f[u][1]), pack2_bf16(f[u][2], f[u][3]), pack2_bf16(f[u][4], f[u][5]), pack2_bf16(f[u][6], f[u][7]));
      if (USE_MC) {
        asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(mc + i * 16), "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
      } else {
#pragma unroll
        for (int r = 0; r < COMM_MAX_RANKS; ++r)
          if (r < world) {
            char* dst = reinterpret_cast<char*>(a.t.buf[(rank + r) % world]) + i * 16;
            asm volatile("st.global.relaxed.sys.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
          }
      }
    }
  }
  if (a.sumsq != nullptr || a.found_inf != nullptr) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ float red[AR_THREADS / 32];
    __shared__ int sbad;
    if (threadIdx.x == 0) sbad = 0;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    if (bad) sbad = 1;
    __syncthreads();
    if (threadIdx.x == 0) {
      float s = 0.0f;
      for (int w = 0; w < (blockDim.x >> 5); ++w) s += red[w];
      exchange_grad_stats(a, s, sbad != 0);
    }
  }
  rank_barrier(a.t, rank, world, ep + 2);
  if (threadIdx.x == 0) a.epoch[blockIdx.x] = ep + 2;
}

// ------------------------------------------------------------------------------------------------------------------
// TMA-driven all-reduce. The LSU kernel above keeps (threads x unroll) 16-byte requests in flight; a load over NVLink
// takes microseconds, so it needs thousands of resident threads to fill the links. Here ONE thread per CTA moves the
// data: for every chunk of this rank's slice it issues `world` bulk loads (cp.async.bulk global -> shared, one per
// rank's copy of the chunk, unicast, peer memory included), the CTA's threads add the copies up in shared memory (fixed
// rank order: deterministic sums), apply scale / unscale, look for non-finite values and accumulate the grad-norm
// partial, and the same thread pushes the finished chunk to every rank's bucket with bulk stores. Three 64 KB stages
// per CTA keep ~190 KB in flight per SM, so a dozen CTAs saturate the links. Per GPU the wire carries (world-1)/world
// of the bucket in each direction — the minimum for an all-reduce — where the multimem path moves the whole bucket out
// and 1 + 1/world of it in.
// ------------------------------------------------------------------------------------------------------------------
constexpr int BK_THREADS = 256;
constexpr int BK_STAGES = 3;
constexpr int BK_STAGE_BYTES = 65536;
constexpr int BK_SMEM_BYTES = BK_STAGES * BK_STAGE_BYTES;

__global__ void __launch_bounds__(BK_THREADS)
fused_allreduce_bulk_kernel(const AllreduceArgs a, const int ch_bytes) {
  extern __shared__ __align__(128) uint8_t bk_smem[];
  __shared__ uint64_t full[BK_STAGES];
  const int rank = a.rank, world = a.world;
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < BK_STAGES; ++s) ptx::mbar_init(ptx::smem_u32(&full[s]), 1);
    ptx::fence_mbar_init();
  }
  uint32_t ep = a.epoch[blockIdx.x];
  rank_barrier(a.t, rank, world, ep + 1);  // every rank's gradients for this bucket are complete (also orders the init)

  const float scale = a.scale * (a.scale_dev ? *a.scale_dev : 1.0f);
  const long long n4 = a.n >> 2;
  const long long per = (n4 + world - 1) / world;
  const long long lo = std::min<long long>(per * rank, n4), hi = std::min<long long>(lo + per, n4);
  const long long slice_bytes = (hi - lo) * 16;
  const long long base_bytes = lo * 16;
  const long long num_chunks = (slice_bytes + ch_bytes - 1) / ch_bytes;
  float acc = 0.0f;
  bool bad = false;

  auto issue_loads = [&](long long c, int stage) {
    const long long off = c * ch_bytes;
    const uint32_t bytes = static_cast<uint32_t>(std::min<long long>(ch_bytes, slice_bytes - off));
    const uint32_t bar = ptx::smem_u32(&full[stage]);
    ptx::mbar_expect_tx(bar, bytes * world);
    for (int r = 0; r < world; ++r)
      ptx::bulk_load(ptx::smem_u32(bk_smem + stage * BK_STAGE_BYTES + r * ch_bytes),
                     reinterpret_cast<const uint8_t*>(a.t.buf[r]) + base_bytes + off, bytes, bar);
  };

  // prologue: STAGES - 1 chunks in flight
  if (tid == 0) {
    ptx::fence_proxy_async_all();   // the barrier's acquire was a generic-proxy read; the bulk loads are async-proxy reads
    for (int j = 0; j < BK_STAGES - 1; ++j) {
      const long long c = blockIdx.x + static_cast<long long>(j) * gridDim.x;
      if (c < num_chunks) issue_loads(c, j);
    }
  }
  uint32_t it = 0;
  for (long long c = blockIdx.x; c < num_chunks; c += gridDim.x, ++it) {
    const int stage = it % BK_STAGES;
    if (tid == 0) {
      // the stage that iteration it-1 used is free once its stores have read it; refill it with chunk it + STAGES - 1
      const long long cn = c + static_cast<long long>(BK_STAGES - 1) * gridDim.x;
      if (cn < num_chunks) {
        ptx::bulk_wait_group_read0();
        issue_loads(cn, (it + BK_STAGES - 1) % BK_STAGES);
      }
    }
    ptx::mbar_wait(ptx::smem_u32(&full[stage]), (it / BK_STAGES) & 1);
    const long long off = c * ch_bytes;
    const int bytes = static_cast<int>(std::min<long long>(ch_bytes, slice_bytes - off));
    uint8_t* sbase = bk_smem + stage * BK_STAGE_BYTES;
    for (int i = tid * 16; i < bytes; i += BK_THREADS * 16) {
      float4 s = *reinterpret_cast<const float4*>(sbase + i);
      for (int r = 1; r < world; ++r) {
        const float4 v = *reinterpret_cast<const float4*>(sbase + r * ch_bytes + i);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      s.x *= scale; s.y *= scale; s.z *= scale; s.w *= scale;
      acc += s.x * s.x + s.y * s.y + s.z * s.z + s.w * s.w;
      bad |= !(isfinite(s.x) && isfinite(s.y) && isfinite(s.z) && isfinite(s.w));
      *reinterpret_cast<float4*>(sbase + i) = s;
    }
    ptx::fence_proxy_async_smem();
    __syncthreads();
    if (tid == 0) {
      for (int r = 0; r < world; ++r)
        ptx::bulk_store(reinterpret_cast<uint8_t*>(a.t.buf[(rank + r) % world]) + base_bytes + off, ptx::smem_u32(sbase),
                        static_cast<uint32_t>(bytes));
      ptx::bulk_commit_group();
    }
  }
  if (tid == 0) {
    ptx::bulk_wait_group0();       // every pushed chunk has been written
    ptx::fence_proxy_async_all();  // ... and is ordered before the flag stores of the closing barrier
  }
  if (a.sumsq != nullptr || a.found_inf != nullptr) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ float red[BK_THREADS / 32];
    __shared__ int sbad;
    if (tid == 0) sbad = 0;
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = acc;
    if (bad) sbad = 1;
    __syncthreads();
    if (tid == 0) {
      float s = 0.0f;
      for (int w = 0; w < BK_THREADS / 32; ++w) s += red[w];
      exchange_grad_stats(a, s, sbad != 0);
    }
  }
  rank_barrier(a.t, rank, world, ep + 2);  // every slice has landed everywhere
  if (tid == 0) a.epoch[blockIdx.x] = ep + 2;
}

struct BcastArgs {
  PeerTable t;
  uint32_t* epoch;
  long long nbytes;  // multiple of 16
  int rank, world, root;
};

__global__ void __launch_bounds__(512)
p2p_broadcast_kernel(const BcastArgs a) {
  uint32_t ep = a.epoch[blockIdx.x];
  rank_barrier(a.t, a.rank, a.world, ep + 1);  // root's data is final, receivers' buffers are idle
  if (a.rank != a.root) {
    const uint4* src = reinterpret_cast<const uint4*>(a.t.buf[a.root]);
    uint4* dst = reinterpret_cast<uint4*>(a.t.buf[a.rank]);
    const long long n16 = a.nbytes >> 4;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n16; i += stride) {
      uint4 v;
      asm volatile("ld.global.relaxed.sys.v4.u32 {%0, %1, %2, %3}, [%4];"
                   : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(src + i) : "memory");
      dst[i] = v;
    }
  }
  rank_barrier(a.t, a.rank, a.world, ep + 2);  // root may overwrite its buffer again
  if (threadIdx.x == 0) a.epoch[blockIdx.x] = ep + 2;
}

struct MetricArgs {
  PeerTable t;
  uint32_t* epoch;
  float* out;            // local result [n]
  const float* weights;  // optional per-rank weights [world] for weighted AVG (batches per process)
  int n, rank, world, op;  // op: 0 AVG, 1 SUM, 2 MAX, 3 MIN
};

__global__ void __launch_bounds__(256)
metric_allreduce_kernel(const MetricArgs a) {
  uint32_t ep = a.epoch[blockIdx.x];
  rank_barrier(a.t, a.rank, a.world, ep + 1);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += gridDim.x * blockDim.x) {
    float acc = (a.op == 2) ? -INFINITY : (a.op == 3 ? INFINITY : 0.0f);
    float wsum = 0.0f;
    for (int r = 0; r < a.world; ++r) {
      float v;
      asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(v) : "l"(reinterpret_cast<const float*>(a.t.buf[r]) + i) : "memory");
      const float w = a.weights ? a.weights[r] : 1.0f;
      if (a.op == 0) { acc += v * w; wsum += w; }
      else if (a.op == 1) acc += v;
      else if (a.op == 2) acc = fmaxf(acc, v);
      else acc = fminf(acc, v);
    }
    a.out[i] = (a.op == 0) ? acc / wsum : acc;
  }
  rank_barrier(a.t, a.rank, a.world, ep + 2);
  if (threadIdx.x == 0) a.epoch[blockIdx.x] = ep + 2;
}

static int fill_table(PeerTable& t, void* const* bufs, void* const* flags, int world) {
  DET_REQUIRE(world >= 1 && world <= COMM_MAX_RANKS, "world size must be in [1, 8]");
  for (int r = 0; r < COMM_MAX_RANKS; ++r) {
    t.buf[r] = r < world ? bufs[r] : nullptr;
    t.flags[r] = r < world ? reinterpret_cast<uint32_t*>(flags[r]) : nullptr;
  }
  return 0;
}

}  // namespace det

// Bytes a flag table needs on each rank (zero-initialised once), and the epoch counter array.
extern "C" long long det_comm_flag_bytes() { return 1LL * det::FLAG_WORDS * 4; }
extern "C" long long det_comm_epoch_bytes() { return 1LL * det::EPOCH_WORDS * 4; }

// peer_bufs / peer_flags: host arrays of `world` device pointers (this rank's view of every rank).
// threads: CTA size. 1024 (default) suits a collective that has the GPU to itself; 128 makes a CTA small enough
// (8 K registers, ~130 B of shared memory) to sit NEXT to a resident GEMM CTA on the same SM, which is how the
// overlapped all-reduce runs under backward without taking SMs away from it.
static int fused_allreduce_launch(void* const* peer_bufs, void* const* peer_flags, void* mc_ptr,
                                  void* epoch_dev, long long n_f32, int rank, int world, float scale,
                                  const float* scale_dev, float* sumsq_dev, int* found_inf_dev,
                                  int max_blocks, int threads, int algo, cudaStream_t stream) {
  using namespace det;
  if (n_f32 <= 0) return 0;
  DET_REQUIRE(n_f32 % 4 == 0, "bucket length must be a multiple of 4 floats");
  AllreduceArgs a;
  int rc = fill_table(a.t, peer_bufs, peer_flags, world);
  if (rc) return rc;
  a.mc_ptr = reinterpret_cast<float*>(mc_ptr);
  a.epoch = reinterpret_cast<uint32_t*>(epoch_dev);
  a.n = n_f32; a.rank = rank; a.world = world; a.scale = scale; a.scale_dev = scale_dev;
  a.sumsq = sumsq_dev; a.found_inf = found_inf_dev;
  const long long per4 = ((n_f32 >> 2) + world - 1) / world;
  if (threads <= 0) threads = AR_THREADS;
  // algo: 0 = automatic, 1 = multimem (NVLS), 2 = LSU peer loads / stores, 3 = TMA bulk copies over peer memory
  static const int algo_env = [] { const char* e = getenv("DET_AR_ALGO"); return e ? atoi(e) : 0; }();
  if (algo == 0) algo = algo_env;
  DET_REQUIRE(threads % 32 == 0 && threads >= 32 && threads <= AR_THREADS, "CTA size must be a multiple of 32 in [32, 1024]");
  // Automatic choice, from the bytes each algorithm puts on a GPU's links: peer copies move 2 (world-1)/world of the
  // bucket per direction, the switch reduction moves 1 + 1/world. Two ranks: 1.0 against 1.5 — bulk copies win (measured
  // 0.70 against 1.15 ms on the 461 MB bucket); from four ranks on the in-switch reduction moves less. CTAs smaller than
  // the default (the co-resident overlapped collective) cannot hold the bulk kernel's stages.
  if (algo == 0) algo = (threads < AR_THREADS || world > 2) ? (mc_ptr != nullptr ? 1 : (threads < AR_THREADS ? 2 : 3)) : 3;
  if (world <= 1) algo = 2;
  if (algo == 1 && mc_ptr == nullptr) algo = 2;
  const long long per_block = static_cast<long long>(threads) * (algo == 1 ? AR_UNROLL_MC : AR_UNROLL_P2P);
  int blocks = static_cast<int>(std::min<long long>((per4 + per_block - 1) / per_block, std::min(sm_count(), COMM_MAX_BLOCKS)));
  // Measured (profiles/allreduce_vs_ctas_8gpu_r1.txt, overlap_sweep_2gpu_r1.txt): with 8 switch reductions in flight
  // per thread the links saturate at 2-8 CTAs and MORE CTAs are slower (8 GPUs: 0.98 ms with 8 CTAs, 1.09 ms with 148;
  // 2 GPUs: 1.15 vs 1.20 ms) — extra CTAs only add flag-barrier traffic. Default to 16.
  if (algo == 3) {
    const int ch_bytes = (BK_STAGE_BYTES / world) & ~127;
    const long long slice_bytes = per4 * 16;
    const long long chunks = (slice_bytes + ch_bytes - 1) / ch_bytes;
    static const int bulk_blocks_env = [] { const char* e = getenv("DET_AR_BULK_BLOCKS"); return e ? atoi(e) : 0; }();
    int bb = max_blocks > 0 ? max_blocks : (bulk_blocks_env > 0 ? bulk_blocks_env : 48);
    bb = static_cast<int>(std::max<long long>(1, std::min<long long>(std::min<long long>(bb, chunks), std::min(sm_count(), COMM_MAX_BLOCKS))));
    static bool attr_set[64] = {false};
    int dev = 0;
    DET_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 64 && !attr_set[dev]) {
      DET_CHECK_CUDA(cudaFuncSetAttribute(fused_allreduce_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BK_SMEM_BYTES));
      attr_set[dev] = true;
    }
    fused_allreduce_bulk_kernel<<<bb, BK_THREADS, BK_SMEM_BYTES, stream>>>(a, ch_bytes);
    DET_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  if (max_blocks <= 0) max_blocks = 16;
  blocks = std::min(blocks, max_blocks);
  blocks = std::max(blocks, 1);
  // Every rank must launch the same grid: blocks depends only on (n, world, SM count, max_blocks).
  if (algo == 1 && world > 1) fused_allreduce_kernel<true><<<blocks, threads, 0, stream>>>(a);
  else fused_allreduce_kernel<false><<<blocks, threads, 0, stream>>>(a);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_fused_allreduce(void* const* peer_bufs, void* const* peer_flags, void* mc_ptr,
                                   void* epoch_dev, long long n_f32, int rank, int world, float scale,
                                   const float* scale_dev, float* sumsq_dev, int* found_inf_dev,
                                   int max_blocks, cudaStream_t stream) {
  return fused_allreduce_launch(peer_bufs, peer_flags, mc_ptr, epoch_dev, n_f32, rank, world, scale, scale_dev, sumsq_dev,
                                found_inf_dev, max_blocks, 0, 0, stream);
}

extern "C" int det_fused_allreduce_cta(void* const* peer_bufs, void* const* peer_flags, void* mc_ptr,
                                       void* epoch_dev, long long n_f32, int rank, int world, float scale,
                                       const float* scale_dev, float* sumsq_dev, int* found_inf_dev,
                                       int max_blocks, int threads, int algo, cudaStream_t stream) {
  return fused_allreduce_launch(peer_bufs, peer_flags, mc_ptr, epoch_dev, n_f32, rank, world, scale, scale_dev, sumsq_dev,
                                found_inf_dev, max_blocks, threads, algo, stream);
}

// Pipelined all-reduce, step 1: reduce-scatter + scale + statistics (sumsq_dev / found_inf_dev describe the WHOLE bucket on
// every rank afterwards). The bucket is `chunks` chunks of chunk = world * sub floats (det_pipe_chunk_floats); rank r's reduced
// share of chunk c, [c * chunk + r * sub, + sub), stays in its own bucket; the rest of the bucket still holds local partial
// sums until det_allgather_send(c) of every rank has landed.
extern "C" long long det_pipe_chunk_floats(long long n_f32, int world, int chunks) {
  return det::pipe_geom(n_f32, world, chunks).chunk4 * 4;
}

extern "C" int det_reduce_scatter_stats(void* const* peer_bufs, void* const* peer_flags, void* mc_ptr, void* epoch_dev,
                                        long long n_f32, int rank, int world, int chunks, float scale, const float* scale_dev,
                                        float* sumsq_dev, int* found_inf_dev, int max_blocks, cudaStream_t stream) {
  using namespace det;
  if (n_f32 <= 0) return 0;
  DET_REQUIRE(n_f32 % 4 == 0, "bucket length must be a multiple of 4 floats");
  DET_REQUIRE(chunks >= 1 && chunks <= PIPE_MAX_CHUNKS, "chunks must be in [1, 8]");
  AllreduceArgs a;
  int rc = fill_table(a.t, peer_bufs, peer_flags, world);
  if (rc) return rc;
  a.mc_ptr = reinterpret_cast<float*>(mc_ptr);
  a.epoch = reinterpret_cast<uint32_t*>(epoch_dev);
  a.n = n_f32; a.rank = rank; a.world = world; a.scale = scale; a.scale_dev = scale_dev;
  a.sumsq = sumsq_dev; a.found_inf = found_inf_dev;
  const bool mc = mc_ptr != nullptr && world > 1;
  const PipeGeom g = pipe_geom(n_f32, world, chunks);
  const long long per_block = static_cast<long long>(AR_THREADS) * (mc ? AR_UNROLL_MC : AR_UNROLL_P2P);
  int blocks = static_cast<int>(std::min<long long>((g.sub4 + per_block - 1) / per_block, std::min(sm_count(), COMM_MAX_BLOCKS)));
  blocks = std::max(1, std::min(blocks, max_blocks > 0 ? max_blocks : 16));
  if (mc) reduce_scatter_stats_kernel<true><<<blocks, AR_THREADS, 0, stream>>>(a, chunks);
  else reduce_scatter_stats_kernel<false><<<blocks, AR_THREADS, 0, stream>>>(a, chunks);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Step 2, every rank, c = 0 .. chunks-1 in order: push this rank's reduced share of chunk c into every rank's bucket, then
// raise its flag for that chunk there.
extern "C" int det_allgather_send(void* const* peer_bufs, void* const* peer_flags, void* mc_ptr, void* epoch_dev,
                                  long long n_f32, int rank, int world, int chunks, int c, int max_blocks, cudaStream_t stream) {
  using namespace det;
  if (n_f32 <= 0) return 0;
  DET_REQUIRE(chunks >= 1 && chunks <= PIPE_MAX_CHUNKS && c >= 0 && c < chunks, "chunk index out of range");
  AllreduceArgs a;
  int rc = fill_table(a.t, peer_bufs, peer_flags, world);
  if (rc) return rc;
  a.mc_ptr = reinterpret_cast<float*>(mc_ptr);
  a.epoch = reinterpret_cast<uint32_t*>(epoch_dev);
  a.n = n_f32; a.rank = rank; a.world = world; a.scale = 1.0f; a.scale_dev = nullptr; a.sumsq = nullptr; a.found_inf = nullptr;
  const bool mc = mc_ptr != nullptr && world > 1;
  const PipeGeom g = pipe_geom(n_f32, world, chunks);
  const long long per_block = static_cast<long long>(AR_THREADS) * 8;
  int blocks = static_cast<int>(std::min<long long>((g.sub4 + per_block - 1) / per_block, std::min(sm_count(), COMM_MAX_BLOCKS)));
  blocks = std::max(1, std::min(blocks, max_blocks > 0 ? max_blocks : 16));
  if (mc) allgather_send_kernel<true><<<blocks, AR_THREADS, 0, stream>>>(a, chunks, c);
  else allgather_send_kernel<false><<<blocks, AR_THREADS, 0, stream>>>(a, chunks, c);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Step 3, every rank: hold `stream` until every OTHER rank's share of chunk c has landed in this rank's bucket (one warp
// spins; put it in front of the kernel that reads the chunk; this rank's own share needs a stream dependency on its send).
extern "C" int det_wait_chunk(void* local_flags, int c, int rank, int world, cudaStream_t stream) {
  using namespace det;
  DET_REQUIRE(c >= 0 && c < PIPE_MAX_CHUNKS, "chunk index out of range");
  wait_chunk_kernel<<<1, 32, 0, stream>>>(reinterpret_cast<const uint32_t*>(local_flags), c, rank, world);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// bf16-wire all-reduce of a fp32 bucket: grad (local fp32, n elements, n % 8 == 0), peer_wire = every rank's bf16 staging
// buffer of n elements in the symmetric arena, mc_wire = its multicast address or null. Three launches on `stream`: pack,
// collective, unpack; the side outputs (sumsq / found_inf) mean the same as in det_fused_allreduce.
extern "C" int det_fused_allreduce_bf16wire(float* grad, void* const* peer_wire, void* const* peer_flags, void* mc_wire,
                                            void* epoch_dev, long long n, int rank, int world, float scale, const float* scale_dev,
                                            float* sumsq_dev, int* found_inf_dev, int max_blocks, cudaStream_t stream) {
  using namespace det;
  if (n <= 0) return 0;
  DET_REQUIRE(n % 8 == 0, "bf16-wire bucket length must be a multiple of 8 elements");
  AllreduceArgs a;
  int rc = fill_table(a.t, peer_wire, peer_flags, world);
  if (rc) return rc;
  a.mc_ptr = reinterpret_cast<float*>(mc_wire);
  a.epoch = reinterpret_cast<uint32_t*>(epoch_dev);
  a.n = n; a.rank = rank; a.world = world; a.scale = scale; a.scale_dev = scale_dev;
  a.sumsq = sumsq_dev; a.found_inf = found_inf_dev;
  const long long n8 = n >> 3;
  const int local_blocks = static_cast<int>(std::max<long long>(1, std::min<long long>((n8 + 255) / 256, 8LL * sm_count())));
  wire_pack_kernel<<<local_blocks, 256, 0, stream>>>(grad, reinterpret_cast<uint4*>(peer_wire[rank]), n8);
  const long long per8 = (n8 + world - 1) / world;
  const bool mc = mc_wire != nullptr && world > 1;
  const long long per_block = static_cast<long long>(AR_THREADS) * (mc ? AR_UNROLL_MC : 1);
  int blocks = static_cast<int>(std::min<long long>((per8 + per_block - 1) / per_block, std::min(sm_count(), COMM_MAX_BLOCKS)));
  blocks = std::max(1, std::min(blocks, max_blocks > 0 ? max_blocks : 16));
  if (mc) fused_allreduce_bf16wire_kernel<true><<<blocks, AR_THREADS, 0, stream>>>(a);
  else fused_allreduce_bf16wire_kernel<false><<<blocks, AR_THREADS, 0, stream>>>(a);
  wire_unpack_kernel<<<local_blocks, 256, 0, stream>>>(reinterpret_cast<const uint4*>(peer_wire[rank]), grad, n8);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_p2p_broadcast(void* const* peer_bufs, void* const* peer_flags, void* epoch_dev,
                                 long long nbytes, int rank, int world, int root, cudaStream_t stream) {
  using namespace det;
  if (nbytes <= 0 || world <= 1) return 0;
  DET_REQUIRE(nbytes % 16 == 0, "broadcast size must be a multiple of 16 bytes");
  BcastArgs a;
  int rc = fill_table(a.t, peer_bufs, peer_flags, world);
  if (rc) return rc;
  a.epoch = reinterpret_cast<uint32_t*>(epoch_dev);
  a.nbytes = nbytes; a.rank = rank; a.world = world; a.root = root;
  const int blocks = static_cast<int>(std::max<long long>(1, std::min<long long>((nbytes / 16 + 511) / 512, std::min(sm_count(), COMM_MAX_BLOCKS))));
  p2p_broadcast_kernel<<<blocks, 512, 0, stream>>>(a);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int det_metric_allreduce(void* const* peer_bufs, void* const* peer_flags, void* epoch_dev,
                                    float* out_dev, const float* weights_dev, int n, int rank, int world,
                                    int op, cudaStream_t stream) {
  using namespace det;
  if (n <= 0) return 0;
  MetricArgs a;
  int rc = fill_table(a.t, peer_bufs, peer_flags, world);
  if (rc) return rc;
  a.epoch = reinterpret_cast<uint32_t*>(epoch_dev);
  a.out = out_dev; a.weights = weights_dev; a.n = n; a.rank = rank; a.world = world; a.op = op;
  const int blocks = std::max(1, std::min((n + 255) / 256, 8));
  metric_allreduce_kernel<<<blocks, 256, 0, stream>>>(a);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---- CUDA IPC plumbing for the symmetric arena (cudaMalloc'd, exported to the other local ranks) ----
extern "C" int det_arena_alloc(long long bytes, void** out) {
  DET_CHECK_CUDA(cudaMalloc(out, static_cast<size_t>(bytes)));
  DET_CHECK_CUDA(cudaMemset(*out, 0, static_cast<size_t>(bytes)));
  return 0;
}
extern "C" int det_arena_free(void* p) {
  DET_CHECK_CUDA(cudaFree(p));
  return 0;
}
extern "C" int det_ipc_export(void* p, unsigned char* handle64) {
  cudaIpcMemHandle_t h;
  DET_CHECK_CUDA(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return 0;
}
extern "C" int det_ipc_import(const unsigned char* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DET_CHECK_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
extern "C" int det_ipc_close(void* p) {
  DET_CHECK_CUDA(cudaIpcCloseMemHandle(p));
  return 0;
}
