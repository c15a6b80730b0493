// gemm_common.cuh — parameter block, fused-epilogue math and the per-chunk epilogue shared by the 1-CTA and the
// CTA-pair (cta_group::2) GEMM kernels.
#pragma once
#include "host_utils.h"
#include "sm100_ptx.cuh"

namespace det {

constexpr int GEMM_BK = 64;
constexpr int GEMM_EPI_WARPS = 16;  // 4 per TMEM lane quarter: the epilogue is issue-bound, so give every SMSP 4 warps
constexpr int GEMM_THREADS = 64 + 32 * GEMM_EPI_WARPS;

enum GemmEpilogue : int {
  EPI_NONE = 0,
  EPI_BIAS = 1,        // C = acc + bias
  EPI_BIAS_GELU = 2,   // Z = acc + bias -> aux_out (optional); C = gelu(Z)
  EPI_DGELU = 3,       // C = acc * gelu'(aux_in)
  EPI_BIAS_RESID = 4,  // C = dropout(acc + bias) + aux_in
};
enum GemmOut : int { OUT_BF16 = 0, OUT_F32 = 1, OUT_F32_ATOMIC_ADD = 2 };

struct GemmParams {
  int M, N, K;
  int tiles_m, tiles_n, splitk, kb_per_split, num_kb;
  void* C;
  long long ldc;
  int out_mode;
  int epi;
  float alpha;
  const float* bias;             // [N] fp32 or null
  const __nv_bfloat16* aux_in;   // [M, ld_aux] bf16 or null
  __nv_bfloat16* aux_out;        // [M, ld_aux] bf16 or null
  long long ld_aux;
  uint32_t drop_threshold;  // 0 = no dropout
  float drop_scale;
  uint32_t drop_seed;
  const uint32_t* drop_seed_dev;  // optional per-step seed in device memory (CUDA-graph friendly)
  float* colstats;                // optional [2 * N] fp32: += column sums and sums of squares of the (bf16-rounded) output
};

// CTA-pair kernel (gemm_pair_sm100.cu). `p` carries the problem, output and epilogue fields; tiling fields are set inside.
bool pair_gemm_supported(const GemmParams& p, const void* A, const void* B, long long lda, long long ldb);
int launch_pair_gemm(const CUtensorMap& ta, const CUtensorMap& tb, GemmParams p, int a_mn, int b_mn, int splitk,
                     cudaStream_t stream);

// GELU in the tanh form, x * 0.5 * (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3))) — the form of the original BERT code — on the
// hardware tanh (one MUFU op, relative error 2^-11): six FP32 instructions and one MUFU per element. The epilogue is
// instruction-issue bound (16 warps share four schedulers with 32 K elements per tile), and the erf form used before —
// a rational erfc with a reciprocal, an exp2 and five FMAs — cost about three times as much: 38.6 against 27.9 us for
// the FFN GEMM with and without it. The two forms differ by at most 5e-4 in absolute value, below the bf16 resolution of
// the outputs for all but tiny activations; forward and backward use the same function, so gradients are consistent.
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float gelu_erf(float x) {   // name kept: the activation every `epi = BIAS_GELU` caller gets
  const float x2 = x * x;
  const float t = tanh_approx(x * fmaf(0.035677408136300125f, x2, 0.7978845608028654f));
  const float hx = 0.5f * x;
  return fmaf(hx, t, hx);
}
__device__ __forceinline__ float dgelu_erf(float x) {
  const float x2 = x * x;
  const float t = tanh_approx(x * fmaf(0.035677408136300125f, x2, 0.7978845608028654f));
  const float q = fmaf(0.10703222440890037f, x2, 0.7978845608028654f);     // d/dx of the tanh argument
  const float s = fmaf(-t, t, 1.0f);                                       // sech^2
  return fmaf(0.5f * x * s, q, fmaf(0.5f, t, 0.5f));
}

// ------------------------------------------------------------------------------------------------
// Column statistics of the output tile, fused into the epilogue (BatchNorm batch statistics of a convolution output
// without re-reading it). tcgen05.ld leaves one ROW per lane, so a column sum is a reduction ACROSS lanes: a butterfly
// that halves the number of live columns per step leaves lane j holding the total of column j after 31 shuffles.
// Totals go to a per-CTA shared-memory accumulator (all tiles of one n-tile share it) and reach global memory with one
// atomic per column per CTA when the n-tile changes or the kernel ends.
// ------------------------------------------------------------------------------------------------
constexpr int GEMM_STATS_SMEM_BYTES = 2 * 256 * 4;

template <int N_>
__device__ __forceinline__ void col_reduce_step(float* a, int lane, int off) {
  const bool up = (lane & off) != 0;
#pragma unroll
  for (int j = 0; j < N_; ++j) {
    const float send = up ? a[j] : a[j + N_];
    const float keep = up ? a[j + N_] : a[j];
    a[j] = keep + __shfl_xor_sync(0xffffffffu, send, off);
  }
}
// in: a[32] = this lane's row; out: a[0] = total of column `lane` over the warp's 32 rows
__device__ __forceinline__ void col_reduce32(float* a, int lane) {
  col_reduce_step<16>(a, lane, 16);
  col_reduce_step<8>(a, lane, 8);
  col_reduce_step<4>(a, lane, 4);
  col_reduce_step<2>(a, lane, 2);
  col_reduce_step<1>(a, lane, 1);
}
__device__ __forceinline__ void epilogue_bar_sync(int threads) { asm volatile("bar.sync 1, %0;" ::"r"(threads) : "memory"); }

// v[CW] final values of this lane's row (valid = row in range); c = first column of the chunk inside the tile
template <int CW>
__device__ __forceinline__ void stats_accumulate(float* s_stat, int bn, const float* v, bool valid, int c, int lane) {
  float a[32], b[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const float x = (valid && j < CW) ? __bfloat162float(__float2bfloat16_rn(v[j < CW ? j : 0])) : 0.0f;
    a[j] = x;
    b[j] = x * x;
  }
  col_reduce32(a, lane);
  col_reduce32(b, lane);
  if (lane < CW) {
    atomicAdd(s_stat + c + lane, a[0]);
    atomicAdd(s_stat + bn + c + lane, b[0]);
  }
}
// all epilogue threads call this between two epilogue_bar_sync(); tid_e = index among the epilogue threads
__device__ __forceinline__ void stats_flush(float* s_stat, int bn, float* colstats, int N, int n0, int tid_e) {
  if (tid_e < bn) {
    const int col = n0 + tid_e;
    if (col < N) {
      atomicAdd(colstats + col, s_stat[tid_e]);
      atomicAdd(colstats + N + col, s_stat[bn + tid_e]);
    }
    s_stat[tid_e] = 0.0f;
    s_stat[bn + tid_e] = 0.0f;
  }
}

// Epilogue of one tcgen05.ld chunk: v[CW] = accumulator values of `row`, columns col0 .. col0+CW-1.
template <int CW>
__device__ __forceinline__ void gemm_epilogue_chunk(const GemmParams& p, float (&v)[CW], const int row, const bool row_ok, const int col0) {
  if (col0 >= p.N) return;  // warp-uniform
  const bool full_cols = (col0 + CW <= p.N);
  if (p.alpha != 1.0f) {
#pragma unroll
    for (int j = 0; j < CW; ++j) v[j] *= p.alpha;
  }
  if (p.bias != nullptr) {
    if (full_cols) {
      const float4* b4 = reinterpret_cast<const float4*>(p.bias + col0);
#pragma unroll
      for (int j = 0; j < CW / 4; ++j) {
        const float4 b = __ldg(b4 + j);
        v[4 * j + 0] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < CW; ++j)
        if (col0 + j < p.N) v[j] += __ldg(p.bias + col0 + j);
    }
  }
  if (row_ok) {
    const long long aux_off = static_cast<long long>(row) * p.ld_aux + col0;
    if (p.epi == EPI_BIAS_GELU) {
      if (p.aux_out != nullptr) {
        if (full_cols) {
          uint4* zo = reinterpret_cast<uint4*>(p.aux_out + aux_off);
#pragma unroll
          for (int j = 0; j < CW / 8; ++j) {
            uint4 u;
            u.x = ptx::pack_bf16x2(v[8 * j + 0], v[8 * j + 1]);
            u.y = ptx::pack_bf16x2(v[8 * j + 2], v[8 * j + 3]);
            u.z = ptx::pack_bf16x2(v[8 * j + 4], v[8 * j + 5]);
            u.w = ptx::pack_bf16x2(v[8 * j + 6], v[8 * j + 7]);
            zo[j] = u;
          }
        } else {
          #pragma unroll
          for (int j = 0; j < CW; ++j)
            if (col0 + j < p.N) p.aux_out[aux_off + j] = __float2bfloat16_rn(v[j]);
        }
      }
#pragma unroll
      for (int j = 0; j < CW; ++j) v[j] = gelu_erf(v[j]);
    } else if (p.epi == EPI_DGELU) {
      if (full_cols) {
        const uint4* zi = reinterpret_cast<const uint4*>(p.aux_in + aux_off);
#pragma unroll
        for (int j = 0; j < CW / 8; ++j) {
          const uint4 u = __ldg(zi + j);
          v[8 * j + 0] *= dgelu_erf(ptx::bf16_lo(u.x)); v[8 * j + 1] *= dgelu_erf(ptx::bf16_hi(u.x));
          v[8 * j + 2] *= dgelu_erf(ptx::bf16_lo(u.y)); v[8 * j + 3] *= dgelu_erf(ptx::bf16_hi(u.y));
          v[8 * j + 4] *= dgelu_erf(ptx::bf16_lo(u.z)); v[8 * j + 5] *= dgelu_erf(ptx::bf16_hi(u.z));
          v[8 * j + 6] *= dgelu_erf(ptx::bf16_lo(u.w)); v[8 * j + 7] *= dgelu_erf(ptx::bf16_hi(u.w));
        }
      } else {
        #pragma unroll
        for (int j = 0; j < CW; ++j)
          if (col0 + j < p.N) v[j] *= dgelu_erf(__bfloat162float(p.aux_in[aux_off + j]));
      }
    } else if (p.epi == EPI_BIAS_RESID) {
      if (p.drop_threshold != 0u) {
        const uint32_t base = static_cast<uint32_t>(row) * static_cast<uint32_t>(p.N) + static_cast<uint32_t>(col0);
        const uint32_t dseed = dropout_effective_seed(p.drop_seed, p.drop_seed_dev);
#pragma unroll
        for (int j = 0; j < CW; ++j)
          v[j] = dropout_keep(dseed, base + j, p.drop_threshold) ? v[j] * p.drop_scale : 0.0f;
      }
      if (full_cols) {
        const uint4* ri = reinterpret_cast<const uint4*>(p.aux_in + aux_off);
#pragma unroll
        for (int j = 0; j < CW / 8; ++j) {
          const uint4 u = __ldg(ri + j);
          v[8 * j + 0] += ptx::bf16_lo(u.x); v[8 * j + 1] += ptx::bf16_hi(u.x);
          v[8 * j + 2] += ptx::bf16_lo(u.y); v[8 * j + 3] += ptx::bf16_hi(u.y);
          v[8 * j + 4] += ptx::bf16_lo(u.z); v[8 * j + 5] += ptx::bf16_hi(u.z);
          v[8 * j + 6] += ptx::bf16_lo(u.w); v[8 * j + 7] += ptx::bf16_hi(u.w);
        }
      } else {
        #pragma unroll
        for (int j = 0; j < CW; ++j)
          if (col0 + j < p.N) v[j] += __bfloat162float(p.aux_in[aux_off + j]);
      }
    }
    // ---- store ----
    const long long c_off = static_cast<long long>(row) * p.ldc + col0;
    if (p.out_mode == OUT_BF16) {
      __nv_bfloat16* cp = reinterpret_cast<__nv_bfloat16*>(p.C) + c_off;
      if (full_cols && ((reinterpret_cast<uintptr_t>(cp) & 15) == 0)) {
        uint4* co = reinterpret_cast<uint4*>(cp);
#pragma unroll
        for (int j = 0; j < CW / 8; ++j) {
          uint4 u;
          u.x = ptx::pack_bf16x2(v[8 * j + 0], v[8 * j + 1]);
          u.y = ptx::pack_bf16x2(v[8 * j + 2], v[8 * j + 3]);
          u.z = ptx::pack_bf16x2(v[8 * j + 4], v[8 * j + 5]);
          u.w = ptx::pack_bf16x2(v[8 * j + 6], v[8 * j + 7]);
          co[j] = u;
        }
      } else {
        #pragma unroll
        for (int j = 0; j < CW; ++j)
          if (col0 + j < p.N) cp[j] = __float2bfloat16_rn(v[j]);
      }
    } else if (p.out_mode == OUT_F32) {
      float* cp = reinterpret_cast<float*>(p.C) + c_off;
      if (full_cols && ((reinterpret_cast<uintptr_t>(cp) & 15) == 0)) {
        float4* co = reinterpret_cast<float4*>(cp);
#pragma unroll
        for (int j = 0; j < CW / 4; ++j) co[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      } else {
        #pragma unroll
        for (int j = 0; j < CW; ++j)
          if (col0 + j < p.N) cp[j] = v[j];
      }
    } else {
      float* cp = reinterpret_cast<float*>(p.C) + c_off;
      if (full_cols && ((reinterpret_cast<uintptr_t>(cp) & 15) == 0)) {
#pragma unroll
        for (int j = 0; j < CW / 4; ++j)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(cp + 4 * j),
                       "f"(v[4 * j]), "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3])
                       : "memory");
      } else {
        #pragma unroll
        for (int j = 0; j < CW; ++j)
          if (col0 + j < p.N) atomicAdd(cp + j, v[j]);
      }
    }
  }
}

}  // namespace det
