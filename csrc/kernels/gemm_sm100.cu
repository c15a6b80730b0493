// gemm_sm100.cu — persistent, warp-specialised bf16 GEMM for sm_100a.
//
//   C[M,N] = epilogue( A[M,K] * B[N,K]^T )        fp32 accumulation in TMEM
//
// Pipeline: warp 0 = TMA producer (cp.async.bulk.tensor into SWIZZLE_128B smem stages),
//           warp 1 = tcgen05.mma issuer (one elected thread) + TMEM owner,
//           warps 2..9 = epilogue (tcgen05.ld -> registers -> fused epilogue -> global).
// Two TMEM accumulator buffers let the epilogue of tile i overlap the MMAs of tile i+1.
//
// Either operand may be K-major (the contraction dim is contiguous in memory) or MN-major (the
// M/N dim is contiguous) so that forward (X W^T), dgrad (dY W) and wgrad (dY^T X) all run without
// any transposed copies.  Split-K with fp32 `red.global.add` output serves wgrad, where the output
// is small and K (= tokens) is long; the same mode accumulates straight into the gradient bucket,
// which is how aggregation_frequency micro-batches are summed without a separate kernel.
//
// This replaces the cuBLAS calls PyTorch makes for nn.Linear in the reference's BERT example
// (reference: examples/nlp/bert_glue_pytorch/model_def.py:27-174 → HF BertModel → torch.addmm).
#include "host_utils.h"
#include "sm100_ptx.cuh"

namespace det {

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
};

// erf GELU through the Abramowitz–Stegun 7.1.26 rational form (|error| < 1.5e-7, far below bf16 resolution): one
// reciprocal, one exp2 and six FMAs per element instead of libdevice erff's branchy ~40 instructions. The epilogue
// is instruction-issue bound, so this is what keeps it under the MMA time of a tile.
__device__ __forceinline__ void phi_terms(float x, float& cdf, float& e) {
  const float ax = fabsf(x) * 0.70710678118654752f;
  const float t = __frcp_rn(fmaf(0.3275911f, ax, 1.0f));
  e = exp2f(-1.4426950408889634f * ax * ax);  // exp(-x^2/2)
  float poly = fmaf(1.061405429f, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  const float half_erfc = 0.5f * poly * t * e;      // 0.5 * erfc(|x|/sqrt2)
  cdf = x >= 0.0f ? 1.0f - half_erfc : half_erfc;
}
__device__ __forceinline__ float gelu_erf(float x) {
  float cdf, e;
  phi_terms(x, cdf, e);
  return x * cdf;
}
__device__ __forceinline__ float dgelu_erf(float x) {
  float cdf, e;
  phi_terms(x, cdf, e);
  return fmaf(x * 0.3989422804014327f, e, cdf);
}

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_EPI_WARPS = 16;  // 4 per TMEM lane quarter: the epilogue is issue-bound, so give every SMSP 4 warps
constexpr int GEMM_THREADS = 64 + 32 * GEMM_EPI_WARPS;

template <int BN>
struct GemmCfg {
  static constexpr int STAGES = (BN == 256) ? 4 : ((BN == 128) ? 6 : 8);
  static constexpr uint32_t A_BYTES = GEMM_BM * GEMM_BK * 2;
  static constexpr uint32_t B_BYTES = BN * GEMM_BK * 2;
  static constexpr uint32_t TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));
  static constexpr uint32_t SMEM_BYTES = STAGES * (A_BYTES + B_BYTES) + 1024 /*align slack*/ + 256 /*barriers*/;
};

template <int BN, int A_MN, int B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a,
                         const __grid_constant__ CUtensorMap tma_b, const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * Cfg::A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + STAGES * Cfg::B_BYTES);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + STAGES;
  uint64_t* tfull_bar = bars + 2 * STAGES;
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&tma_a);
    ptx::tma_prefetch_desc(&tma_b);
    for (int i = 0; i < STAGES; ++i) {
      ptx::mbar_init(ptx::smem_u32(&full_bar[i]), 1);
      ptx::mbar_init(ptx::smem_u32(&empty_bar[i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(ptx::smem_u32(&tfull_bar[i]), 1);
      ptx::mbar_init(ptx::smem_u32(&tempty_bar[i]), GEMM_EPI_WARPS);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(ptx::smem_u32(tmem_slot), Cfg::TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_work = p.tiles_m * p.tiles_n * p.splitk;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t it = 0;
      for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
        const int ks = w % p.splitk;
        const int tile = w / p.splitk;
        const int m0 = (tile % p.tiles_m) * GEMM_BM;
        const int n0 = (tile / p.tiles_m) * BN;
        const int kb0 = ks * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.num_kb);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const uint32_t s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          ptx::mbar_wait(ptx::smem_u32(&empty_bar[s]), ph ^ 1);
          const uint32_t fb = ptx::smem_u32(&full_bar[s]);
          ptx::mbar_expect_tx(fb, Cfg::A_BYTES + Cfg::B_BYTES);
          const uint32_t a_dst = ptx::smem_u32(sA + s * Cfg::A_BYTES);
          const uint32_t b_dst = ptx::smem_u32(sB + s * Cfg::B_BYTES);
          const int k0 = kb * GEMM_BK;
          if (A_MN == 0) {
            ptx::tma_load_2d(a_dst, &tma_a, fb, k0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < GEMM_BM / 64; ++j)
              ptx::tma_load_2d(a_dst + j * 8192, &tma_a, fb, m0 + 64 * j, k0);
          }
          if (B_MN == 0) {
            ptx::tma_load_2d(b_dst, &tma_b, fb, k0, n0);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              ptx::tma_load_2d(b_dst + j * 8192, &tma_b, fb, n0 + 64 * j, k0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(GEMM_BM, BN, A_MN, B_MN);
      uint32_t it = 0;
      uint32_t lt = 0;
      for (int w = blockIdx.x; w < total_work; w += gridDim.x, ++lt) {
        const int ks = w % p.splitk;
        const int kb0 = ks * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.num_kb);
        const uint32_t buf = lt & 1;
        const uint32_t aph = (lt >> 1) & 1;
        ptx::mbar_wait(ptx::smem_u32(&tempty_bar[buf]), aph ^ 1);
        ptx::tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * BN;
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const uint32_t s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          ptx::mbar_wait(ptx::smem_u32(&full_bar[s]), ph);
          ptx::tc_fence_after();
          const uint32_t a_base = ptx::smem_u32(sA + s * Cfg::A_BYTES);
          const uint32_t b_base = ptx::smem_u32(sB + s * Cfg::B_BYTES);
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k) {
            const uint64_t adesc = A_MN ? ptx::make_smem_desc_sw128(a_base + k * 2048, 8192, 1024)
                                        : ptx::make_smem_desc_sw128(a_base + k * 32, 16, 1024);
            const uint64_t bdesc = B_MN ? ptx::make_smem_desc_sw128(b_base + k * 2048, 8192, 1024)
                                        : ptx::make_smem_desc_sw128(b_base + k * 32, 16, 1024);
            ptx::umma_bf16_ss(tmem_d, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          ptx::umma_commit(ptx::smem_u32(&empty_bar[s]));  // frees the smem stage when MMAs finish
        }
        ptx::umma_commit(ptx::smem_u32(&tfull_bar[buf]));  // accumulator ready for the epilogue
      }
    }
  } else {
    // ===================== Epilogue warps =====================
    const int ew = warp - 2;
    const int q = warp & 3;            // TMEM lane quarter this warp may access
    const int half = ew >> 2;          // which quarter of the BN columns (0..3)
    constexpr int COLS_PER_WARP = BN / 4;
    constexpr int CW = COLS_PER_WARP < 32 ? COLS_PER_WARP : 32;  // columns per tcgen05.ld chunk (16 when BN == 64)
    uint32_t lt = 0;
    for (int w = blockIdx.x; w < total_work; w += gridDim.x, ++lt) {
      const int tile = w / p.splitk;
      const int m0 = (tile % p.tiles_m) * GEMM_BM;
      const int n0 = (tile / p.tiles_m) * BN;
      const uint32_t buf = lt & 1;
      const uint32_t aph = (lt >> 1) & 1;
      ptx::mbar_wait(ptx::smem_u32(&tfull_bar[buf]), aph);
      ptx::tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < p.M;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN;
#pragma unroll 1
      for (int c = half * COLS_PER_WARP; c < (half + 1) * COLS_PER_WARP; c += CW) {
        float v[CW];
        if constexpr (CW == 32) ptx::tmem_ld_32x32b_x32(taddr + c, v);
        else ptx::tmem_ld_32x32b_x16(taddr + c, v);
        ptx::tmem_ld_wait();
        const int col0 = n0 + c;
        if (col0 >= p.N) continue;  // warp-uniform
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
              const uint32_t dseed = p.drop_seed + (p.drop_seed_dev ? __ldg(p.drop_seed_dev) * 0x9E3779B9u : 0u);
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
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&tempty_bar[buf]));
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

template <int BN, int A_MN, int B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                       cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  auto kfn = gemm_bf16_tcgen05_kernel<BN, A_MN, B_MN>;
  static bool attr_set[64] = {false};
  int dev = 0;
  DET_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DET_CHECK_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set[dev] = true;
  }
  kfn<<<grid, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(ta, tb, p);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static int pick_bn(int M, int N, int sms, bool splitk_ok) {
  // with split-K available the grid is filled along K, so take the tile with the best arithmetic intensity
  if (splitk_ok && N >= 192) return 256;
  // Choose the N-tile that wastes the fewest SM-rounds (persistent grid of `sms` CTAs).
  const int tiles_m = (M + GEMM_BM - 1) / GEMM_BM;
  int best = 128;
  double best_cost = 1e30;
  const int cands[3] = {256, 128, 64};
  for (int bn : cands) {
    if (bn > 64 && N < bn / 2) continue;
    const int tiles = tiles_m * ((N + bn - 1) / bn);
    const int rounds = (tiles + sms - 1) / sms;
    // cost ~ rounds * per-tile time; smaller tiles have lower MMA efficiency (smem-bound B reuse)
    const double eff = (bn == 256) ? 1.0 : (bn == 128 ? 0.8 : 0.55);
    const double cost = rounds * (bn / eff);
    if (cost < best_cost) { best_cost = cost; best = bn; }
  }
  return best;
}

}  // namespace det

// C ABI ------------------------------------------------------------------------------------------
// A: a_mn == 0 -> row-major [M, K] (ld = lda); a_mn == 1 -> stored [K, M] with M contiguous.
// B: b_mn == 0 -> row-major [N, K] (ld = ldb); b_mn == 1 -> stored [K, N] with N contiguous.
// splitk <= 0 picks automatically (only > 1 when out_mode == OUT_F32_ATOMIC_ADD).
extern "C" int det_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int a_mn,
                             int b_mn, long long lda, long long ldb, long long ldc, int out_mode,
                             int epi, float alpha, const float* bias, const void* aux_in,
                             void* aux_out, long long ld_aux, float dropout_p,
                             unsigned int dropout_seed, const unsigned int* dropout_seed_dev, int splitk,
                             int force_bn,
                             cudaStream_t stream) {
  using namespace det;
  DET_REQUIRE(M > 0 && N > 0 && K > 0, "empty GEMM");
  DET_REQUIRE(out_mode >= 0 && out_mode <= 2, "bad out_mode");
  const int sms = sm_count();
  int bn = force_bn > 0 ? force_bn : pick_bn(M, N, sms, out_mode == OUT_F32_ATOMIC_ADD);
  DET_REQUIRE(bn == 64 || bn == 128 || bn == 256, "BN must be 64, 128 or 256");
  GemmParams p;
  p.M = M; p.N = N; p.K = K;
  p.tiles_m = (M + GEMM_BM - 1) / GEMM_BM;
  p.tiles_n = (N + bn - 1) / bn;
  p.num_kb = (K + GEMM_BK - 1) / GEMM_BK;
  const int tiles = p.tiles_m * p.tiles_n;
  if (out_mode != OUT_F32_ATOMIC_ADD) {
    splitk = 1;
  } else if (splitk <= 0) {
    splitk = 1;
    if (tiles < sms) splitk = max(1, min(p.num_kb / 4, (sms + tiles / 2) / tiles));
  }
  splitk = max(1, min(splitk, p.num_kb));
  p.kb_per_split = (p.num_kb + splitk - 1) / splitk;
  p.splitk = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
  p.C = C; p.ldc = ldc; p.out_mode = out_mode; p.epi = epi; p.alpha = alpha;
  p.bias = bias;
  p.aux_in = reinterpret_cast<const __nv_bfloat16*>(aux_in);
  p.aux_out = reinterpret_cast<__nv_bfloat16*>(aux_out);
  p.ld_aux = ld_aux;
  if (epi == EPI_DGELU || epi == EPI_BIAS_RESID) DET_REQUIRE(aux_in != nullptr, "epilogue needs aux_in");
  if (dropout_p > 0.0f) {
    DET_REQUIRE(dropout_p < 1.0f, "dropout p must be < 1");
    p.drop_threshold = static_cast<uint32_t>(static_cast<double>(dropout_p) * 4294967296.0);
    p.drop_scale = 1.0f / (1.0f - dropout_p);
  } else {
    p.drop_threshold = 0; p.drop_scale = 1.0f;
  }
  p.drop_seed = dropout_seed;
  p.drop_seed_dev = dropout_seed_dev;

  CUtensorMap ta, tb;
  int rc;
  if (a_mn == 0) rc = make_tma_2d_bf16(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, 64, GEMM_BM);
  else           rc = make_tma_2d_bf16(&ta, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, 64);
  if (rc) return rc;
  if (b_mn == 0) rc = make_tma_2d_bf16(&tb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, 64, (uint32_t)bn);
  else           rc = make_tma_2d_bf16(&tb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, 64);
  if (rc) return rc;

  const int grid = min(tiles * p.splitk, sms);
#define DET_GEMM_DISPATCH(BN_)                                                     \
  if (a_mn == 0 && b_mn == 0) return launch_gemm<BN_, 0, 0>(ta, tb, p, grid, stream); \
  if (a_mn == 0 && b_mn == 1) return launch_gemm<BN_, 0, 1>(ta, tb, p, grid, stream); \
  if (a_mn == 1 && b_mn == 0) return launch_gemm<BN_, 1, 0>(ta, tb, p, grid, stream); \
  return launch_gemm<BN_, 1, 1>(ta, tb, p, grid, stream);
  if (bn == 256) { DET_GEMM_DISPATCH(256) }
  if (bn == 128) { DET_GEMM_DISPATCH(128) }
  DET_GEMM_DISPATCH(64)
#undef DET_GEMM_DISPATCH
}
