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
#include "gemm_common.cuh"

namespace det {

constexpr int GEMM_BM = 128;

template <int BN>
struct GemmCfg {
  static constexpr int STAGES = (BN == 256) ? 4 : ((BN == 128) ? 6 : 8);
  static constexpr uint32_t A_BYTES = GEMM_BM * GEMM_BK * 2;
  static constexpr uint32_t B_BYTES = BN * GEMM_BK * 2;
  static constexpr uint32_t TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));
  // 193.25 KB for BN = 256: together with the 1 KB the system reserves per CTA this still fits the 196 KB shared-memory
  // carve-out and leaves 60 KB of L1 for the epilogue's global traffic. Two more KB tip it into the 228 KB carve-out
  // (28 KB of L1) and cost the BERT step 1.9 % — so only the STATS specialisation pays for its accumulators.
  static constexpr uint32_t SMEM_BYTES = STAGES * (A_BYTES + B_BYTES) + 1024 /*align slack*/ + 256 /*barriers*/;
  static constexpr uint32_t SMEM_BYTES_STATS = SMEM_BYTES + GEMM_STATS_SMEM_BYTES /*column statistics*/;
};

// STATS = true additionally accumulates per-column sum / sum of squares of the output (BatchNorm statistics); it is a
// template parameter, not a run-time branch: even a never-taken branch in the epilogue cost the BERT GEMMs 4 %.
template <int BN, int A_MN, int B_MN, bool STATS = false>
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
  ptx::pdl_launch_dependents();   // the next kernel may start its prologue as soon as SMs free up

  const int total_work = p.tiles_m * p.tiles_n * p.splitk;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      ptx::pdl_wait();            // operands may be the previous kernel's output
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
    constexpr int EPI_THREADS = GEMM_EPI_WARPS * 32;
    constexpr bool stats = STATS;
    float* s_stat = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [2][BN]
    const int tid_e = ew * 32 + lane;
    int stat_n0 = -1;                               // n-tile whose sums currently sit in s_stat
    if (stats) {
      for (int i = tid_e; i < 2 * BN; i += EPI_THREADS) s_stat[i] = 0.0f;
      epilogue_bar_sync(EPI_THREADS);
    }
    ptx::pdl_wait();              // bias / aux_in / C belong to earlier kernels
    uint32_t lt = 0;
    for (int w = blockIdx.x; w < total_work; w += gridDim.x, ++lt) {
      const int tile = w / p.splitk;
      const int m0 = (tile % p.tiles_m) * GEMM_BM;
      const int n0 = (tile / p.tiles_m) * BN;
      if (stats && stat_n0 != n0) {
        if (stat_n0 >= 0) {
          epilogue_bar_sync(EPI_THREADS);
          stats_flush(s_stat, BN, p.colstats, p.N, stat_n0, tid_e);
          epilogue_bar_sync(EPI_THREADS);
        }
        stat_n0 = n0;
      }
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
        gemm_epilogue_chunk<CW>(p, v, row, row_ok, n0 + c);
        if (stats && n0 + c < p.N) stats_accumulate<CW>(s_stat, BN, v, row_ok, c, lane);
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&tempty_bar[buf]));
    }
    if (stats && stat_n0 >= 0) {
      epilogue_bar_sync(EPI_THREADS);
      stats_flush(s_stat, BN, p.colstats, p.N, stat_n0, tid_e);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

template <int BN, int A_MN, int B_MN, bool STATS = false>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid,
                       cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  auto kfn = gemm_bf16_tcgen05_kernel<BN, A_MN, B_MN, STATS>;
  constexpr uint32_t SMEM = STATS ? Cfg::SMEM_BYTES_STATS : Cfg::SMEM_BYTES;
  static bool attr_set[64] = {false};
  int dev = 0;
  DET_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DET_CHECK_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr_set[dev] = true;
  }
  static const bool pdl = [] { const char* e = getenv("DET_PDL"); return e && e[0] == '1'; }();
  if (pdl) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(GEMM_THREADS); cfg.dynamicSmemBytes = SMEM; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    DET_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kfn, ta, tb, p));
  } else {
    kfn<<<grid, GEMM_THREADS, SMEM, stream>>>(ta, tb, p);
  }
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
static int gemm_launch_impl(const void* A, const void* B, void* C, int M, int N, int K, int a_mn,
                            int b_mn, long long lda, long long ldb, long long ldc, int out_mode,
                            int epi, float alpha, const float* bias, const void* aux_in,
                            void* aux_out, long long ld_aux, float dropout_p,
                            unsigned int dropout_seed, const unsigned int* dropout_seed_dev, int splitk,
                            int force_bn, float* colstats,
                            cudaStream_t stream) {
  using namespace det;
  if (colstats != nullptr) {
    DET_REQUIRE(out_mode == OUT_BF16, "column statistics are defined for bf16 outputs");
    if (force_bn <= 0 || force_bn == 512) force_bn = N <= 64 ? 64 : (N <= 128 ? 128 : 256);   // the pair kernel has no statistics path
  }
  DET_REQUIRE(M > 0 && N > 0 && K > 0, "empty GEMM");
  DET_REQUIRE(out_mode >= 0 && out_mode <= 2, "bad out_mode");
  const int sms = sm_count();
  // force_bn == 512 selects the CTA-pair kernel (gemm_pair_sm100.cu: 256 x 256 tile over two SMs, TMA-store epilogue).
  // Automatic choice: every problem large enough to give most pairs a tile; DET_GEMM_2CTA=0 switches it off.
  static const int pair_auto = [] { const char* e = getenv("DET_GEMM_2CTA"); return (e && e[0] == '0') ? 0 : 1; }();
  const bool pair_wanted = force_bn == 512 || (force_bn <= 0 && pair_auto && (M >= 1024 || out_mode == OUT_F32_ATOMIC_ADD) && M >= 256 && N >= 256 &&
                                               static_cast<long long>(M) * N >= (out_mode == OUT_F32_ATOMIC_ADD ? (1LL << 19) : (1LL << 21)));
  GemmParams p;
  p.M = M; p.N = N; p.K = K;
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
  p.colstats = colstats;

  const bool pair = pair_wanted && pair_gemm_supported(p, A, B, lda, ldb);
  DET_REQUIRE(pair || force_bn != 512,
              "the CTA-pair kernel needs 16-byte aligned outputs whose row pitch is a multiple of 16 bytes");
  const int bn = pair ? 256 : (force_bn > 0 ? force_bn : pick_bn(M, N, sms, out_mode == OUT_F32_ATOMIC_ADD));
  DET_REQUIRE(bn == 64 || bn == 128 || bn == 256, "BN must be 64, 128 or 256 (512 = CTA pair)");

  CUtensorMap ta, tb;
  int rc;
  if (a_mn == 0) rc = make_tma_2d_bf16(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, 64, GEMM_BM);
  else           rc = make_tma_2d_bf16(&ta, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, 64);
  if (rc) return rc;
  if (b_mn == 0) rc = make_tma_2d_bf16(&tb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, 64, pair ? 128u : (uint32_t)bn);
  else           rc = make_tma_2d_bf16(&tb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, 64);
  if (rc) return rc;
  if (pair) return launch_pair_gemm(ta, tb, p, a_mn, b_mn, splitk, stream);

  p.tiles_m = (M + GEMM_BM - 1) / GEMM_BM;
  p.tiles_n = (N + bn - 1) / bn;
  p.num_kb = (K + GEMM_BK - 1) / GEMM_BK;
  const int tiles = p.tiles_m * p.tiles_n;
  if (out_mode != OUT_F32_ATOMIC_ADD) {
    splitk = 1;
  } else if (splitk <= 0) {
    splitk = 1;
    // largest split that still fits ONE wave: rounding up instead (54 tiles x 3 = 162 items on 148 SMs) costs a second,
    // nearly empty wave — 35.3 vs 25.6 us on the 2304 x 768 x 6144 weight gradient (profiles/wgrad_sweep_r1.json)
    if (tiles < sms) splitk = max(1, min(p.num_kb / 4, sms / tiles));
  }
  splitk = max(1, min(splitk, p.num_kb));
  p.kb_per_split = (p.num_kb + splitk - 1) / splitk;
  p.splitk = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;

  const int grid = min(tiles * p.splitk, sms);
  if (colstats != nullptr) {
    DET_REQUIRE(a_mn == 0 && b_mn == 0, "column statistics need K-major operands");
    if (bn == 256) return launch_gemm<256, 0, 0, true>(ta, tb, p, grid, stream);
    if (bn == 128) return launch_gemm<128, 0, 0, true>(ta, tb, p, grid, stream);
    return launch_gemm<64, 0, 0, true>(ta, tb, p, grid, stream);
  }
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

extern "C" int det_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int a_mn,
                             int b_mn, long long lda, long long ldb, long long ldc, int out_mode,
                             int epi, float alpha, const float* bias, const void* aux_in,
                             void* aux_out, long long ld_aux, float dropout_p,
                             unsigned int dropout_seed, const unsigned int* dropout_seed_dev, int splitk,
                             int force_bn, cudaStream_t stream) {
  return gemm_launch_impl(A, B, C, M, N, K, a_mn, b_mn, lda, ldb, ldc, out_mode, epi, alpha, bias, aux_in, aux_out, ld_aux,
                          dropout_p, dropout_seed, dropout_seed_dev, splitk, force_bn, nullptr, stream);
}

// Same GEMM, and colstats[0:N] += column sums, colstats[N:2N] += column sums of squares of the bf16 output (the caller
// clears colstats). Serves BatchNorm batch statistics of 1x1 convolutions without a second pass over the output.
extern "C" int det_gemm_bf16_stats(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb,
                                   long long ldc, int epi, const float* bias, float* colstats, cudaStream_t stream) {
  return gemm_launch_impl(A, B, C, M, N, K, 0, 0, lda, ldb, ldc, 0, epi, 1.0f, bias, nullptr, nullptr, 0, 0.0f, 0, nullptr, 0, 0,
                          colstats, stream);
}
