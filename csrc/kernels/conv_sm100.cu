// conv_sm100.cu — implicit-GEMM 2-D convolution (NHWC bf16, unit stride) on tcgen05 for sm_100a.
//
// No im2col buffer exists anywhere: the TMA unit fetches, for every filter tap (r, s) and every 64-channel
// slice, a 4-D box [channels=64, BW, BH, BN] of the activation tensor shifted by the tap offset; coordinates
// that fall outside the image arrive as zeros, which is the padding halo. The box rows land in shared memory
// exactly like the rows of a K-major GEMM operand (128 B per pixel, SWIZZLE_128B), so the MMA side is the
// same single-thread tcgen05.mma pipeline as the plain GEMM (gemm_sm100.cu).
//
//   FPROP  y[n,h,w,co]  = sum_{r,s,ci} x[n,h+r-p,w+s-p,ci] * W[co,r,s,ci]      M = output pixels, K = (r,s,ci)
//   DGRAD  dx[n,h,w,ci] = sum_{r,s,co} dy[n,h-r+p,w-s+p,co] * W[co,r,s,ci]     M = input pixels,  K = (r,s,co)
//   WGRAD  dW[co,r,s,ci] += sum_{n,h,w} dy[n,h,w,co] * x[n,h+r-p,w+s-p,ci]     M = co, N = (r,s,ci), K = pixels
//
// WGRAD contracts over pixels, so both operands are MN-major boxes of <= 64 pixels and the result is added
// into the fp32 gradient buffer with split-K `red.global.add` (aggregation_frequency micro-batches accumulate
// for free, as in the Linear layers).
//
// The reference has no convolution code of its own: its vision examples call torch.nn.Conv2d → cuDNN
// (reference: examples/computer_vision/cifar10_pytorch/model_def.py:49-74,
// examples/tutorials/mnist_pytorch/model_def.py:34-53). Strided convolutions and the few-channel first layers
// take the explicit gather path in conv_ops.cu + the plain GEMM.
#include "gemm_common.cuh"

namespace det {

enum ConvMode : int { CONV_FPROP = 0, CONV_DGRAD = 1, CONV_WGRAD = 2 };

struct ConvParams {
  int PN, PH, PW;             // pixel grid cut into boxes (fprop: output, dgrad: input, wgrad: dy pixels)
  int BW, BH, BNb, rows_valid;
  int tw, th, tn;             // boxes along w, h, n
  int S, taps;
  int sgn, off_h, off_w;      // source pixel = box pixel + sgn * (r, s) + off
  int cchunks;                // fprop/dgrad: 64-wide slices of the contraction channels
  int Cn;                     // channels along GEMM-N (fprop: Cout, dgrad/wgrad: Cin)
  int Cm;                     // wgrad: Cout (GEMM-M)
  int tiles_m, tiles_n, n_per_tap;
  int num_kb, splitk, kb_per_split;
  void* out;
  long long ldc;
  const float* bias;          // fprop only, fp32 [Cout] or null
  int relu;                   // fprop only
  float* colstats;            // fprop only, optional [2 * Cout]: += per-channel sum and sum of squares of the output
};

constexpr int CONV_BM = 128;
constexpr int CONV_BK = 64;
constexpr int CONV_EPI_WARPS = 16;
constexpr int CONV_THREADS = 64 + 32 * CONV_EPI_WARPS;

template <int BN>
struct ConvCfg {
  static constexpr int STAGES = (BN == 256) ? 4 : ((BN == 128) ? 6 : 8);
  static constexpr uint32_t A_BYTES = CONV_BM * CONV_BK * 2;
  static constexpr uint32_t B_BYTES = BN * CONV_BK * 2;
  static constexpr uint32_t TMEM_COLS = (2 * BN <= 128) ? 128 : (2 * BN <= 256 ? 256 : 512);
  static constexpr uint32_t SMEM_BYTES = STAGES * (A_BYTES + B_BYTES) + 1024 + 256;   // see GemmCfg on the carve-out
  static constexpr uint32_t SMEM_BYTES_STATS = SMEM_BYTES + GEMM_STATS_SMEM_BYTES;
};

template <int BN, int MODE, bool STATS = false>
__global__ void __launch_bounds__(CONV_THREADS, 1)
conv_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                    const ConvParams p) {
  using Cfg = ConvCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int A_MN = (MODE == CONV_WGRAD) ? 1 : 0;
  constexpr int B_MN = (MODE == CONV_FPROP) ? 0 : 1;
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

  if (MODE == CONV_WGRAD) {
    // A pixel box may hold fewer than 64 pixels; the K rows it never writes must read as zeros.
    uint4* z = reinterpret_cast<uint4*>(smem);
    const int n16 = STAGES * (Cfg::A_BYTES + Cfg::B_BYTES) / 16;
    for (int i = threadIdx.x; i < n16; i += CONV_THREADS) z[i] = make_uint4(0, 0, 0, 0);
    ptx::fence_proxy_async_smem();
  }
  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&tma_a);
    ptx::tma_prefetch_desc(&tma_b);
    for (int i = 0; i < STAGES; ++i) {
      ptx::mbar_init(ptx::smem_u32(&full_bar[i]), 1);
      ptx::mbar_init(ptx::smem_u32(&empty_bar[i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(ptx::smem_u32(&tfull_bar[i]), 1);
      ptx::mbar_init(ptx::smem_u32(&tempty_bar[i]), CONV_EPI_WARPS);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(ptx::smem_u32(tmem_slot), Cfg::TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_work = p.tiles_m * p.tiles_n * p.splitk;
  const uint32_t box_bytes = static_cast<uint32_t>(p.rows_valid) * 128u;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t it = 0;
      for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
        const int ks = w % p.splitk;
        const int tile = w / p.splitk;
        const int mt = tile % p.tiles_m;
        const int nt = tile / p.tiles_m;
        const int kb0 = ks * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.num_kb);
        if (MODE != CONV_WGRAD) {
          const int w0 = (mt % p.tw) * p.BW;
          const int h0 = ((mt / p.tw) % p.th) * p.BH;
          const int n0 = (mt / (p.tw * p.th)) * p.BNb;
          const int cn0 = nt * BN;
          for (int kb = kb0; kb < kb1; ++kb, ++it) {
            const uint32_t s = it % STAGES;
            const uint32_t ph = (it / STAGES) & 1;
            ptx::mbar_wait(ptx::smem_u32(&empty_bar[s]), ph ^ 1);
            const uint32_t fb = ptx::smem_u32(&full_bar[s]);
            ptx::mbar_expect_tx(fb, box_bytes + Cfg::B_BYTES);
            const uint32_t a_dst = ptx::smem_u32(sA + s * Cfg::A_BYTES);
            const uint32_t b_dst = ptx::smem_u32(sB + s * Cfg::B_BYTES);
            const int tap = kb / p.cchunks;
            const int cc = kb - tap * p.cchunks;
            const int fr = tap / p.S, fs = tap - fr * p.S;
            ptx::tma_load_4d(a_dst, &tma_a, fb, cc * 64, w0 + p.sgn * fs + p.off_w, h0 + p.sgn * fr + p.off_h, n0);
            if (MODE == CONV_FPROP) {
              ptx::tma_load_3d(b_dst, &tma_b, fb, cc * 64, tap, cn0);
            } else {
#pragma unroll
              for (int j = 0; j < BN / 64; ++j) ptx::tma_load_3d(b_dst + j * 8192, &tma_b, fb, cn0 + 64 * j, tap, cc * 64);
            }
          }
        } else {
          const int co0 = mt * CONV_BM;
          const int tap = nt / p.n_per_tap;
          const int ci0 = (nt - tap * p.n_per_tap) * BN;
          const int fr = tap / p.S, fs = tap - fr * p.S;
          for (int kb = kb0; kb < kb1; ++kb, ++it) {
            const uint32_t s = it % STAGES;
            const uint32_t ph = (it / STAGES) & 1;
            ptx::mbar_wait(ptx::smem_u32(&empty_bar[s]), ph ^ 1);
            const uint32_t fb = ptx::smem_u32(&full_bar[s]);
            ptx::mbar_expect_tx(fb, box_bytes * (2 + BN / 64));
            const uint32_t a_dst = ptx::smem_u32(sA + s * Cfg::A_BYTES);
            const uint32_t b_dst = ptx::smem_u32(sB + s * Cfg::B_BYTES);
            const int w0 = (kb % p.tw) * p.BW;
            const int h0 = ((kb / p.tw) % p.th) * p.BH;
            const int n0 = (kb / (p.tw * p.th)) * p.BNb;
#pragma unroll
            for (int j = 0; j < 2; ++j) ptx::tma_load_4d(a_dst + j * 8192, &tma_a, fb, co0 + 64 * j, w0, h0, n0);
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              ptx::tma_load_4d(b_dst + j * 8192, &tma_b, fb, ci0 + 64 * j, w0 + fs + p.off_w, h0 + fr + p.off_h, n0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(CONV_BM, BN, A_MN, B_MN);
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
          for (int k = 0; k < CONV_BK / 16; ++k) {
            const uint64_t adesc = A_MN ? ptx::make_smem_desc_sw128(a_base + k * 2048, 8192, 1024)
                                        : ptx::make_smem_desc_sw128(a_base + k * 32, 16, 1024);
            const uint64_t bdesc = B_MN ? ptx::make_smem_desc_sw128(b_base + k * 2048, 8192, 1024)
                                        : ptx::make_smem_desc_sw128(b_base + k * 32, 16, 1024);
            ptx::umma_bf16_ss(tmem_d, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          ptx::umma_commit(ptx::smem_u32(&empty_bar[s]));
        }
        ptx::umma_commit(ptx::smem_u32(&tfull_bar[buf]));
      }
    }
  } else {
    // ===================== Epilogue warps =====================
    const int ew = warp - 2;
    const int q = warp & 3;
    const int part = ew >> 2;
    constexpr int COLS_PER_WARP = BN / 4;
    constexpr int CW = COLS_PER_WARP < 32 ? COLS_PER_WARP : 32;
    constexpr int EPI_THREADS = CONV_EPI_WARPS * 32;
    constexpr bool stats = STATS && MODE == CONV_FPROP;
    float* s_stat = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [2][BN]
    const int tid_e = ew * 32 + lane;
    int stat_nt = -1;
    if (stats) {
      for (int i = tid_e; i < 2 * BN; i += EPI_THREADS) s_stat[i] = 0.0f;
      epilogue_bar_sync(EPI_THREADS);
    }
    uint32_t lt = 0;
    for (int w = blockIdx.x; w < total_work; w += gridDim.x, ++lt) {
      const int tile = w / p.splitk;
      const int mt = tile % p.tiles_m;
      const int nt = tile / p.tiles_m;
      if (stats && stat_nt != nt) {
        if (stat_nt >= 0) {
          epilogue_bar_sync(EPI_THREADS);
          stats_flush(s_stat, BN, p.colstats, p.Cn, stat_nt * BN, tid_e);
          epilogue_bar_sync(EPI_THREADS);
        }
        stat_nt = nt;
      }
      const uint32_t buf = lt & 1;
      const uint32_t aph = (lt >> 1) & 1;
      const int r = q * 32 + lane;
      bool row_ok;
      long long out_row;
      int cn0, col_lim;        // first GEMM-N channel of the tile, channel bound
      long long col_base;      // offset of channel cn0 inside an output row
      if (MODE != CONV_WGRAD) {
        const int iw = r % p.BW;
        const int t = r / p.BW;
        const int ih = t % p.BH;
        const int in = t / p.BH;
        const int pw = (mt % p.tw) * p.BW + iw;
        const int ph = ((mt / p.tw) % p.th) * p.BH + ih;
        const int pn = (mt / (p.tw * p.th)) * p.BNb + in;
        row_ok = (r < p.rows_valid) && pw < p.PW && ph < p.PH && pn < p.PN;
        out_row = (static_cast<long long>(pn) * p.PH + ph) * p.PW + pw;
        cn0 = nt * BN;
        col_base = cn0;
      } else {
        const int tap = nt / p.n_per_tap;
        cn0 = (nt - tap * p.n_per_tap) * BN;
        const int co = mt * CONV_BM + r;
        row_ok = co < p.Cm;
        out_row = co;
        col_base = static_cast<long long>(tap) * p.Cn + cn0;
      }
      col_lim = p.Cn;
      ptx::mbar_wait(ptx::smem_u32(&tfull_bar[buf]), aph);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN;
#pragma unroll 1
      for (int c = part * COLS_PER_WARP; c < (part + 1) * COLS_PER_WARP; c += CW) {
        float v[CW];
        if constexpr (CW == 32) ptx::tmem_ld_32x32b_x32(taddr + c, v);
        else ptx::tmem_ld_32x32b_x16(taddr + c, v);
        ptx::tmem_ld_wait();
        const int ch0 = cn0 + c;
        if (ch0 >= col_lim) continue;  // warp-uniform
        const bool full_cols = (ch0 + CW <= col_lim);
        const long long off = out_row * p.ldc + col_base + c;
        if (!row_ok) {
          // nothing to store for this lane; stay in the loop so the warp-wide TMEM loads remain converged
        } else if (MODE == CONV_WGRAD) {
          float* cp = reinterpret_cast<float*>(p.out) + off;
          if (full_cols && ((reinterpret_cast<uintptr_t>(cp) & 15) == 0)) {
#pragma unroll
            for (int j = 0; j < CW / 4; ++j)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(cp + 4 * j), "f"(v[4 * j]),
                           "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3])
                           : "memory");
          } else {
#pragma unroll
            for (int j = 0; j < CW; ++j)
              if (ch0 + j < col_lim) atomicAdd(cp + j, v[j]);
          }
        } else {
          if (MODE == CONV_FPROP) {
            if (p.bias != nullptr) {
#pragma unroll
              for (int j = 0; j < CW; ++j)
                if (ch0 + j < col_lim) v[j] += __ldg(p.bias + ch0 + j);
            }
            if (p.relu) {
#pragma unroll
              for (int j = 0; j < CW; ++j) v[j] = fmaxf(v[j], 0.0f);
            }
          }
          __nv_bfloat16* cp = reinterpret_cast<__nv_bfloat16*>(p.out) + off;
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
              if (ch0 + j < col_lim) cp[j] = __float2bfloat16_rn(v[j]);
          }
        }
        if (stats) stats_accumulate<CW>(s_stat, BN, v, row_ok, c, lane);   // every lane takes part in the shuffles
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&tempty_bar[buf]));
    }
    if (stats && stat_nt >= 0) {
      epilogue_bar_sync(EPI_THREADS);
      stats_flush(s_stat, BN, p.colstats, p.Cn, stat_nt * BN, tid_e);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

template <int BN, int MODE, bool STATS = false>
static int launch_conv(const CUtensorMap& ta, const CUtensorMap& tb, const ConvParams& p, int grid, cudaStream_t stream) {
  using Cfg = ConvCfg<BN>;
  auto kfn = conv_tcgen05_kernel<BN, MODE, STATS>;
  constexpr uint32_t SMEM = STATS ? Cfg::SMEM_BYTES_STATS : Cfg::SMEM_BYTES;
  static bool attr_set[64] = {false};
  int dev = 0;
  DET_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DET_CHECK_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr_set[dev] = true;
  }
  kfn<<<grid, CONV_THREADS, SMEM, stream>>>(ta, tb, p);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Cut the pixel grid [PN, PH, PW] into boxes of at most `maxrows` pixels: whole rows first, then whole images.
static void pick_box(int PN, int PH, int PW, int maxrows, ConvParams& p) {
  int tw = (PW + maxrows - 1) / maxrows;
  int BW = (PW + tw - 1) / tw;
  int bh_cap = maxrows / BW;
  if (bh_cap < 1) bh_cap = 1;
  int th = (PH + bh_cap - 1) / bh_cap;
  int BH = (PH + th - 1) / th;
  int BNb = 1;
  if (tw == 1 && th == 1) {
    BNb = maxrows / (BW * BH);
    if (BNb < 1) BNb = 1;
    if (BNb > PN) BNb = PN;
  }
  p.BW = BW; p.BH = BH; p.BNb = BNb;
  p.tw = tw; p.th = th; p.tn = (PN + BNb - 1) / BNb;
  p.rows_valid = BW * BH * BNb;
  p.PN = PN; p.PH = PH; p.PW = PW;
}

}  // namespace det

// C ABI ------------------------------------------------------------------------------------------
// Unit-stride convolution, NHWC bf16 activations, weight bf16 [Cout, R, S, Cin].
//   mode 0 (fprop): src = x [N,H,W,Cin],     other = weight, out = y  bf16 [N,Ho,Wo,Cout]   (+bias, relu)
//   mode 1 (dgrad): src = dy [N,Ho,Wo,Cout], other = weight, out = dx bf16 [N,H,W,Cin]
//   mode 2 (wgrad): src = dy [N,Ho,Wo,Cout], other = x,      out = dW fp32 [Cout,R,S,Cin]  (accumulated)
// Ho = H + 2*pad_h - R + 1, Wo = W + 2*pad_w - S + 1. Cin and Cout must be multiples of 8.
static int conv_launch_impl(int mode, const void* src, const void* other, void* out, int N, int H, int W,
                            int Cin, int Cout, int R, int S, int pad_h, int pad_w, const float* bias,
                            int relu, int splitk, int force_bn, float* colstats, cudaStream_t stream) {
  using namespace det;
  DET_REQUIRE(colstats == nullptr || mode == CONV_FPROP, "column statistics exist for the forward convolution only");
  DET_REQUIRE(mode >= 0 && mode <= 2, "conv mode must be 0 (fprop), 1 (dgrad) or 2 (wgrad)");
  DET_REQUIRE(N > 0 && H > 0 && W > 0 && Cin > 0 && Cout > 0 && R > 0 && S > 0, "empty convolution");
  DET_REQUIRE(Cin % 8 == 0 && Cout % 8 == 0, "implicit conv needs channel counts that are multiples of 8");
  const int Ho = H + 2 * pad_h - R + 1, Wo = W + 2 * pad_w - S + 1;
  DET_REQUIRE(Ho > 0 && Wo > 0, "convolution output is empty");
  const int sms = sm_count();
  ConvParams p;
  std::memset(&p, 0, sizeof(p));
  p.S = S; p.taps = R * S;
  p.out = out; p.bias = bias; p.relu = relu; p.colstats = colstats;
  const int Cn = (mode == CONV_FPROP) ? Cout : Cin;
  p.Cn = Cn;
  int bn = force_bn;
  if (bn <= 0) bn = Cn <= 64 ? 64 : (Cn <= 128 ? 128 : 256);
  DET_REQUIRE(bn == 64 || bn == 128 || bn == 256, "BN must be 64, 128 or 256");

  CUtensorMap ta, tb;
  int rc;
  const uint64_t xdims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  const uint64_t xstr[3] = {(uint64_t)Cin * 2, (uint64_t)W * Cin * 2, (uint64_t)H * W * Cin * 2};
  const uint64_t ydims[4] = {(uint64_t)Cout, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)N};
  const uint64_t ystr[3] = {(uint64_t)Cout * 2, (uint64_t)Wo * Cout * 2, (uint64_t)Ho * Wo * Cout * 2};
  const uint64_t wdims[3] = {(uint64_t)Cin, (uint64_t)(R * S), (uint64_t)Cout};
  const uint64_t wstr[2] = {(uint64_t)Cin * 2, (uint64_t)R * S * Cin * 2};

  if (mode == CONV_FPROP || mode == CONV_DGRAD) {
    if (mode == CONV_FPROP) {
      pick_box(N, Ho, Wo, 128, p);
      p.sgn = 1; p.off_h = -pad_h; p.off_w = -pad_w;
      p.cchunks = (Cin + 63) / 64;
      p.ldc = Cout;
    } else {
      pick_box(N, H, W, 128, p);
      p.sgn = -1; p.off_h = pad_h; p.off_w = pad_w;
      p.cchunks = (Cout + 63) / 64;
      p.ldc = Cin;
    }
    const uint32_t abox[4] = {64, (uint32_t)p.BW, (uint32_t)p.BH, (uint32_t)p.BNb};
    rc = (mode == CONV_FPROP) ? make_tma_nd_bf16(&ta, src, 4, xdims, xstr, abox)
                              : make_tma_nd_bf16(&ta, src, 4, ydims, ystr, abox);
    if (rc) return rc;
    const uint32_t wbox_f[3] = {64, 1, (uint32_t)bn};
    const uint32_t wbox_d[3] = {64, 1, 64};
    rc = make_tma_nd_bf16(&tb, other, 3, wdims, wstr, mode == CONV_FPROP ? wbox_f : wbox_d);
    if (rc) return rc;
    p.tiles_m = p.tw * p.th * p.tn;
    p.tiles_n = (Cn + bn - 1) / bn;
    p.n_per_tap = 1;
    p.num_kb = p.taps * p.cchunks;
    p.splitk = 1;
    p.kb_per_split = p.num_kb;
  } else {
    pick_box(N, Ho, Wo, 64, p);
    p.sgn = 1; p.off_h = -pad_h; p.off_w = -pad_w;
    p.Cm = Cout;
    p.ldc = static_cast<long long>(R) * S * Cin;
    const uint32_t box[4] = {64, (uint32_t)p.BW, (uint32_t)p.BH, (uint32_t)p.BNb};
    rc = make_tma_nd_bf16(&ta, src, 4, ydims, ystr, box);
    if (rc) return rc;
    rc = make_tma_nd_bf16(&tb, other, 4, xdims, xstr, box);
    if (rc) return rc;
    p.tiles_m = (Cout + CONV_BM - 1) / CONV_BM;
    p.n_per_tap = (Cin + bn - 1) / bn;
    p.tiles_n = p.taps * p.n_per_tap;
    p.num_kb = p.tw * p.th * p.tn;
    const int tiles = p.tiles_m * p.tiles_n;
    if (splitk <= 0) {
      splitk = 1;
      // largest split that keeps the work items within ONE wave of the persistent grid: aiming at two waves and rounding up
      // (9 tiles x 33 = 297 items on 148 SMs) leaves a third, nearly empty round — 228 vs 167 us on the 56x56x64 layer
      // (profiles/conv_sweep_r1.json)
      if (tiles < sms) splitk = max(1, min(p.num_kb / 4, sms / tiles));
    }
    splitk = max(1, min(splitk, p.num_kb));
    p.kb_per_split = (p.num_kb + splitk - 1) / splitk;
    p.splitk = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
  }
  const int grid = min(p.tiles_m * p.tiles_n * p.splitk, sms);
  if (colstats != nullptr) {
    if (bn == 256) return launch_conv<256, CONV_FPROP, true>(ta, tb, p, grid, stream);
    if (bn == 128) return launch_conv<128, CONV_FPROP, true>(ta, tb, p, grid, stream);
    return launch_conv<64, CONV_FPROP, true>(ta, tb, p, grid, stream);
  }
#define DET_CONV_DISPATCH(BN_)                                                              \
  if (mode == CONV_FPROP) return launch_conv<BN_, CONV_FPROP>(ta, tb, p, grid, stream);     \
  if (mode == CONV_DGRAD) return launch_conv<BN_, CONV_DGRAD>(ta, tb, p, grid, stream);     \
  return launch_conv<BN_, CONV_WGRAD>(ta, tb, p, grid, stream);
  if (bn == 256) { DET_CONV_DISPATCH(256) }
  if (bn == 128) { DET_CONV_DISPATCH(128) }
  DET_CONV_DISPATCH(64)
#undef DET_CONV_DISPATCH
}

extern "C" int det_conv2d_implicit(int mode, const void* src, const void* other, void* out, int N, int H, int W,
                                   int Cin, int Cout, int R, int S, int pad_h, int pad_w, const float* bias,
                                   int relu, int splitk, int force_bn, cudaStream_t stream) {
  return conv_launch_impl(mode, src, other, out, N, H, W, Cin, Cout, R, S, pad_h, pad_w, bias, relu, splitk, force_bn, nullptr, stream);
}

// Forward convolution that also accumulates colstats[0:Cout] += per-channel sum and colstats[Cout:2*Cout] += sum of
// squares of the bf16 output (the caller clears colstats): BatchNorm batch statistics without a pass over the output.
extern "C" int det_conv2d_fprop_stats(const void* x, const void* weight, void* y, int N, int H, int W, int Cin, int Cout, int R,
                                      int S, int pad_h, int pad_w, const float* bias, float* colstats, cudaStream_t stream) {
  return conv_launch_impl(0, x, weight, y, N, H, W, Cin, Cout, R, S, pad_h, pad_w, bias, 0, 0, 0, colstats, stream);
}
