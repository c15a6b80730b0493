// gemm_pair_sm100.cu — the CTA-pair bf16 GEMM: a 2-CTA cluster (one TPC) computes a 256 x 256 output tile with
// tcgen05.mma.cta_group::2 and writes it back through shared memory with TMA stores.
//
//   C[M,N] = epilogue( A[M,K] * B[N,K]^T )        fp32 accumulation in TMEM (2 x 256 columns: double buffered)
//
// Why a pair: the GEMMs of a transformer layer at K = 768 are bound by the L2 -> SM operand feed, not by the tensor
// pipe. A 128 x 256 single-CTA tile pulls 48 KB per 64-deep k-block (96 B / cycle / SM at the MMA's pace, 14 KB / cycle
// over 148 SMs — more than twice what the L2 delivers); in a pair each CTA stages its own 128 rows of A and HALF of the
// B tile, 32 KB per k-block, so the same MMA work needs a third less L2 traffic.
//
// Why TMA stores: tcgen05.ld leaves one output ROW per thread, so direct global stores are 32 different cache lines per
// warp instruction (the round-1 epilogue: +66 % on the FFN GEMM with bias + GELU). Here every epilogue warp packs its
// 32 rows x 32 columns into a SWIZZLE_128B staging tile (conflict-free 16-byte st.shared), and one thread per 128 x 64
// column slab issues cp.async.bulk.tensor stores: full 128-byte lines, asynchronous, ragged edges clipped by the TMA
// unit. Tiles that the epilogue has to READ (residual, GELU input for dGELU) are prefetched by TMA into the same
// staging slab while the MMAs of the tile are still running, and are overwritten in place. fp32 outputs (split-K
// weight gradients, aggregation_frequency accumulation) leave as cp.reduce.async.bulk.tensor .add boxes: the L2 does the
// read-modify-write on whole lines instead of 32 scattered red.global per warp instruction.
//
// Warp roles per CTA (576 threads): warps 0..15 = epilogue (warp w owns TMEM lane quarter w % 4 and column slab w / 4 of
// the 256 columns), warp 16 = TMA producer, warp 17 = MMA issuer (leader CTA only) + TMEM owner. The two single-thread
// roles sit on the HIGHEST warp ids on purpose: the warp scheduler favours higher ids among eligible warps, and a
// producer or MMA issuer that waits behind sixteen busy epilogue warps stalls the tensor pipe.
//
// All TMA loads here are plain (non-multicast) loads through tensor maps over this GPU's buffers.
//
// Replaces the cuBLAS calls behind nn.Linear in the reference's BERT example
// (reference: examples/nlp/bert_glue_pytorch/model_def.py:27-174 -> HF BertModel -> torch.addmm) and the apex /
// Horovod gradient accumulation kernels (reference: harness/determined/pytorch/_pytorch_context.py:470-477).
#include "gemm_common.cuh"

namespace det {

constexpr int P_STAGES = 5;   // measured: 3, 4 and 5 stages time the same at K = 768; 5 is worth 5 % at 4096^3
constexpr uint32_t P_A_BYTES = 128 * GEMM_BK * 2;   // this CTA's 128 rows of A
constexpr uint32_t P_B_BYTES = 128 * GEMM_BK * 2;   // this CTA's 128 of the tile's 256 columns of B
constexpr uint32_t P_STG_BYTES = 128 * 128;         // one staging slab: 128 rows x 128 B (64 bf16 or 32 fp32)
constexpr uint32_t P_SMEM_BYTES = P_STAGES * (P_A_BYTES + P_B_BYTES) + 4 * P_STG_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr int P_BN = 256;
constexpr int P_WARP_TMA = GEMM_EPI_WARPS;       // see "Warp roles" above
constexpr int P_WARP_MMA = GEMM_EPI_WARPS + 1;
static_assert(P_SMEM_BYTES <= 232448, "exceeds the 227 KB of shared memory a CTA can have");

// Epilogue flavours, compiled separately: a run-time switch inside the chunk loop costs issue slots in exactly the
// code that has to stay under the MMA time of a tile.
enum PairEpi : int {
  PE_PLAIN = 0,     // bf16 out: alpha, optional bias
  PE_GELU = 1,      // bf16 out: gelu(acc + bias)
  PE_GELU_AUX = 2,  // bf16 out: Z = acc + bias -> aux_out, C = gelu(Z)   (two passes over the staging slab)
  PE_DGELU = 3,     // bf16 out: acc * gelu'(aux_in)                       (aux tile prefetched by TMA)
  PE_RESID = 4,     // bf16 out: dropout(acc + bias) + aux_in              (aux tile prefetched by TMA)
  PE_F32 = 5,       // fp32 out: store, or reduce-add when out_mode == OUT_F32_ATOMIC_ADD
};

// Work decomposition shared by the three roles. Static mode: items = tiles x splitk, dealt round-robin to the clusters.
// Stream mode (fp32 reduce-add outputs only): the tiles' k-blocks form one long line that is cut into equal ranges, one
// per cluster; a range that crosses a tile boundary becomes two segments, each finished by its own reduce-add epilogue.
// Every cluster then runs the same number of k-blocks (+-1) whatever the tile count — 27 weight-gradient tiles on 74
// pairs no longer mean 54 busy pairs and 20 idle ones.
struct PairSched {
  int cluster_id, num_clusters;
  int tiles_m, num_kb, splitk, kb_per_split, total_work;
  bool stream;
  long long u, uend;   // stream mode: current / last k-block unit of this cluster
  int w;               // static mode: next work item
  __device__ __forceinline__ PairSched(const GemmParams& p, int cid, int nc)
      : cluster_id(cid), num_clusters(nc), tiles_m(p.tiles_m), num_kb(p.num_kb), splitk(p.splitk),
        kb_per_split(p.kb_per_split), total_work(p.tiles_m * p.tiles_n * p.splitk), stream(p.splitk < 0), w(cid) {
    if (stream) {
      const long long total = static_cast<long long>(p.tiles_m) * p.tiles_n * p.num_kb;
      const long long share = (total + nc - 1) / nc;
      u = share * cid;
      uend = u + share < total ? u + share : total;
    } else {
      u = uend = 0;
    }
  }
  // next segment: tile index and k-block range [kb0, kb1)
  __device__ __forceinline__ bool next(int& tile, int& kb0, int& kb1) {
    if (stream) {
      if (u >= uend) return false;
      tile = static_cast<int>(u / num_kb);
      kb0 = static_cast<int>(u - static_cast<long long>(tile) * num_kb);
      const long long left = uend - u;
      kb1 = (num_kb - kb0 < left) ? num_kb : kb0 + static_cast<int>(left);
      u += kb1 - kb0;
      return true;
    }
    if (w >= total_work) return false;
    const int ks = w % splitk;
    tile = w / splitk;
    kb0 = ks * kb_per_split;
    kb1 = min(kb0 + kb_per_split, num_kb);
    w += num_clusters;
    return true;
  }
};

// ---- per-chunk epilogue math: v[32] = accumulators of (row, col0 .. col0+31) --------------------------------------------
__device__ __forceinline__ void pe_add_bias(float (&v)[32], const float* __restrict__ bias, int col0, int N) {
  if (col0 + 32 <= N) {
    const float4* b4 = reinterpret_cast<const float4*>(bias + col0);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 b = __ldg(b4 + j);
      v[4 * j + 0] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (col0 + j < N) v[j] += __ldg(bias + col0 + j);
  }
}

// write this thread's 32 bf16 (64 B = four 16-byte units, units u0..u0+3 of its 128-byte row) into the swizzled slab
__device__ __forceinline__ void pe_store_bf16(uint8_t* stg, uint32_t row, uint32_t u0, const float (&v)[32]) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 u;
    u.x = ptx::pack_bf16x2(v[8 * j + 0], v[8 * j + 1]);
    u.y = ptx::pack_bf16x2(v[8 * j + 2], v[8 * j + 3]);
    u.z = ptx::pack_bf16x2(v[8 * j + 4], v[8 * j + 5]);
    u.w = ptx::pack_bf16x2(v[8 * j + 6], v[8 * j + 7]);
    *reinterpret_cast<uint4*>(stg + ptx::sw128_offset(row, u0 + j)) = u;
  }
}

template <int A_MN, int B_MN, int PE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_pair_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         const __grid_constant__ CUtensorMap tma_c, const __grid_constant__ CUtensorMap tma_x,
                         const GemmParams p) {
  constexpr int STAGES = P_STAGES;
  constexpr bool AUX_LOAD = (PE == PE_DGELU || PE == PE_RESID);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * P_A_BYTES;
  uint8_t* sStg = sB + STAGES * P_B_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sStg + 4 * P_STG_BYTES);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + STAGES;
  uint64_t* tfull_bar = bars + 2 * STAGES;
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;
  uint64_t* aux_bar = bars + 2 * STAGES + 4;      // [4]: one per column slab
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 8);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = ptx::cluster_ctarank();
  const bool leader = cta_rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  if (warp == P_WARP_TMA && lane == 0) {
    ptx::tma_prefetch_desc(&tma_a);
    ptx::tma_prefetch_desc(&tma_b);
    ptx::tma_prefetch_desc(&tma_c);
    if (PE == PE_GELU_AUX || AUX_LOAD) ptx::tma_prefetch_desc(&tma_x);
    for (int i = 0; i < STAGES; ++i) {
      ptx::mbar_init(ptx::smem_u32(&full_bar[i]), 1);
      ptx::mbar_init(ptx::smem_u32(&empty_bar[i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(ptx::smem_u32(&tfull_bar[i]), 1);
      ptx::mbar_init(ptx::smem_u32(&tempty_bar[i]), 2 * GEMM_EPI_WARPS);
    }
    for (int i = 0; i < 4; ++i) ptx::mbar_init(ptx::smem_u32(&aux_bar[i]), 1);
    ptx::fence_mbar_init();
  }
  if (warp == P_WARP_MMA) ptx::tmem_alloc_2cta(ptx::smem_u32(tmem_slot), 512);
  ptx::tc_fence_before();
  ptx::cluster_sync_all();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Programmatic dependent launch: everything above (descriptor prefetch, barrier init, TMEM allocation, the cluster
  // handshake) ran while the previous kernel of the stream was still draining; nothing below may start before its
  // results are visible. A no-op when the kernel was launched without the attribute.
  ptx::pdl_launch_dependents();
  ptx::pdl_wait();

  PairSched sched(p, cluster_id, num_clusters);
  int tile, kb0, kb1;

  if (warp == P_WARP_TMA) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      uint32_t it = 0;
      while (sched.next(tile, kb0, kb1)) {
        const int m0 = (tile % p.tiles_m) * 256 + static_cast<int>(cta_rank) * 128;
        const int n0 = (tile / p.tiles_m) * P_BN + static_cast<int>(cta_rank) * 128;
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const uint32_t s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          ptx::mbar_wait(ptx::smem_u32(&empty_bar[s]), ph ^ 1);
          if (leader) ptx::mbar_expect_tx(ptx::smem_u32(&full_bar[s]), 2 * (P_A_BYTES + P_B_BYTES));
          const uint32_t fb = ptx::mapa_shared(ptx::smem_u32(&full_bar[s]), 0);
          const uint32_t a_dst = ptx::smem_u32(sA + s * P_A_BYTES);
          const uint32_t b_dst = ptx::smem_u32(sB + s * P_B_BYTES);
          const int k0 = kb * GEMM_BK;
          if (A_MN == 0) {
            ptx::tma_load_2d_2cta(a_dst, &tma_a, fb, k0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < 2; ++j) ptx::tma_load_2d_2cta(a_dst + j * 8192, &tma_a, fb, m0 + 64 * j, k0);
          }
          if (B_MN == 0) {
            ptx::tma_load_2d_2cta(b_dst, &tma_b, fb, k0, n0);
          } else {
#pragma unroll
            for (int j = 0; j < 2; ++j) ptx::tma_load_2d_2cta(b_dst + j * 8192, &tma_b, fb, n0 + 64 * j, k0);
          }
        }
      }
    }
  } else if (warp == P_WARP_MMA) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && lane == 0) {
      constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(256, P_BN, A_MN, B_MN);
      uint32_t it = 0;
      uint32_t lt = 0;
      while (sched.next(tile, kb0, kb1)) {
        const uint32_t buf = lt & 1;
        const uint32_t aph = (lt >> 1) & 1;
        ++lt;
        ptx::mbar_wait(ptx::smem_u32(&tempty_bar[buf]), aph ^ 1);
        ptx::tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * P_BN;
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const uint32_t s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          ptx::mbar_wait(ptx::smem_u32(&full_bar[s]), ph);
          ptx::tc_fence_after();
          const uint32_t a_base = ptx::smem_u32(sA + s * P_A_BYTES);
          const uint32_t b_base = ptx::smem_u32(sB + s * P_B_BYTES);
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k) {
            const uint64_t adesc = A_MN ? ptx::make_smem_desc_sw128(a_base + k * 2048, 8192, 1024)
                                        : ptx::make_smem_desc_sw128(a_base + k * 32, 16, 1024);
            const uint64_t bdesc = B_MN ? ptx::make_smem_desc_sw128(b_base + k * 2048, 8192, 1024)
                                        : ptx::make_smem_desc_sw128(b_base + k * 32, 16, 1024);
            ptx::umma_bf16_ss_2cta(tmem_d, adesc, bdesc, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          ptx::umma_commit_2cta(ptx::smem_u32(&empty_bar[s]), 3);   // frees the stage in both CTAs
        }
        ptx::umma_commit_2cta(ptx::smem_u32(&tfull_bar[buf]), 3);   // accumulators ready in both CTAs
      }
    }
  } else {
    // ===================== Epilogue warps (both CTAs, each on its own 128 rows) =====================
    const int ew = warp;
    const int q = warp & 3;            // TMEM lane quarter this warp may read
    const int part = ew >> 2;          // 64-column slab of the tile
    const bool slab_leader = (ew & 3) == 0 && lane == 0;
    const int bar_id = 1 + part;
    uint8_t* stg = sStg + part * P_STG_BYTES;
    const uint32_t stg_u32 = ptx::smem_u32(stg);
    const uint32_t aux_u32 = ptx::smem_u32(&aux_bar[part]);
    const uint32_t r = static_cast<uint32_t>(q * 32 + lane);   // row inside this CTA's 128 rows
    const uint32_t leader_tempty0 = ptx::mapa_shared(ptx::smem_u32(&tempty_bar[0]), 0);
    const uint32_t leader_tempty1 = ptx::mapa_shared(ptx::smem_u32(&tempty_bar[1]), 0);
    uint32_t lt = 0;
    uint32_t aux_n = 0;                // aux tiles loaded so far by this slab (mbarrier parity)
    while (sched.next(tile, kb0, kb1)) {
      const int m0 = (tile % p.tiles_m) * 256 + static_cast<int>(cta_rank) * 128;
      const int nc0 = (tile / p.tiles_m) * P_BN + part * 64;     // first column of this slab
      const uint32_t buf = lt & 1;
      const uint32_t aph = (lt >> 1) & 1;
      ++lt;
      const bool active = m0 < p.M && nc0 < p.N;                 // uniform over the slab's four warps
      // ---- the staging slab is free once the previous tile's store has read it; start the aux prefetch
      if (active) {
        if (AUX_LOAD) {
          if (slab_leader) {
            ptx::bulk_wait_group_read0();
            ptx::mbar_expect_tx(aux_u32, P_STG_BYTES);
            ptx::tma_load_2d(stg_u32, &tma_x, aux_u32, nc0, m0);
          }
        } else {
          if (slab_leader) ptx::bulk_wait_group_read0();
          ptx::named_bar_sync(bar_id, 128);
        }
      }
      ptx::mbar_wait(ptx::smem_u32(&tfull_bar[buf]), aph);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * P_BN + part * 64;
      const uint32_t tempty = buf ? leader_tempty1 : leader_tempty0;
      if (!active) {
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive_cluster_nofence(tempty);
        continue;
      }
      const int row_g = m0 + static_cast<int>(r);
      if (AUX_LOAD) {
        ptx::mbar_wait(aux_u32, aux_n & 1);
        ++aux_n;
      }

      if (PE == PE_F32) {
        // two rounds of 32 fp32 columns (one 128-byte row each)
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          if (c == 1) {
            if (slab_leader) ptx::bulk_wait_group_read0();
            ptx::named_bar_sync(bar_id, 128);
          }
          float v[32];
          ptx::tmem_ld_32x32b_x32(taddr + c * 32, v);
          ptx::tmem_ld_wait();
          if (c == 1) {
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_cluster_nofence(tempty);
          }
          if (p.alpha != 1.0f) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] *= p.alpha;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(stg + ptx::sw128_offset(r, j)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          ptx::fence_proxy_async_smem();
          ptx::named_bar_sync(bar_id, 128);
          if (slab_leader && nc0 + c * 32 < p.N) {
            if (p.out_mode == OUT_F32_ATOMIC_ADD) ptx::tma_reduce_add_2d(&tma_c, stg_u32, nc0 + c * 32, m0);
            else ptx::tma_store_2d(&tma_c, stg_u32, nc0 + c * 32, m0);
            ptx::bulk_commit_group();
          }
        }
        continue;
      }

      if (PE == PE_GELU_AUX) {
        // pass 1: Z = acc + bias -> aux_out
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          float v[32];
          ptx::tmem_ld_32x32b_x32(taddr + c * 32, v);
          ptx::tmem_ld_wait();
          pe_add_bias(v, p.bias, nc0 + c * 32, p.N);
          pe_store_bf16(stg, r, 4 * c, v);
        }
        ptx::fence_proxy_async_smem();
        ptx::named_bar_sync(bar_id, 128);
        if (slab_leader) {
          ptx::tma_store_2d(&tma_x, stg_u32, nc0, m0);
          ptx::bulk_commit_group();
          ptx::bulk_wait_group_read0();
        }
        ptx::named_bar_sync(bar_id, 128);
      }

      // ---- bf16 output pass: two chunks of 32 columns fill the slab's 128-byte rows
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float v[32];
        ptx::tmem_ld_32x32b_x32(taddr + c * 32, v);
        ptx::tmem_ld_wait();
        if (c == 1) {               // last read of this accumulator buffer: hand it back to the MMA warp
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive_cluster_nofence(tempty);
        }
        const int col0 = nc0 + c * 32;
        if (PE == PE_PLAIN) {
          if (p.alpha != 1.0f) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] *= p.alpha;
          }
          if (p.bias != nullptr) pe_add_bias(v, p.bias, col0, p.N);
        } else if (PE == PE_GELU || PE == PE_GELU_AUX) {
          pe_add_bias(v, p.bias, col0, p.N);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
        } else if (PE == PE_DGELU) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint4 u = *reinterpret_cast<const uint4*>(stg + ptx::sw128_offset(r, 4 * c + j));
            v[8 * j + 0] *= dgelu_erf(ptx::bf16_lo(u.x)); v[8 * j + 1] *= dgelu_erf(ptx::bf16_hi(u.x));
            v[8 * j + 2] *= dgelu_erf(ptx::bf16_lo(u.y)); v[8 * j + 3] *= dgelu_erf(ptx::bf16_hi(u.y));
            v[8 * j + 4] *= dgelu_erf(ptx::bf16_lo(u.z)); v[8 * j + 5] *= dgelu_erf(ptx::bf16_hi(u.z));
            v[8 * j + 6] *= dgelu_erf(ptx::bf16_lo(u.w)); v[8 * j + 7] *= dgelu_erf(ptx::bf16_hi(u.w));
          }
        } else if (PE == PE_RESID) {
          if (p.bias != nullptr) pe_add_bias(v, p.bias, col0, p.N);
          if (p.drop_threshold != 0u) {
            const uint32_t base = static_cast<uint32_t>(row_g) * static_cast<uint32_t>(p.N) + static_cast<uint32_t>(col0);
            const uint32_t dseed = dropout_effective_seed(p.drop_seed, p.drop_seed_dev);
#pragma unroll
            for (int j = 0; j < 32; ++j)
              v[j] = dropout_keep(dseed, base + j, p.drop_threshold) ? v[j] * p.drop_scale : 0.0f;
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint4 u = *reinterpret_cast<const uint4*>(stg + ptx::sw128_offset(r, 4 * c + j));
            v[8 * j + 0] += ptx::bf16_lo(u.x); v[8 * j + 1] += ptx::bf16_hi(u.x);
            v[8 * j + 2] += ptx::bf16_lo(u.y); v[8 * j + 3] += ptx::bf16_hi(u.y);
            v[8 * j + 4] += ptx::bf16_lo(u.z); v[8 * j + 5] += ptx::bf16_hi(u.z);
            v[8 * j + 6] += ptx::bf16_lo(u.w); v[8 * j + 7] += ptx::bf16_hi(u.w);
          }
        }
        pe_store_bf16(stg, r, 4 * c, v);
      }
      ptx::fence_proxy_async_smem();
      ptx::named_bar_sync(bar_id, 128);
      if (slab_leader) {
        ptx::tma_store_2d(&tma_c, stg_u32, nc0, m0);
        ptx::bulk_commit_group();
      }
    }
    if (slab_leader) ptx::bulk_wait_group_read0();   // the slab must outlive the last store's read of it
  }

  ptx::tc_fence_before();
  ptx::cluster_sync_all();
  if (warp == P_WARP_MMA) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_2cta(tmem_base, 512);
  }
}

template <int A_MN, int B_MN, int PE>
static int launch_pair(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& tx,
                       const GemmParams& p, int grid, cudaStream_t stream) {
  auto kfn = gemm_pair_tcgen05_kernel<A_MN, B_MN, PE>;
  static bool attr_set[64] = {false};
  int dev = 0;
  DET_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DET_CHECK_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, P_SMEM_BYTES));
    attr_set[dev] = true;
  }
  static const bool pdl = [] { const char* e = getenv("DET_PDL"); return e != nullptr && e[0] == '1'; }();
  if (pdl) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(GEMM_THREADS); cfg.dynamicSmemBytes = P_SMEM_BYTES; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    DET_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kfn, ta, tb, tc, tx, p));
  } else {
    kfn<<<grid, GEMM_THREADS, P_SMEM_BYTES, stream>>>(ta, tb, tc, tx, p);
  }
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <int PE>
static int launch_pair_majors(int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
                              const CUtensorMap& tx, const GemmParams& p, int grid, cudaStream_t stream) {
  if (a_mn == 0 && b_mn == 0) return launch_pair<0, 0, PE>(ta, tb, tc, tx, p, grid, stream);
  if (a_mn == 0 && b_mn == 1) return launch_pair<0, 1, PE>(ta, tb, tc, tx, p, grid, stream);
  if (a_mn == 1 && b_mn == 0) return launch_pair<1, 0, PE>(ta, tb, tc, tx, p, grid, stream);
  return launch_pair<1, 1, PE>(ta, tb, tc, tx, p, grid, stream);
}

bool pair_gemm_supported(const GemmParams& p, const void* A, const void* B, long long lda, long long ldb) {
  const uintptr_t c = reinterpret_cast<uintptr_t>(p.C);
  if (p.colstats != nullptr) return false;
  if (p.out_mode == OUT_BF16) {
    if ((c & 15) != 0 || (p.ldc % 8) != 0) return false;
  } else {
    if ((c & 15) != 0 || (p.ldc % 4) != 0) return false;
    if (p.epi != EPI_NONE || p.bias != nullptr) return false;
  }
  if (p.epi == EPI_BIAS_GELU && p.bias == nullptr) return false;
  const __nv_bfloat16* aux = p.aux_in != nullptr ? p.aux_in : p.aux_out;
  if (aux != nullptr && ((reinterpret_cast<uintptr_t>(aux) & 15) != 0 || (p.ld_aux % 8) != 0)) return false;
  if (p.epi == EPI_BIAS_RESID && p.out_mode != OUT_BF16) return false;
  (void)A; (void)B; (void)lda; (void)ldb;
  return true;
}

// p is complete except tiles_m / tiles_n / splitk / kb_per_split, which are (re)computed here for the 256 x 256 pair tile.
// splitk: 0 = automatic, > 0 = that many K slices, < 0 = stream-K (fp32 reduce-add outputs only).
int launch_pair_gemm(const CUtensorMap& ta, const CUtensorMap& tb, GemmParams p, int a_mn, int b_mn, int splitk,
                     cudaStream_t stream) {
  const int units = max(1, sm_count() / 2);
  p.tiles_m = (p.M + 255) / 256;
  p.tiles_n = (p.N + P_BN - 1) / P_BN;
  p.num_kb = (p.K + GEMM_BK - 1) / GEMM_BK;
  const int tiles = p.tiles_m * p.tiles_n;
  int work;
  if (p.out_mode != OUT_F32_ATOMIC_ADD) {
    p.splitk = 1;
    p.kb_per_split = p.num_kb;
    work = tiles;
  } else {
    static const int stream_auto = [] { const char* e = getenv("DET_GEMM_STREAMK"); return (e && e[0] == '0') ? 0 : 1; }();
    static const int one_wave_auto = [] { const char* e = getenv("DET_GEMM_ONEWAVE"); return (e && e[0] == '0') ? 0 : 1; }();
    if (splitk == 0 && one_wave_auto && tiles < units) {
      // a static split that fills one wave almost exactly beats stream-K: every pair runs ONE segment with one epilogue
      // instead of two partial tiles (3072 x 768 x 6144: 25.2 us against 29.0; profiles/gemm_layout_probe_r2.json)
      const int s1 = units / tiles;
      if (s1 >= 2 && s1 <= 4 && tiles * s1 * 20 >= units * 19 && p.num_kb / s1 >= 8) splitk = s1;
    }
    if (splitk == 0) {
      // whole tiles when they fill the pairs evenly, otherwise cut the k-block line into equal ranges
      const long long units_kb = static_cast<long long>(tiles) * p.num_kb;
      const bool even = tiles >= units && tiles % units == 0;
      splitk = (!even && stream_auto && units_kb >= 4LL * units) ? -1 : 0;
      if (splitk == 0) {
        splitk = 1;
        if (tiles < units) splitk = max(1, min(p.num_kb / 4, units / tiles));
      }
    }
    if (splitk < 0) {
      p.splitk = -1;
      p.kb_per_split = p.num_kb;
      const long long units_kb = static_cast<long long>(tiles) * p.num_kb;
      work = static_cast<int>(units_kb < units ? units_kb : units);
    } else {
      splitk = max(1, min(splitk, p.num_kb));
      p.kb_per_split = (p.num_kb + splitk - 1) / splitk;
      p.splitk = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
      work = tiles * p.splitk;
    }
  }
  const int grid = 2 * min(work, units);

  CUtensorMap tc, tx;
  int rc;
  if (p.out_mode == OUT_BF16) rc = make_tma_2d_bf16(&tc, p.C, (uint64_t)p.N, (uint64_t)p.M, (uint64_t)p.ldc, 64, 128);
  else rc = make_tma_2d_f32(&tc, p.C, (uint64_t)p.N, (uint64_t)p.M, (uint64_t)p.ldc, 32, 128);
  if (rc) return rc;
  const __nv_bfloat16* aux = p.aux_in != nullptr ? p.aux_in : p.aux_out;
  const bool uses_aux = (p.epi == EPI_DGELU || p.epi == EPI_BIAS_RESID || (p.epi == EPI_BIAS_GELU && p.aux_out != nullptr));
  if (uses_aux) {
    rc = make_tma_2d_bf16(&tx, aux, (uint64_t)p.N, (uint64_t)p.M, (uint64_t)p.ld_aux, 64, 128);
    if (rc) return rc;
  } else {
    tx = tc;
  }
  if (p.out_mode != OUT_BF16) return launch_pair_majors<PE_F32>(a_mn, b_mn, ta, tb, tc, tx, p, grid, stream);
  switch (p.epi) {
    case EPI_BIAS_GELU:
      if (p.aux_out != nullptr) return launch_pair_majors<PE_GELU_AUX>(a_mn, b_mn, ta, tb, tc, tx, p, grid, stream);
      return launch_pair_majors<PE_GELU>(a_mn, b_mn, ta, tb, tc, tx, p, grid, stream);
    case EPI_DGELU:
      return launch_pair_majors<PE_DGELU>(a_mn, b_mn, ta, tb, tc, tx, p, grid, stream);
    case EPI_BIAS_RESID:
      return launch_pair_majors<PE_RESID>(a_mn, b_mn, ta, tb, tc, tx, p, grid, stream);
    default:
      return launch_pair_majors<PE_PLAIN>(a_mn, b_mn, ta, tb, tc, tx, p, grid, stream);
  }
}

}  // namespace det
