// attention_flash_sm100.cu — multi-head self-attention for sequences LONGER than 128 tokens (head_dim 64):
// the tiled, online-softmax form of attention_sm100.cu, same packed-QKV interface, same dropout hash.
//
//   forward : one CTA per (batch, head, 128-query tile); K/V tiles of 128 keys stream through shared memory by TMA.
//             Per key tile: S = Q K^T on tcgen05 -> running max / running sum in registers -> P (bf16, swizzled smem)
//             -> O_j = P V on tcgen05 -> accumulated in registers with the usual exp2(m_old - m_new) rescale.
//             K_{j+1} is fetched while the softmax of tile j runs, V_{j+1} while S_{j+1} is computed; two CTAs
//             share an SM so one CTA's softmax overlaps the other's tensor work.
//   backward: one CTA per (batch, head, 128-key tile) keeps K, V resident and dK, dV in TMEM across the loop over
//             query tiles; per query tile S and dP' are recomputed, P' and dS go to smem once and feed
//             dV += P'^T dO, dK += dS^T Q, dQ_tile = dS K. dQ tiles are summed over key tiles with
//             red.global.add.v4.f32 into an fp32 scratch and converted to bf16 by a last small kernel.
//             delta = rowsum(dO * O) comes from a pre-pass over the saved forward output.
//
// Used by the SQuAD-shape BERT / ALBERT examples (seq 384; reference: examples/nlp/bert_squad_pytorch/model_def.py,
// examples/nlp/albert_squad_pytorch/model_def.py, which reach attention through HF BertSelfAttention → cuBLAS +
// softmax/dropout kernels).
#include "host_utils.h"
#include "sm100_ptx.cuh"

namespace det {

struct FlashParams {
  int B, S, H, hidden;
  int nq, nkv;             // 128-row tiles per sequence (both ceil(S / 128))
  float scale;
  const float* mask;       // additive key mask [B, S] fp32, or null
  float* lse;              // [B, H, S] fp32 (natural log)
  const float* delta;      // bwd: [B, H, S] fp32
  __nv_bfloat16* out;      // fwd: ctx [T, hidden]; bwd: dqkv [T, 3*hidden]
  float* dq_accum;         // bwd: fp32 [T, hidden], zeroed
  uint32_t drop_threshold;
  float drop_scale;
  uint32_t drop_seed;
  const uint32_t* drop_seed_dev;
};

constexpr int FA_T = 128;                           // tile rows (queries or keys)
constexpr int FA_D = 64;
constexpr uint32_t FA_TILE = FA_T * FA_D * 2;       // 16 KB
constexpr float FA_LOG2E = 1.4426950408889634f;

__device__ __forceinline__ uint4 fa_pack8(const float* v) {
  uint4 pk;
  pk.x = ptx::pack_bf16x2(v[0], v[1]); pk.y = ptx::pack_bf16x2(v[2], v[3]);
  pk.z = ptx::pack_bf16x2(v[4], v[5]); pk.w = ptx::pack_bf16x2(v[6], v[7]);
  return pk;
}

// ------------------------------------------------------------------------------------------------
// Forward: 256 threads = 4 TMEM lane quarters x 2 key halves (64 keys each); 2 CTAs per SM.
// ------------------------------------------------------------------------------------------------
constexpr int FA_FWD_THREADS = 256;
constexpr uint32_t FA_FWD_RED = 2 * 2 * FA_T * 4 + FA_T * 4;
constexpr uint32_t FA_FWD_SMEM = 5 * FA_TILE + 1024 + 64 + FA_FWD_RED;

__global__ void __launch_bounds__(FA_FWD_THREADS, 2)
flash_fwd_kernel(const __grid_constant__ CUtensorMap tma_qkv, const FlashParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + FA_TILE;
  uint8_t* sV = smem + 2 * FA_TILE;
  uint8_t* sP = smem + 3 * FA_TILE;   // 2 chunks of [128 q][64 keys]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 5 * FA_TILE);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);
  float* red_max = reinterpret_cast<float*>(smem + 5 * FA_TILE + 64);   // [2][128]
  float* red_sum = red_max + 2 * FA_T;                                  // [2][128]
  float* smask = red_sum + 2 * FA_T;                                    // [128]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int qt = blockIdx.x % p.nq;
  const int bh = blockIdx.x / p.nq;
  const int b = bh / p.H, h = bh % p.H;
  const uint32_t k_bar = ptx::smem_u32(&bars[0]);
  const uint32_t v_bar = ptx::smem_u32(&bars[1]);
  const uint32_t mma_bar = ptx::smem_u32(&bars[2]);

  if (tid == 0) {
    ptx::tma_prefetch_desc(&tma_qkv);
    ptx::mbar_init(k_bar, 1);
    ptx::mbar_init(v_bar, 1);
    ptx::mbar_init(mma_bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 256);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem;          // 128 columns
  const uint32_t tO = tmem + 128;    // 64 columns

  const int row0 = b * p.S;
  if (tid == 0) {
    ptx::mbar_expect_tx(k_bar, 2 * FA_TILE);
    ptx::tma_load_2d(ptx::smem_u32(sQ), &tma_qkv, k_bar, h * FA_D, row0 + qt * FA_T);
    ptx::tma_load_2d(ptx::smem_u32(sK), &tma_qkv, k_bar, p.hidden + h * FA_D, row0);
    ptx::mbar_expect_tx(v_bar, FA_TILE);
    ptx::tma_load_2d(ptx::smem_u32(sV), &tma_qkv, v_bar, 2 * p.hidden + h * FA_D, row0);
  }

  const int q = (warp & 3) * 32 + lane;   // query row inside the tile == TMEM lane
  const int cg = warp >> 2;               // key half
  const int qg = qt * FA_T + q;           // query position in the sequence
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const float sc = p.scale * FA_LOG2E;
  const uint32_t dseed = dropout_effective_seed(p.drop_seed, p.drop_seed_dev);
  const uint32_t drow = ((static_cast<uint32_t>(b) * p.H + h) * p.S + qg) * p.S;

  float o[32];
#pragma unroll
  for (int e = 0; e < 32; ++e) o[e] = 0.0f;
  float m_run = -INFINITY, l_run = 0.0f;

  for (int j = 0; j < p.nkv; ++j) {
    if (tid < FA_T) {
      const int key = j * FA_T + tid;
      smask[tid] = key < p.S ? (p.mask != nullptr ? __ldg(p.mask + b * p.S + key) * FA_LOG2E : 0.0f) : -INFINITY;
    }
    if (tid == 0) {
      ptx::mbar_wait(k_bar, j & 1);
      ptx::tc_fence_after();
      constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(128, 128, 0, 0);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sQ) + k * 32, 16, 1024);
        const uint64_t bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sK) + k * 32, 16, 1024);
        ptx::umma_bf16_ss(tS, ad, bd, idesc, k > 0);
      }
      ptx::umma_commit(mma_bar);
    }
    __syncthreads();   // key mask staged
    ptx::mbar_wait(mma_bar, 0);
    ptx::tc_fence_after();
    if (tid == 0 && j + 1 < p.nkv) {   // S is in TMEM: the K buffer is free for the next tile
      ptx::mbar_expect_tx(k_bar, FA_TILE);
      ptx::tma_load_2d(ptx::smem_u32(sK), &tma_qkv, k_bar, p.hidden + h * FA_D, row0 + (j + 1) * FA_T);
    }

    // pass 1: row maximum over this thread's 64 keys (two TMEM reads of 32 columns keep registers low)
    float mx = -INFINITY;
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      float s[32];
      ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 64 + hf * 32, s);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 32; ++e) mx = fmaxf(mx, fmaf(s[e], sc, smask[cg * 64 + hf * 32 + e]));
    }
    red_max[cg * FA_T + q] = mx;
    __syncthreads();
    const float m_new = fmaxf(m_run, fmaxf(red_max[q], red_max[FA_T + q]));
    const float m_use = (m_new == -INFINITY) ? 0.0f : m_new;          // a fully masked prefix keeps everything at zero
    const float alpha = (m_run == -INFINITY) ? 0.0f : fast_exp2(m_run - m_use);
    m_run = m_new;

    // pass 2: probabilities, running sum, P' -> smem
    float lsum = 0.0f;
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      float s[32];
      ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 64 + hf * 32, s);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        s[e] = fast_exp2(fmaf(s[e], sc, smask[cg * 64 + hf * 32 + e]) - m_use);
        lsum += s[e];
      }
      if (p.drop_threshold != 0u) {
        const uint32_t didx0 = drow + j * FA_T + cg * 64 + hf * 32;
#pragma unroll
        for (int e = 0; e < 32; ++e) s[e] = dropout_keep(dseed, didx0 + e, p.drop_threshold) ? s[e] * p.drop_scale : 0.0f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        *reinterpret_cast<uint4*>(sP + cg * FA_TILE + ptx::sw128_offset(q, hf * 4 + u)) = fa_pack8(s + 8 * u);
    }
    l_run = fmaf(l_run, alpha, lsum);
#pragma unroll
    for (int e = 0; e < 32; ++e) o[e] *= alpha;
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();

    if (tid == 0) {
      ptx::mbar_wait(v_bar, j & 1);
      ptx::tc_fence_after();
      constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(128, 64, 0, 1);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sP) + (k >> 2) * FA_TILE + (k & 3) * 32, 16, 1024);
        const uint64_t bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sV) + k * 2048, 8192, 1024);
        ptx::umma_bf16_ss(tO, ad, bd, idesc, k > 0);
      }
      ptx::umma_commit(mma_bar);
    }
    ptx::mbar_wait(mma_bar, 1);
    ptx::tc_fence_after();
    if (tid == 0 && j + 1 < p.nkv) {
      ptx::mbar_expect_tx(v_bar, FA_TILE);
      ptx::tma_load_2d(ptx::smem_u32(sV), &tma_qkv, v_bar, 2 * p.hidden + h * FA_D, row0 + (j + 1) * FA_T);
    }
    {
      float t[32];
      ptx::tmem_ld_32x32b_x32(tO + lane_base + cg * 32, t);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 32; ++e) o[e] += t[e];
    }
    ptx::tc_fence_before();
    __syncthreads();   // S / O / P / exchange buffers are free for the next key tile
  }

  red_sum[cg * FA_T + q] = l_run;
  __syncthreads();
  const float l = red_sum[q] + red_sum[FA_T + q];
  const float inv = l > 0.0f ? 1.0f / l : 0.0f;
  if (qg < p.S) {
    if (cg == 0) p.lse[(static_cast<long long>(b) * p.H + h) * p.S + qg] = (m_run + log2f(l)) * (1.0f / FA_LOG2E);
#pragma unroll
    for (int e = 0; e < 32; ++e) o[e] *= inv;
    uint4* dst = reinterpret_cast<uint4*>(p.out + (static_cast<long long>(row0) + qg) * p.hidden + h * FA_D + cg * 32);
#pragma unroll
    for (int u = 0; u < 4; ++u) dst[u] = fa_pack8(o + 8 * u);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 256);
  }
}

// ------------------------------------------------------------------------------------------------
// Backward pre-pass: delta[b, h, s] = sum_d dO[s, h, d] * O[s, h, d]; a thread per 16-byte vector, 8 lanes per head.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
flash_delta_kernel(const __nv_bfloat16* __restrict__ dctx, const __nv_bfloat16* __restrict__ ctx,
                   float* __restrict__ delta, int B, int S, int H) {
  const long long v = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;   // vector index over [T, H*64] / 8
  const long long total = static_cast<long long>(B) * S * H * 8;
  float acc = 0.0f;
  if (v < total) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(dctx) + v);
    const uint4 c = __ldg(reinterpret_cast<const uint4*>(ctx) + v);
    acc = ptx::bf16_lo(a.x) * ptx::bf16_lo(c.x) + ptx::bf16_hi(a.x) * ptx::bf16_hi(c.x) +
          ptx::bf16_lo(a.y) * ptx::bf16_lo(c.y) + ptx::bf16_hi(a.y) * ptx::bf16_hi(c.y) +
          ptx::bf16_lo(a.z) * ptx::bf16_lo(c.z) + ptx::bf16_hi(a.z) * ptx::bf16_hi(c.z) +
          ptx::bf16_lo(a.w) * ptx::bf16_lo(c.w) + ptx::bf16_hi(a.w) * ptx::bf16_hi(c.w);
  }
  acc += __shfl_xor_sync(0xffffffffu, acc, 4);
  acc += __shfl_xor_sync(0xffffffffu, acc, 2);
  acc += __shfl_xor_sync(0xffffffffu, acc, 1);
  if (v < total && (threadIdx.x & 7) == 0) {
    const long long w = v >> 3;                 // (token, head)
    const long long tok = w / H;
    const int h = static_cast<int>(w % H);
    delta[((tok / S) * H + h) * S + tok % S] = acc;
  }
}

// dq_accum fp32 [T, hidden] -> the Q third of dqkv (bf16, row stride 3 * hidden)
__global__ void flash_dq_convert_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dqkv, long long T, int hidden) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;   // one thread per 8 elements
  const int per_row = hidden / 8;
  if (i >= T * per_row) return;
  const long long row = i / per_row;
  const int c = static_cast<int>(i % per_row) * 8;
  const float4 a = *reinterpret_cast<const float4*>(acc + row * hidden + c);
  const float4 b = *reinterpret_cast<const float4*>(acc + row * hidden + c + 4);
  const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  *reinterpret_cast<uint4*>(dqkv + row * (3LL * hidden) + c) = fa_pack8(v);
}

// ------------------------------------------------------------------------------------------------
// Backward: 512 threads = 4 lane quarters x 4 column groups (32 keys each); 1 CTA per SM; one CTA per key tile.
//
// Software pipeline over the query tiles (Q/dO in a 3-stage TMA ring). Per tile i:
//   A  wait S_i, dP'_i (MMA1)        -> registers;            barrier: TMEM S/dP free
//   B  issue MMA1 of tile i+1        (tensor pipe works while the CUDA cores do C)
//   C  P'_i, dS_i in registers;  wait MMA2 of tile i-1;  refill the ring stage it freed;  dQ_{i-1} -> red.add;
//      P'_i, dS_i -> smem;                                    barrier: smem operands visible
//   D  issue MMA2 of tile i:  dV += P'^T dO,  dK += dS^T Q,  dQ_i = dS K
// ------------------------------------------------------------------------------------------------
constexpr int FA_BWD_THREADS = 512;
constexpr int FA_BWD_STAGES = 3;
constexpr uint32_t FA_BWD_SMEM = (6 + 2 * FA_BWD_STAGES) * FA_TILE + 1024 + 64 + FA_T * 4;

__device__ __forceinline__ void fa_dq_flush(const FlashParams& p, uint32_t tdQ, uint32_t lane_base, int cg, int row0, int h, int qg) {
  float d[16];
  ptx::tmem_ld_32x32b_x16(tdQ + lane_base + cg * 16, d);
  ptx::tmem_ld_wait();
  if (qg < p.S) {
    float* dst = p.dq_accum + (static_cast<long long>(row0) + qg) * p.hidden + h * FA_D + cg * 16;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4 * u), "f"(d[4 * u]), "f"(d[4 * u + 1]),
                   "f"(d[4 * u + 2]), "f"(d[4 * u + 3]) : "memory");
  }
}

__global__ void __launch_bounds__(FA_BWD_THREADS, 1)
flash_bwd_kernel(const __grid_constant__ CUtensorMap tma_qkv, const __grid_constant__ CUtensorMap tma_do,
                 const FlashParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = smem + FA_TILE;
  uint8_t* sP = smem + 2 * FA_TILE;    // 2 chunks
  uint8_t* sdS = smem + 4 * FA_TILE;   // 2 chunks
  uint8_t* sQ = smem + 6 * FA_TILE;    // FA_BWD_STAGES tiles
  uint8_t* sdO = sQ + FA_BWD_STAGES * FA_TILE;
  uint8_t* tail = sdO + FA_BWD_STAGES * FA_TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);   // kv, mma1, mma2, q[STAGES]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 + FA_BWD_STAGES);
  float* smask = reinterpret_cast<float*>(tail + 64);   // [128], this CTA's keys

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kt = blockIdx.x % p.nkv;
  const int bh = blockIdx.x / p.nkv;
  const int b = bh / p.H, h = bh % p.H;
  const uint32_t kv_bar = ptx::smem_u32(&bars[0]);
  const uint32_t bar1 = ptx::smem_u32(&bars[1]);
  const uint32_t bar2 = ptx::smem_u32(&bars[2]);
  const uint32_t q_bar0 = ptx::smem_u32(&bars[3]);

  if (tid < FA_T) {
    const int key = kt * FA_T + tid;
    smask[tid] = key < p.S ? (p.mask != nullptr ? __ldg(p.mask + b * p.S + key) * FA_LOG2E : 0.0f) : -INFINITY;
  }
  if (tid == 0) {
    ptx::tma_prefetch_desc(&tma_qkv);
    ptx::tma_prefetch_desc(&tma_do);
    for (int i = 0; i < 3 + FA_BWD_STAGES; ++i) ptx::mbar_init(ptx::smem_u32(&bars[i]), 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tdP = tmem + 128, tdV = tmem + 256, tdQ = tmem + 320, tdK = tmem + 384;

  const int row0 = b * p.S;
  auto load_q = [&](int tile) {   // tid 0 only
    const int st = tile % FA_BWD_STAGES;
    const uint32_t qb = q_bar0 + 8 * st;
    ptx::mbar_expect_tx(qb, 2 * FA_TILE);
    ptx::tma_load_2d(ptx::smem_u32(sQ + st * FA_TILE), &tma_qkv, qb, h * FA_D, row0 + tile * FA_T);
    ptx::tma_load_2d(ptx::smem_u32(sdO + st * FA_TILE), &tma_do, qb, h * FA_D, row0 + tile * FA_T);
  };
  auto issue_mma1 = [&](int tile) {   // tid 0 only: S = Q K^T, dP' = dO V^T
    const int st = tile % FA_BWD_STAGES;
    ptx::mbar_wait(q_bar0 + 8 * st, (tile / FA_BWD_STAGES) & 1);
    ptx::tc_fence_after();
    constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(128, 128, 0, 0);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sQ + st * FA_TILE) + k * 32, 16, 1024);
      const uint64_t bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sK) + k * 32, 16, 1024);
      ptx::umma_bf16_ss(tS, ad, bd, idesc, k > 0);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sdO + st * FA_TILE) + k * 32, 16, 1024);
      const uint64_t bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sV) + k * 32, 16, 1024);
      ptx::umma_bf16_ss(tdP, ad, bd, idesc, k > 0);
    }
    ptx::umma_commit(bar1);
  };

  if (tid == 0) {
    ptx::mbar_expect_tx(kv_bar, 2 * FA_TILE);
    ptx::tma_load_2d(ptx::smem_u32(sK), &tma_qkv, kv_bar, p.hidden + h * FA_D, row0 + kt * FA_T);
    ptx::tma_load_2d(ptx::smem_u32(sV), &tma_qkv, kv_bar, 2 * p.hidden + h * FA_D, row0 + kt * FA_T);
    for (int t = 0; t < FA_BWD_STAGES && t < p.nq; ++t) load_q(t);
    ptx::mbar_wait(kv_bar, 0);
    issue_mma1(0);
  }

  const int q = (warp & 3) * 32 + lane;
  const int cg = warp >> 2;   // keys [32*cg, 32*cg + 32) of the tile
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const float sc = p.scale * FA_LOG2E;
  const uint32_t dseed = dropout_effective_seed(p.drop_seed, p.drop_seed_dev);
  const long long stat_base = (static_cast<long long>(b) * p.H + h) * p.S;

  for (int i = 0; i < p.nq; ++i) {
    const int st = i % FA_BWD_STAGES;
    const int qg = i * FA_T + q;
    const bool q_ok = qg < p.S;
    const float lse2 = q_ok ? __ldg(p.lse + stat_base + qg) * FA_LOG2E : 0.0f;
    const float delta = q_ok ? __ldg(p.delta + stat_base + qg) : 0.0f;
    const uint32_t didx0 = ((static_cast<uint32_t>(b) * p.H + h) * p.S + qg) * p.S + kt * FA_T + cg * 32;

    // ---- A: scores of tile i into registers ----
    float pr[32], dp[32];
    ptx::mbar_wait(bar1, i & 1);
    ptx::tc_fence_after();
    ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 32, pr);
    ptx::tmem_ld_32x32b_x32(tdP + lane_base + cg * 32, dp);
    ptx::tmem_ld_wait();
    ptx::tc_fence_before();
    __syncthreads();

    // ---- B: the tensor pipe starts on tile i+1 while this tile's softmax math runs ----
    if (tid == 0 && i + 1 < p.nq) issue_mma1(i + 1);

    // ---- C: P' and dS (packed to bf16 right away to keep registers for the dQ read-out below) ----
    uint4 pk_p[4], pk_s[4];
    {
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        const float pv = q_ok ? fast_exp2(fmaf(pr[e], sc, smask[cg * 32 + e]) - lse2) : 0.0f;
        float keep = 1.0f;
        if (p.drop_threshold != 0u) keep = dropout_keep(dseed, didx0 + e, p.drop_threshold) ? p.drop_scale : 0.0f;
        pr[e] = pv * keep;                                // P'
        dp[e] = pv * (dp[e] * keep - delta) * p.scale;    // dS
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) { pk_p[u] = fa_pack8(pr + 8 * u); pk_s[u] = fa_pack8(dp + 8 * u); }
    }
    if (i > 0) {
      ptx::mbar_wait(bar2, (i - 1) & 1);   // MMA2 of tile i-1 done: sP / sdS / its ring stage are free, dQ_{i-1} is in TMEM
      ptx::tc_fence_after();
      if (tid == 0 && i + 2 < p.nq) load_q(i + 2);   // (i + 2) % 3 == (i - 1) % 3
      fa_dq_flush(p, tdQ, lane_base, cg, row0, h, qg - FA_T);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t gu = cg * 4 + u;   // 16-byte unit 0..15 across the 128 keys
      *reinterpret_cast<uint4*>(sP + (gu >> 3) * FA_TILE + ptx::sw128_offset(q, gu & 7)) = pk_p[u];
      *reinterpret_cast<uint4*>(sdS + (gu >> 3) * FA_TILE + ptx::sw128_offset(q, gu & 7)) = pk_s[u];
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();

    // ---- D: accumulate into dV, dK; fresh dQ tile ----
    if (tid == 0) {
      ptx::tc_fence_after();
      constexpr uint32_t idesc_tn = ptx::make_idesc_bf16_f32(128, 64, 1, 1);
      constexpr uint32_t idesc_nn = ptx::make_idesc_bf16_f32(128, 64, 0, 1);
      const uint32_t q_s = ptx::smem_u32(sQ + st * FA_TILE), do_s = ptx::smem_u32(sdO + st * FA_TILE);
#pragma unroll
      for (int k = 0; k < 8; ++k) {   // dV[key, d] += sum_q P'[q, key] dO[q, d]
        const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sP) + k * 2048, FA_TILE, 1024);
        const uint64_t bd = ptx::make_smem_desc_sw128(do_s + k * 2048, 8192, 1024);
        ptx::umma_bf16_ss(tdV, ad, bd, idesc_tn, (i > 0 || k > 0) ? 1u : 0u);
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {   // dK[key, d] += sum_q dS[q, key] Q[q, d]
        const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sdS) + k * 2048, FA_TILE, 1024);
        const uint64_t bd = ptx::make_smem_desc_sw128(q_s + k * 2048, 8192, 1024);
        ptx::umma_bf16_ss(tdK, ad, bd, idesc_tn, (i > 0 || k > 0) ? 1u : 0u);
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {   // dQ_tile[q, d] = sum_key dS[q, key] K[key, d]
        const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sdS) + (k >> 2) * FA_TILE + (k & 3) * 32, 16, 1024);
        const uint64_t bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sK) + k * 2048, 8192, 1024);
        ptx::umma_bf16_ss(tdQ, ad, bd, idesc_nn, k > 0);
      }
      ptx::umma_commit(bar2);
    }
  }

  ptx::mbar_wait(bar2, (p.nq - 1) & 1);
  ptx::tc_fence_after();
  fa_dq_flush(p, tdQ, lane_base, cg, row0, h, (p.nq - 1) * FA_T + q);
  // dK, dV of this key tile: 128 columns [dK | dV], column group cg takes 32 of them.
  {
    const int key = kt * FA_T + q;
    const int which = cg >> 1;            // 0 dK, 1 dV
    const int col = (cg & 1) * 32;
    float o[32];
    ptx::tmem_ld_32x32b_x32((which == 0 ? tdK : tdV) + lane_base + col, o);
    ptx::tmem_ld_wait();
    if (key < p.S) {
      uint4* dst = reinterpret_cast<uint4*>(p.out + (static_cast<long long>(row0) + key) * (3LL * p.hidden) +
                                            static_cast<long long>(which + 1) * p.hidden + h * FA_D + col);
#pragma unroll
      for (int u = 0; u < 4; ++u) dst[u] = fa_pack8(o + 8 * u);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 512);
  }
}

static int fa_fill(FlashParams& p, int B, int S, int H, const float* mask, float* lse, void* out, float dropout_p,
                   unsigned seed, const unsigned* seed_dev) {
  DET_REQUIRE(B >= 1 && S >= 1 && H >= 1, "bad attention shape");
  DET_REQUIRE(static_cast<long long>(B) * S < (1LL << 31), "too many tokens");
  p.B = B; p.S = S; p.H = H; p.hidden = H * FA_D;
  p.nq = p.nkv = (S + FA_T - 1) / FA_T;
  p.scale = 0.125f;
  p.mask = mask; p.lse = lse; p.delta = nullptr; p.dq_accum = nullptr;
  p.out = reinterpret_cast<__nv_bfloat16*>(out);
  if (dropout_p > 0.0f) {
    DET_REQUIRE(dropout_p < 1.0f, "dropout p must be < 1");
    p.drop_threshold = static_cast<uint32_t>(static_cast<double>(dropout_p) * 4294967296.0);
    p.drop_scale = 1.0f / (1.0f - dropout_p);
  } else {
    p.drop_threshold = 0; p.drop_scale = 1.0f;
  }
  p.drop_seed = seed;
  p.drop_seed_dev = seed_dev;
  return 0;
}

}  // namespace det

// Same tensors as det_attention_fwd, any sequence length.
extern "C" int det_attention_flash_fwd(const void* qkv, const float* mask, void* ctx, float* lse, int B, int S, int H,
                                       float dropout_p, unsigned int seed, const unsigned int* seed_dev,
                                       cudaStream_t stream) {
  using namespace det;
  FlashParams p;
  int rc = fa_fill(p, B, S, H, mask, lse, ctx, dropout_p, seed, seed_dev);
  if (rc) return rc;
  CUtensorMap tq;
  rc = make_tma_2d_bf16(&tq, qkv, 3ull * p.hidden, (uint64_t)B * S, 3ull * p.hidden, 64, FA_T);
  if (rc) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  DET_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DET_CHECK_CUDA(cudaFuncSetAttribute(flash_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FA_FWD_SMEM));
    attr_set[dev] = true;
  }
  flash_fwd_kernel<<<B * H * p.nq, FA_FWD_THREADS, FA_FWD_SMEM, stream>>>(tq, p);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ctx: the forward output (for delta); dq_accum: fp32 scratch [B*S, H*64]; delta: fp32 scratch [B, H, S].
extern "C" int det_attention_flash_bwd(const void* qkv, const void* dctx, const void* ctx, const float* mask,
                                       const float* lse, void* dqkv, float* dq_accum, float* delta, int B, int S, int H,
                                       float dropout_p, unsigned int seed, const unsigned int* seed_dev,
                                       cudaStream_t stream) {
  using namespace det;
  FlashParams p;
  int rc = fa_fill(p, B, S, H, mask, const_cast<float*>(lse), dqkv, dropout_p, seed, seed_dev);
  if (rc) return rc;
  p.delta = delta;
  p.dq_accum = dq_accum;
  const long long T = static_cast<long long>(B) * S;
  CUtensorMap tq, td;
  rc = make_tma_2d_bf16(&tq, qkv, 3ull * p.hidden, (uint64_t)T, 3ull * p.hidden, 64, FA_T);
  if (rc) return rc;
  rc = make_tma_2d_bf16(&td, dctx, (uint64_t)p.hidden, (uint64_t)T, (uint64_t)p.hidden, 64, FA_T);
  if (rc) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  DET_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DET_CHECK_CUDA(cudaFuncSetAttribute(flash_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FA_BWD_SMEM));
    attr_set[dev] = true;
  }
  DET_CHECK_CUDA(cudaMemsetAsync(dq_accum, 0, sizeof(float) * T * p.hidden, stream));
  {
    const long long vecs = T * H * 8;
    const int threads = 256;
    flash_delta_kernel<<<static_cast<unsigned>((vecs + threads - 1) / threads), threads, 0, stream>>>(
        reinterpret_cast<const __nv_bfloat16*>(dctx), reinterpret_cast<const __nv_bfloat16*>(ctx), delta, B, S, H);
    DET_CHECK_CUDA(cudaGetLastError());
  }
  flash_bwd_kernel<<<B * H * p.nkv, FA_BWD_THREADS, FA_BWD_SMEM, stream>>>(tq, td, p);
  DET_CHECK_CUDA(cudaGetLastError());
  {
    const long long n = T * (p.hidden / 8);
    const int threads = 256;
    flash_dq_convert_kernel<<<static_cast<unsigned>((n + threads - 1) / threads), threads, 0, stream>>>(
        dq_accum, reinterpret_cast<__nv_bfloat16*>(dqkv), T, p.hidden);
    DET_CHECK_CUDA(cudaGetLastError());
  }
  return 0;
}
