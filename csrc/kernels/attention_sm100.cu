// attention_sm100.cu — fused multi-head self-attention forward and backward for sequences of up to
// 128 tokens with head_dim 64 (BERT-base GLUE shape: seq 128, 12 heads), one CTA per (batch, head).
//
// All five matrix products run on tcgen05 tensor cores with TMEM accumulators; Q/K/V/dO tiles are
// fetched by TMA straight out of the packed [tokens, 3*hidden] QKV activation (no head split /
// transpose kernels), and the probability tile never leaves the SM:
//
//   fwd:  S = Q K^T  ->  P = dropout(softmax(S*scale + mask))  ->  O = P V
//   bwd:  S, dP' = dO V^T  ->  P, dS (registers)  ->  dV = P'^T dO, dQ = dS K, dK = dS^T Q
//
// The transposed products use MN-major shared-memory descriptors over the SAME swizzled smem tile
// that the row-major products read, so P and dS are written once.
//
// Thread mapping: a 128x128 score tile lives in TMEM as 128 lanes (query rows) x 128 columns (keys).
// A warp may only touch the 32 lanes of its quarter (warp_id % 4), so the CTA uses several warps per
// quarter, each owning a slice of the columns; row reductions (max / sum / delta) are combined
// through a small smem exchange.
//
// Replaces HF BertSelfAttention's matmul/softmax/dropout/matmul chain used by the reference's BERT
// example (reference: examples/nlp/bert_glue_pytorch/model_def.py:60-68).
#include "host_utils.h"
#include "sm100_ptx.cuh"

namespace det {

struct AttnParams {
  int B, S, H;          // batch, sequence length (<=128), heads
  int hidden;           // H * 64
  float scale;          // 1/sqrt(64)
  const float* mask;    // additive key mask [B, S] fp32, or null
  float* lse;           // [B, H, S] fp32 (fwd writes, bwd reads)
  __nv_bfloat16* out;   // fwd: ctx [T, hidden]; bwd: dqkv [T, 3*hidden]
  uint32_t drop_threshold;
  float drop_scale;
  uint32_t drop_seed;
  const uint32_t* drop_seed_dev;
};

constexpr int ATT_S = 128;
constexpr int ATT_D = 64;
constexpr uint32_t ATT_TILE = ATT_S * ATT_D * 2;  // 16 KB
constexpr float LOG2E = 1.4426950408889634f;

__device__ __forceinline__ uint4 pack8_bf16(const float* v) {
  uint4 pk;
  pk.x = ptx::pack_bf16x2(v[0], v[1]); pk.y = ptx::pack_bf16x2(v[2], v[3]);
  pk.z = ptx::pack_bf16x2(v[4], v[5]); pk.w = ptx::pack_bf16x2(v[6], v[7]);
  return pk;
}

// ------------------------------------------------------------------------------------------------
// Forward: 256 threads = 4 lane quarters x 2 column halves (64 keys each); 2 CTAs per SM.
// ------------------------------------------------------------------------------------------------
constexpr int ATT_FWD_THREADS = 256;
constexpr uint32_t ATT_FWD_RED = 2 * 2 * ATT_S * 4 + ATT_S * 4;  // max + sum exchange + staged key mask
constexpr uint32_t ATT_FWD_SMEM = 5 * ATT_TILE + 1024 + 64 + ATT_FWD_RED;

__global__ void __launch_bounds__(ATT_FWD_THREADS, 2)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tma_qkv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + ATT_TILE;
  uint8_t* sV = smem + 2 * ATT_TILE;
  uint8_t* sP = smem + 3 * ATT_TILE;  // 2 chunks of [128 q][64 keys]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 5 * ATT_TILE);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  float* red_max = reinterpret_cast<float*>(smem + 5 * ATT_TILE + 64);  // [2][128]
  float* red_sum = red_max + 2 * ATT_S;                                  // [2][128]
  float* smask = red_sum + 2 * ATT_S;                                    // [128] additive key mask * log2(e), -inf past S

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x / p.H, h = blockIdx.x % p.H;
  const uint32_t tma_bar = ptx::smem_u32(&bars[0]);
  const uint32_t mma_bar = ptx::smem_u32(&bars[1]);

  if (tid < ATT_S) smask[tid] = tid < p.S ? (p.mask != nullptr ? __ldg(p.mask + b * p.S + tid) * LOG2E : 0.0f) : -INFINITY;

  if (tid == 0) {
    ptx::tma_prefetch_desc(&tma_qkv);
    ptx::mbar_init(tma_bar, 1);
    ptx::mbar_init(mma_bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 256);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem;        // 128 columns
  const uint32_t tO = tmem + 128;  // 64 columns

  if (tid == 0) {
    ptx::mbar_expect_tx(tma_bar, 3 * ATT_TILE);
    const int row0 = b * p.S;
    ptx::tma_load_2d(ptx::smem_u32(sQ), &tma_qkv, tma_bar, h * ATT_D, row0);
    ptx::tma_load_2d(ptx::smem_u32(sK), &tma_qkv, tma_bar, p.hidden + h * ATT_D, row0);
    ptx::tma_load_2d(ptx::smem_u32(sV), &tma_qkv, tma_bar, 2 * p.hidden + h * ATT_D, row0);
    ptx::mbar_wait(tma_bar, 0);
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
  ptx::mbar_wait(mma_bar, 0);
  ptx::tc_fence_after();

  // ---- softmax: this thread owns query row q, keys [64*cg, 64*cg + 64) ----
  const int q = (warp & 3) * 32 + lane;  // query row == TMEM lane
  const int cg = warp >> 2;              // column half
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  float s[64];
  ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 64, s);
  ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 64 + 32, s + 32);
  ptx::tmem_ld_wait();
  const float sc = p.scale * LOG2E;
  float mx = -INFINITY;
#pragma unroll
  for (int e = 0; e < 64; ++e) {
    const int j = cg * 64 + e;
    s[e] = fmaf(s[e], sc, smask[j]);  // -inf for keys past the sequence end
    mx = fmaxf(mx, s[e]);
  }
  red_max[cg * ATT_S + q] = mx;
  __syncthreads();
  mx = fmaxf(red_max[q], red_max[ATT_S + q]);
  float sum = 0.0f;
#pragma unroll
  for (int e = 0; e < 64; ++e) {
    s[e] = fast_exp2(s[e] - mx);
    sum += s[e];
  }
  red_sum[cg * ATT_S + q] = sum;
  __syncthreads();
  sum = red_sum[q] + red_sum[ATT_S + q];
  const float inv = 1.0f / sum;
  if (cg == 0 && q < p.S) p.lse[(static_cast<long long>(b) * p.H + h) * p.S + q] = (mx + log2f(sum)) * (1.0f / LOG2E);
  const uint32_t didx0 = ((static_cast<uint32_t>(b) * p.H + h) * p.S + q) * p.S + cg * 64;
  const uint32_t dseed = dropout_effective_seed(p.drop_seed, p.drop_seed_dev);
#pragma unroll
  for (int u = 0; u < 8; ++u) {  // 16-byte units of 8 keys; this half is exactly smem chunk `cg`
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float pv = s[u * 8 + e] * inv;
      if (p.drop_threshold != 0u)
        pv = dropout_keep(dseed, didx0 + u * 8 + e, p.drop_threshold) ? pv * p.drop_scale : 0.0f;
      v[e] = pv;
    }
    *reinterpret_cast<uint4*>(sP + cg * ATT_TILE + ptx::sw128_offset(q, u)) = pack8_bf16(v);
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();

  if (tid == 0) {
    ptx::tc_fence_after();
    constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(128, 64, 0, 1);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sP) + (k >> 2) * ATT_TILE + (k & 3) * 32, 16, 1024);
      const uint64_t bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sV) + k * 2048, 8192, 1024);
      ptx::umma_bf16_ss(tO, ad, bd, idesc, k > 0);
    }
    ptx::umma_commit(mma_bar);
  }
  ptx::mbar_wait(mma_bar, 1);
  ptx::tc_fence_after();
  float o[32];
  ptx::tmem_ld_32x32b_x32(tO + lane_base + cg * 32, o);
  ptx::tmem_ld_wait();
  if (q < p.S) {
    uint4* dst = reinterpret_cast<uint4*>(p.out + (static_cast<long long>(b) * p.S + q) * p.hidden + h * ATT_D + cg * 32);
#pragma unroll
    for (int u = 0; u < 4; ++u) dst[u] = pack8_bf16(o + 8 * u);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 256);
  }
}

// ------------------------------------------------------------------------------------------------
// Forward, persistent form: two CTAs per SM walk (batch, head) items. The Q/K buffers are free as soon as S sits in
// TMEM and the V buffer as soon as O does, so the next item's tiles are fetched under this item's softmax / PV / store
// and TMEM allocation and barrier setup are paid once per CTA instead of once per item.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ATT_FWD_THREADS, 2)
attention_fwd_persistent_kernel(const __grid_constant__ CUtensorMap tma_qkv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + ATT_TILE;
  uint8_t* sV = smem + 2 * ATT_TILE;
  uint8_t* sP = smem + 3 * ATT_TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 5 * ATT_TILE);   // qk, v, mma
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);
  float* red_max = reinterpret_cast<float*>(smem + 5 * ATT_TILE + 64);
  float* red_sum = red_max + 2 * ATT_S;
  float* smask = red_sum + 2 * ATT_S;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int total = p.B * p.H;
  const int n_items = (total - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const uint32_t qk_bar = ptx::smem_u32(&bars[0]);
  const uint32_t v_bar = ptx::smem_u32(&bars[1]);
  const uint32_t mma_bar = ptx::smem_u32(&bars[2]);

  if (tid == 0) {
    ptx::tma_prefetch_desc(&tma_qkv);
    ptx::mbar_init(qk_bar, 1);
    ptx::mbar_init(v_bar, 1);
    ptx::mbar_init(mma_bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 256);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tO = tmem + 128;

  auto item_of = [&](int i) { return static_cast<int>(blockIdx.x) + i * static_cast<int>(gridDim.x); };
  auto load_qk = [&](int i) {   // tid 0 only
    const int w = item_of(i), b = w / p.H, h = w % p.H, row0 = b * p.S;
    ptx::mbar_expect_tx(qk_bar, 2 * ATT_TILE);
    ptx::tma_load_2d(ptx::smem_u32(sQ), &tma_qkv, qk_bar, h * ATT_D, row0);
    ptx::tma_load_2d(ptx::smem_u32(sK), &tma_qkv, qk_bar, p.hidden + h * ATT_D, row0);
  };
  auto load_v = [&](int i) {    // tid 0 only
    const int w = item_of(i), b = w / p.H, h = w % p.H, row0 = b * p.S;
    ptx::mbar_expect_tx(v_bar, ATT_TILE);
    ptx::tma_load_2d(ptx::smem_u32(sV), &tma_qkv, v_bar, 2 * p.hidden + h * ATT_D, row0);
  };
  if (tid == 0 && n_items > 0) { load_qk(0); load_v(0); }

  const int q = (warp & 3) * 32 + lane;
  const int cg = warp >> 2;
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const float sc = p.scale * LOG2E;
  const uint32_t dseed = dropout_effective_seed(p.drop_seed, p.drop_seed_dev);

  for (int i = 0; i < n_items; ++i) {
    const int w = item_of(i), b = w / p.H, h = w % p.H;
    if (tid < ATT_S) smask[tid] = tid < p.S ? (p.mask != nullptr ? __ldg(p.mask + b * p.S + tid) * LOG2E : 0.0f) : -INFINITY;
    if (tid == 0) {
      ptx::mbar_wait(qk_bar, i & 1);
      ptx::tc_fence_after();
      constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(128, 128, 0, 0);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        ptx::umma_bf16_ss(tS, ptx::make_smem_desc_sw128(ptx::smem_u32(sQ) + k * 32, 16, 1024),
                          ptx::make_smem_desc_sw128(ptx::smem_u32(sK) + k * 32, 16, 1024), idesc, k > 0);
      ptx::umma_commit(mma_bar);
    }
    __syncthreads();   // key mask staged
    ptx::mbar_wait(mma_bar, 0);
    ptx::tc_fence_after();
    if (tid == 0 && i + 1 < n_items) load_qk(i + 1);   // S is in TMEM: Q / K buffers are free

    float s[64];
    ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 64, s);
    ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 64 + 32, s + 32);
    ptx::tmem_ld_wait();
    float mx = -INFINITY;
#pragma unroll
    for (int e = 0; e < 64; ++e) {
      s[e] = fmaf(s[e], sc, smask[cg * 64 + e]);
      mx = fmaxf(mx, s[e]);
    }
    red_max[cg * ATT_S + q] = mx;
    __syncthreads();
    mx = fmaxf(red_max[q], red_max[ATT_S + q]);
    float sum = 0.0f;
#pragma unroll
    for (int e = 0; e < 64; ++e) {
      s[e] = fast_exp2(s[e] - mx);
      sum += s[e];
    }
    red_sum[cg * ATT_S + q] = sum;
    __syncthreads();
    sum = red_sum[q] + red_sum[ATT_S + q];
    const float inv = 1.0f / sum;
    if (cg == 0 && q < p.S) p.lse[(static_cast<long long>(b) * p.H + h) * p.S + q] = (mx + log2f(sum)) * (1.0f / LOG2E);
    const uint32_t didx0 = ((static_cast<uint32_t>(b) * p.H + h) * p.S + q) * p.S + cg * 64;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      float v[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        float pv = s[u * 8 + e] * inv;
        if (p.drop_threshold != 0u)
          pv = dropout_keep(dseed, didx0 + u * 8 + e, p.drop_threshold) ? pv * p.drop_scale : 0.0f;
        v[e] = pv;
      }
      *reinterpret_cast<uint4*>(sP + cg * ATT_TILE + ptx::sw128_offset(q, u)) = pack8_bf16(v);
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();

    if (tid == 0) {
      ptx::mbar_wait(v_bar, i & 1);
      ptx::tc_fence_after();
      constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(128, 64, 0, 1);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        ptx::umma_bf16_ss(tO, ptx::make_smem_desc_sw128(ptx::smem_u32(sP) + (k >> 2) * ATT_TILE + (k & 3) * 32, 16, 1024),
                          ptx::make_smem_desc_sw128(ptx::smem_u32(sV) + k * 2048, 8192, 1024), idesc, k > 0);
      ptx::umma_commit(mma_bar);
    }
    ptx::mbar_wait(mma_bar, 1);
    ptx::tc_fence_after();
    if (tid == 0 && i + 1 < n_items) load_v(i + 1);    // O is in TMEM: the V buffer is free
    float o[32];
    ptx::tmem_ld_32x32b_x32(tO + lane_base + cg * 32, o);
    ptx::tmem_ld_wait();
    if (q < p.S) {
      uint4* dst = reinterpret_cast<uint4*>(p.out + (static_cast<long long>(b) * p.S + q) * p.hidden + h * ATT_D + cg * 32);
#pragma unroll
      for (int u = 0; u < 4; ++u) dst[u] = pack8_bf16(o + 8 * u);
    }
    ptx::tc_fence_before();
    __syncthreads();   // S / O / P / exchange buffers are free for the next item
  }
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 256);
  }
}

// ------------------------------------------------------------------------------------------------
// Forward, dense form: FOUR CTAs per SM, one (batch, head) item each. An item is a chain of short dependent steps
// (TMA -> QK^T -> softmax -> PV -> store) that cannot fill an SM by itself; what fills it is several items side by
// side. To fit four: 48 KB of tiles (P overwrites the Q and K tiles, which are dead once S sits in TMEM), 128 TMEM
// columns (O reuses S's columns, dead once P is in shared memory) and <= 64 registers — the softmax walks its 64 keys
// in two 32-column reads of S, twice (max, then exp / sum / pack), instead of holding 64 scores, and P is stored
// un-normalised: O = P_un V, and each output row is scaled by 1/sum (and the dropout scale) on its way out.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t ATT_FWD_DENSE_SMEM = 3 * ATT_TILE + 1024 + 64 + ATT_FWD_RED;

__global__ void __launch_bounds__(ATT_FWD_THREADS, 4)
attention_fwd_dense_kernel(const __grid_constant__ CUtensorMap tma_qkv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + ATT_TILE;
  uint8_t* sV = smem + 2 * ATT_TILE;
  uint8_t* sP = smem;                 // 2 chunks of [128 q][64 keys] over Q | K
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 3 * ATT_TILE);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  float* red_max = reinterpret_cast<float*>(smem + 3 * ATT_TILE + 64);
  float* red_sum = red_max + 2 * ATT_S;
  float* smask = red_sum + 2 * ATT_S;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x / p.H, h = blockIdx.x % p.H;
  const uint32_t tma_bar = ptx::smem_u32(&bars[0]);
  const uint32_t mma_bar = ptx::smem_u32(&bars[1]);

  if (tid < ATT_S) smask[tid] = tid < p.S ? (p.mask != nullptr ? __ldg(p.mask + b * p.S + tid) * LOG2E : 0.0f) : -INFINITY;
  if (tid == 0) {
    ptx::tma_prefetch_desc(&tma_qkv);
    ptx::mbar_init(tma_bar, 1);
    ptx::mbar_init(mma_bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 128);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tS = *tmem_slot;     // 128 columns; O takes the first 64 once S has been consumed

  if (tid == 0) {
    ptx::mbar_expect_tx(tma_bar, 3 * ATT_TILE);
    const int row0 = b * p.S;
    ptx::tma_load_2d(ptx::smem_u32(sQ), &tma_qkv, tma_bar, h * ATT_D, row0);
    ptx::tma_load_2d(ptx::smem_u32(sK), &tma_qkv, tma_bar, p.hidden + h * ATT_D, row0);
    ptx::tma_load_2d(ptx::smem_u32(sV), &tma_qkv, tma_bar, 2 * p.hidden + h * ATT_D, row0);
    ptx::mbar_wait(tma_bar, 0);
    ptx::tc_fence_after();
    constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(128, 128, 0, 0);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      ptx::umma_bf16_ss(tS, ptx::make_smem_desc_sw128(ptx::smem_u32(sQ) + k * 32, 16, 1024),
                        ptx::make_smem_desc_sw128(ptx::smem_u32(sK) + k * 32, 16, 1024), idesc, k > 0);
    ptx::umma_commit(mma_bar);
  }
  ptx::mbar_wait(mma_bar, 0);
  ptx::tc_fence_after();

  const int q = (warp & 3) * 32 + lane;
  const int cg = warp >> 2;
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const float sc = p.scale * LOG2E;
  // ---- pass 1: row maximum over this thread's 64 keys, 32 at a time
  float mx = -INFINITY;
#pragma unroll 1
  for (int c = 0; c < 2; ++c) {
    float s[32];
    ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 64 + c * 32, s);
    ptx::tmem_ld_wait();
#pragma unroll
    for (int e = 0; e < 32; ++e) mx = fmaxf(mx, fmaf(s[e], sc, smask[cg * 64 + c * 32 + e]));
  }
  red_max[cg * ATT_S + q] = mx;
  __syncthreads();
  mx = fmaxf(red_max[q], red_max[ATT_S + q]);
  // ---- pass 2: exp, row sum, dropout, un-normalised P -> shared memory (over Q | K: S is complete, they are dead)
  const uint32_t dseed = dropout_effective_seed(p.drop_seed, p.drop_seed_dev);
  const uint32_t didx0 = ((static_cast<uint32_t>(b) * p.H + h) * p.S + q) * p.S + cg * 64;
  float sum = 0.0f;
#pragma unroll 1
  for (int c = 0; c < 2; ++c) {
    float s[32];
    ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 64 + c * 32, s);
    ptx::tmem_ld_wait();
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float v[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int k = c * 32 + u * 8 + e;
        float pv = fast_exp2(fmaf(s[u * 8 + e], sc, smask[cg * 64 + k]) - mx);
        sum += pv;                      // the normaliser counts dropped entries too (dropout acts on probabilities)
        if (p.drop_threshold != 0u && !dropout_keep(dseed, didx0 + k, p.drop_threshold)) pv = 0.0f;
        v[e] = pv;
      }
      *reinterpret_cast<uint4*>(sP + cg * ATT_TILE + ptx::sw128_offset(q, c * 4 + u)) = pack8_bf16(v);
    }
  }
  red_sum[cg * ATT_S + q] = sum;
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  sum = red_sum[q] + red_sum[ATT_S + q];
  if (cg == 0 && q < p.S) p.lse[(static_cast<long long>(b) * p.H + h) * p.S + q] = (mx + log2f(sum)) * (1.0f / LOG2E);

  if (tid == 0) {
    ptx::tc_fence_after();
    constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(128, 64, 0, 1);
#pragma unroll
    for (int k = 0; k < 8; ++k)
      ptx::umma_bf16_ss(tS, ptx::make_smem_desc_sw128(ptx::smem_u32(sP) + (k >> 2) * ATT_TILE + (k & 3) * 32, 16, 1024),
                        ptx::make_smem_desc_sw128(ptx::smem_u32(sV) + k * 2048, 8192, 1024), idesc, k > 0);
    ptx::umma_commit(mma_bar);
  }
  ptx::mbar_wait(mma_bar, 1);
  ptx::tc_fence_after();
  float o[32];
  ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 32, o);
  ptx::tmem_ld_wait();
  const float row_scale = (p.drop_threshold != 0u ? p.drop_scale : 1.0f) / sum;
  if (q < p.S) {
#pragma unroll
    for (int e = 0; e < 32; ++e) o[e] *= row_scale;
    uint4* dst = reinterpret_cast<uint4*>(p.out + (static_cast<long long>(b) * p.S + q) * p.hidden + h * ATT_D + cg * 32);
#pragma unroll
    for (int u = 0; u < 4; ++u) dst[u] = pack8_bf16(o + 8 * u);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tS, 128);
  }
}

// ------------------------------------------------------------------------------------------------
// Backward: 512 threads = 4 lane quarters x 4 column groups (32 keys each); 1 CTA per SM.
// ------------------------------------------------------------------------------------------------
constexpr int ATT_BWD_THREADS = 512;
constexpr uint32_t ATT_BWD_RED = 4 * ATT_S * 4 + ATT_S * 4;
constexpr uint32_t ATT_BWD_SMEM = 8 * ATT_TILE + 1024 + 64 + ATT_BWD_RED;

__global__ void __launch_bounds__(ATT_BWD_THREADS, 1)
attention_bwd_kernel(const __grid_constant__ CUtensorMap tma_qkv,
                     const __grid_constant__ CUtensorMap tma_do, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + ATT_TILE;
  uint8_t* sV = smem + 2 * ATT_TILE;
  uint8_t* sdO = smem + 3 * ATT_TILE;
  uint8_t* sP = smem + 4 * ATT_TILE;   // 2 chunks
  uint8_t* sdS = smem + 6 * ATT_TILE;  // 2 chunks
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 8 * ATT_TILE);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  float* red_delta = reinterpret_cast<float*>(smem + 8 * ATT_TILE + 64);  // [4][128]
  float* smask = red_delta + 4 * ATT_S;                                   // [128]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x / p.H, h = blockIdx.x % p.H;
  const uint32_t tma_bar = ptx::smem_u32(&bars[0]);
  const uint32_t mma_bar = ptx::smem_u32(&bars[1]);

  if (tid < ATT_S) smask[tid] = tid < p.S ? (p.mask != nullptr ? __ldg(p.mask + b * p.S + tid) * LOG2E : 0.0f) : -INFINITY;

  if (tid == 0) {
    ptx::tma_prefetch_desc(&tma_qkv);
    ptx::tma_prefetch_desc(&tma_do);
    ptx::mbar_init(tma_bar, 1);
    ptx::mbar_init(mma_bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tdP = tmem + 128, tdV = tmem + 256, tdQ = tmem + 320, tdK = tmem + 384;

  if (tid == 0) {
    ptx::mbar_expect_tx(tma_bar, 4 * ATT_TILE);
    const int row0 = b * p.S;
    ptx::tma_load_2d(ptx::smem_u32(sQ), &tma_qkv, tma_bar, h * ATT_D, row0);
    ptx::tma_load_2d(ptx::smem_u32(sK), &tma_qkv, tma_bar, p.hidden + h * ATT_D, row0);
    ptx::tma_load_2d(ptx::smem_u32(sV), &tma_qkv, tma_bar, 2 * p.hidden + h * ATT_D, row0);
    ptx::tma_load_2d(ptx::smem_u32(sdO), &tma_do, tma_bar, h * ATT_D, row0);
    ptx::mbar_wait(tma_bar, 0);
    ptx::tc_fence_after();
    constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(128, 128, 0, 0);
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // S = Q K^T
      const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sQ) + k * 32, 16, 1024);
      const uint64_t bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sK) + k * 32, 16, 1024);
      ptx::umma_bf16_ss(tS, ad, bd, idesc, k > 0);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // dP' = dO V^T
      const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sdO) + k * 32, 16, 1024);
      const uint64_t bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sV) + k * 32, 16, 1024);
      ptx::umma_bf16_ss(tdP, ad, bd, idesc, k > 0);
    }
    ptx::umma_commit(mma_bar);
  }
  ptx::mbar_wait(mma_bar, 0);
  ptx::tc_fence_after();

  const int q = (warp & 3) * 32 + lane;
  const int cg = warp >> 2;  // keys [32*cg, 32*cg + 32)
  const bool q_ok = q < p.S;
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const float sc = p.scale * LOG2E;
  const float lse2 = q_ok ? p.lse[(static_cast<long long>(b) * p.H + h) * p.S + q] * LOG2E : 0.0f;
  const uint32_t didx0 = ((static_cast<uint32_t>(b) * p.H + h) * p.S + q) * p.S + cg * 32;
  const uint32_t dseed = dropout_effective_seed(p.drop_seed, p.drop_seed_dev);

  float pr[32], dp[32];
  ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 32, pr);
  ptx::tmem_ld_32x32b_x32(tdP + lane_base + cg * 32, dp);
  ptx::tmem_ld_wait();
  // P (pre-dropout probabilities), masked dP, P' = dropout(P) -> smem, partial delta
  float delta = 0.0f;
  {
    float pd[32];
#pragma unroll
    for (int e = 0; e < 32; ++e) {
      const int j = cg * 32 + e;
      const float pv = q_ok ? fast_exp2(fmaf(pr[e], sc, smask[j]) - lse2) : 0.0f;  // exp2(-inf) = 0 past the sequence end
      float keep = 1.0f;
      if (p.drop_threshold != 0u) keep = dropout_keep(dseed, didx0 + e, p.drop_threshold) ? p.drop_scale : 0.0f;
      pr[e] = pv;
      pd[e] = pv * keep;
      dp[e] *= keep;               // dP = dP' * mask * scale
      delta += pv * dp[e];         // sum_j P'_ij dP'_ij  ==  sum_j P_ij dP_ij
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t gu = cg * 4 + u;  // global 16-byte unit index 0..15
      *reinterpret_cast<uint4*>(sP + (gu >> 3) * ATT_TILE + ptx::sw128_offset(q, gu & 7)) = pack8_bf16(pd + 8 * u);
    }
  }
  red_delta[cg * ATT_S + q] = delta;
  __syncthreads();
  delta = red_delta[q] + red_delta[ATT_S + q] + red_delta[2 * ATT_S + q] + red_delta[3 * ATT_S + q];
  // dS = P * (dP - delta) * scale -> smem
#pragma unroll
  for (int e = 0; e < 32; ++e) pr[e] = pr[e] * (dp[e] - delta) * p.scale;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const uint32_t gu = cg * 4 + u;
    *reinterpret_cast<uint4*>(sdS + (gu >> 3) * ATT_TILE + ptx::sw128_offset(q, gu & 7)) = pack8_bf16(pr + 8 * u);
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();

  if (tid == 0) {
    ptx::tc_fence_after();
    constexpr uint32_t idesc_tn = ptx::make_idesc_bf16_f32(128, 64, 1, 1);  // A^T (MN-major), B MN-major
    constexpr uint32_t idesc_nn = ptx::make_idesc_bf16_f32(128, 64, 0, 1);  // A K-major,     B MN-major
#pragma unroll
    for (int k = 0; k < 8; ++k) {  // dV[key, d] = sum_q P'[q, key] dO[q, d]
      const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sP) + k * 2048, ATT_TILE, 1024);
      const uint64_t bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sdO) + k * 2048, 8192, 1024);
      ptx::umma_bf16_ss(tdV, ad, bd, idesc_tn, k > 0);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {  // dQ[q, d] = sum_key dS[q, key] K[key, d]
      const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sdS) + (k >> 2) * ATT_TILE + (k & 3) * 32, 16, 1024);
      const uint64_t bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sK) + k * 2048, 8192, 1024);
      ptx::umma_bf16_ss(tdQ, ad, bd, idesc_nn, k > 0);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {  // dK[key, d] = sum_q dS[q, key] Q[q, d]
      const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sdS) + k * 2048, ATT_TILE, 1024);
      const uint64_t bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sQ) + k * 2048, 8192, 1024);
      ptx::umma_bf16_ss(tdK, ad, bd, idesc_tn, k > 0);
    }
    ptx::umma_commit(mma_bar);
  }
  ptx::mbar_wait(mma_bar, 1);
  ptx::tc_fence_after();

  // Readout: 3 matrices x 64 columns = 192 columns; column group cg takes 48 of them (3 x 16).
  const long long out_row = (static_cast<long long>(b) * p.S + q) * (3LL * p.hidden) + h * ATT_D;
#pragma unroll
  for (int t = 0; t < 3; ++t) {
    const int c16 = cg * 3 + t;          // 16-column slab index 0..11 over [dQ | dK | dV]
    const int which = c16 >> 2;          // 0 dQ, 1 dK, 2 dV
    const int col = (c16 & 3) * 16;
    const uint32_t src = (which == 0) ? tdQ : (which == 1 ? tdK : tdV);
    float o[16];
    ptx::tmem_ld_32x32b_x16(src + lane_base + col, o);
    ptx::tmem_ld_wait();
    if (q_ok) {
      uint4* dst = reinterpret_cast<uint4*>(p.out + out_row + static_cast<long long>(which) * p.hidden + col);
      dst[0] = pack8_bf16(o);
      dst[1] = pack8_bf16(o + 8);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 512);
  }
}

// ------------------------------------------------------------------------------------------------
// Backward, persistent form: one CTA per SM walks (batch, head) items, so TMEM allocation, barrier setup and the
// TMA round trip of an item are paid under the previous item's work instead of once per 12 us CTA (ncu on the
// one-shot kernel: half of all warp samples wait on TMA / MMA barriers, tensor pipe 5 % busy). Per item i:
//   A  wait S_i, dP'_i                 -> registers;  barrier: TMEM S/dP free
//   C  P'_i, delta, dS_i in registers;  wait MMA2 of item i-1 -> its Q/K/V/dO stage is refilled with item i+1 and its
//      dQ/dK/dV are stored;  P'_i, dS_i -> smem;  barrier
//   D  issue MMA2(i), then MMA1(i+1) as soon as its tiles have landed
// Two stages of {Q, K, V, dO} + P + dS = 12 tiles (192 KB).
// ------------------------------------------------------------------------------------------------
constexpr uint32_t ATT_BWDP_SMEM = 12 * ATT_TILE + 1024 + 64 + 4 * ATT_S * 4 + 2 * ATT_S * 4;

__global__ void __launch_bounds__(ATT_BWD_THREADS, 1)
attention_bwd_persistent_kernel(const __grid_constant__ CUtensorMap tma_qkv, const __grid_constant__ CUtensorMap tma_do,
                                const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sIn = smem;                       // stage s: Q, K, V, dO at sIn + (4 * s + {0,1,2,3}) * TILE
  uint8_t* sP = smem + 8 * ATT_TILE;         // 2 chunks
  uint8_t* sdS = smem + 10 * ATT_TILE;       // 2 chunks
  uint8_t* tail = smem + 12 * ATT_TILE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);      // ld[0], ld[1], mma1, mma2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
  float* red_delta = reinterpret_cast<float*>(tail + 64);  // [4][128]
  float* smask = red_delta + 4 * ATT_S;                    // [2][128]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int total = p.B * p.H;
  const int n_items = (total - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const uint32_t ld_bar0 = ptx::smem_u32(&bars[0]);
  const uint32_t bar1 = ptx::smem_u32(&bars[2]);
  const uint32_t bar2 = ptx::smem_u32(&bars[3]);

  if (tid == 0) {
    ptx::tma_prefetch_desc(&tma_qkv);
    ptx::tma_prefetch_desc(&tma_do);
    for (int i = 0; i < 4; ++i) ptx::mbar_init(ptx::smem_u32(&bars[i]), 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tdP = tmem + 128, tdV = tmem + 256, tdQ = tmem + 320, tdK = tmem + 384;

  auto item_of = [&](int i) { return static_cast<int>(blockIdx.x) + i * static_cast<int>(gridDim.x); };
  auto load_item = [&](int i) {   // tid 0 only
    const int w = item_of(i), st = i & 1;
    const int b = w / p.H, h = w % p.H, row0 = b * p.S;
    const uint32_t lb = ld_bar0 + 8 * st;
    uint8_t* base = sIn + 4 * st * ATT_TILE;
    ptx::mbar_expect_tx(lb, 4 * ATT_TILE);
    ptx::tma_load_2d(ptx::smem_u32(base), &tma_qkv, lb, h * ATT_D, row0);
    ptx::tma_load_2d(ptx::smem_u32(base + ATT_TILE), &tma_qkv, lb, p.hidden + h * ATT_D, row0);
    ptx::tma_load_2d(ptx::smem_u32(base + 2 * ATT_TILE), &tma_qkv, lb, 2 * p.hidden + h * ATT_D, row0);
    ptx::tma_load_2d(ptx::smem_u32(base + 3 * ATT_TILE), &tma_do, lb, h * ATT_D, row0);
  };
  auto issue_mma1 = [&](int i) {  // tid 0 only
    const int st = i & 1;
    ptx::mbar_wait(ld_bar0 + 8 * st, (i >> 1) & 1);
    ptx::tc_fence_after();
    const uint32_t q_s = ptx::smem_u32(sIn + 4 * st * ATT_TILE), k_s = q_s + ATT_TILE, v_s = q_s + 2 * ATT_TILE, do_s = q_s + 3 * ATT_TILE;
    constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(128, 128, 0, 0);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      ptx::umma_bf16_ss(tS, ptx::make_smem_desc_sw128(q_s + k * 32, 16, 1024), ptx::make_smem_desc_sw128(k_s + k * 32, 16, 1024), idesc, k > 0);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      ptx::umma_bf16_ss(tdP, ptx::make_smem_desc_sw128(do_s + k * 32, 16, 1024), ptx::make_smem_desc_sw128(v_s + k * 32, 16, 1024), idesc, k > 0);
    ptx::umma_commit(bar1);
  };

  const int q = (warp & 3) * 32 + lane;
  const int cg = warp >> 2;
  const bool q_ok = q < p.S;
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const float sc = p.scale * LOG2E;
  const uint32_t dseed = dropout_effective_seed(p.drop_seed, p.drop_seed_dev);

  auto stage_mask = [&](int i) {  // threads 0..127
    const int b = item_of(i) / p.H;
    smask[(i & 1) * ATT_S + tid] = tid < p.S ? (p.mask != nullptr ? __ldg(p.mask + b * p.S + tid) * LOG2E : 0.0f) : -INFINITY;
  };
  auto store_item = [&](int i) {  // read dQ | dK | dV of item i out of TMEM (192 columns, 48 per column group)
    const int w = item_of(i);
    const int b = w / p.H, h = w % p.H;
    const long long out_row = (static_cast<long long>(b) * p.S + q) * (3LL * p.hidden) + h * ATT_D;
#pragma unroll
    for (int t = 0; t < 3; ++t) {
      const int c16 = cg * 3 + t;
      const int which = c16 >> 2;
      const int col = (c16 & 3) * 16;
      const uint32_t src = (which == 0) ? tdQ : (which == 1 ? tdK : tdV);
      float o[16];
      ptx::tmem_ld_32x32b_x16(src + lane_base + col, o);
      ptx::tmem_ld_wait();
      if (q_ok) {
        uint4* dst = reinterpret_cast<uint4*>(p.out + out_row + static_cast<long long>(which) * p.hidden + col);
        dst[0] = pack8_bf16(o);
        dst[1] = pack8_bf16(o + 8);
      }
    }
  };

  if (n_items > 0) {
    if (tid < ATT_S) stage_mask(0);
    if (tid == 0) {
      load_item(0);
      if (n_items > 1) load_item(1);
      issue_mma1(0);
    }
  }

  for (int i = 0; i < n_items; ++i) {
    const int w = item_of(i);
    const int b = w / p.H, h = w % p.H;
    if (tid < ATT_S && i + 1 < n_items) stage_mask(i + 1);
    const float lse2 = q_ok ? __ldg(p.lse + (static_cast<long long>(b) * p.H + h) * p.S + q) * LOG2E : 0.0f;
    const uint32_t didx0 = ((static_cast<uint32_t>(b) * p.H + h) * p.S + q) * p.S + cg * 32;
    const float* mk = smask + (i & 1) * ATT_S;

    // ---- A ----
    float pr[32], dp[32];
    ptx::mbar_wait(bar1, i & 1);
    ptx::tc_fence_after();
    ptx::tmem_ld_32x32b_x32(tS + lane_base + cg * 32, pr);
    ptx::tmem_ld_32x32b_x32(tdP + lane_base + cg * 32, dp);
    ptx::tmem_ld_wait();
    ptx::tc_fence_before();
    __syncthreads();   // also publishes this item's staged mask (written one iteration ago, or in the prologue)

    // ---- C ----
    uint4 pk_p[4];
    float delta = 0.0f;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float pd[8];
#pragma unroll
      for (int v = 0; v < 8; ++v) {
        const int e = u * 8 + v;
        const float pv = q_ok ? fast_exp2(fmaf(pr[e], sc, mk[cg * 32 + e]) - lse2) : 0.0f;
        float keep = 1.0f;
        if (p.drop_threshold != 0u) keep = dropout_keep(dseed, didx0 + e, p.drop_threshold) ? p.drop_scale : 0.0f;
        dp[e] *= keep;
        delta = fmaf(pv, dp[e], delta);
        pr[e] = pv;
        pd[v] = pv * keep;
      }
      pk_p[u] = pack8_bf16(pd);
    }
    red_delta[cg * ATT_S + q] = delta;
    if (i > 0) {
      ptx::mbar_wait(bar2, (i - 1) & 1);   // item i-1 is complete: its input stage, sP / sdS and the dQ/dK/dV columns are free
      ptx::tc_fence_after();
      if (tid == 0 && i + 1 < n_items) load_item(i + 1);
      store_item(i - 1);
    }
    __syncthreads();   // delta exchange
    delta = red_delta[q] + red_delta[ATT_S + q] + red_delta[2 * ATT_S + q] + red_delta[3 * ATT_S + q];
#pragma unroll
    for (int e = 0; e < 32; ++e) pr[e] = pr[e] * (dp[e] - delta) * p.scale;   // dS
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t gu = cg * 4 + u;
      *reinterpret_cast<uint4*>(sP + (gu >> 3) * ATT_TILE + ptx::sw128_offset(q, gu & 7)) = pk_p[u];
      *reinterpret_cast<uint4*>(sdS + (gu >> 3) * ATT_TILE + ptx::sw128_offset(q, gu & 7)) = pack8_bf16(pr + 8 * u);
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();

    // ---- D ----
    if (tid == 0) {
      ptx::tc_fence_after();
      const int st = i & 1;
      const uint32_t q_s = ptx::smem_u32(sIn + 4 * st * ATT_TILE), k_s = q_s + ATT_TILE, do_s = q_s + 3 * ATT_TILE;
      constexpr uint32_t idesc_tn = ptx::make_idesc_bf16_f32(128, 64, 1, 1);
      constexpr uint32_t idesc_nn = ptx::make_idesc_bf16_f32(128, 64, 0, 1);
#pragma unroll
      for (int k = 0; k < 8; ++k)   // dV = P'^T dO
        ptx::umma_bf16_ss(tdV, ptx::make_smem_desc_sw128(ptx::smem_u32(sP) + k * 2048, ATT_TILE, 1024),
                          ptx::make_smem_desc_sw128(do_s + k * 2048, 8192, 1024), idesc_tn, k > 0);
#pragma unroll
      for (int k = 0; k < 8; ++k)   // dQ = dS K
        ptx::umma_bf16_ss(tdQ, ptx::make_smem_desc_sw128(ptx::smem_u32(sdS) + (k >> 2) * ATT_TILE + (k & 3) * 32, 16, 1024),
                          ptx::make_smem_desc_sw128(k_s + k * 2048, 8192, 1024), idesc_nn, k > 0);
#pragma unroll
      for (int k = 0; k < 8; ++k)   // dK = dS^T Q
        ptx::umma_bf16_ss(tdK, ptx::make_smem_desc_sw128(ptx::smem_u32(sdS) + k * 2048, ATT_TILE, 1024),
                          ptx::make_smem_desc_sw128(q_s + k * 2048, 8192, 1024), idesc_tn, k > 0);
      ptx::umma_commit(bar2);
      if (i + 1 < n_items) issue_mma1(i + 1);
    }
  }
  if (n_items > 0) {
    ptx::mbar_wait(bar2, (n_items - 1) & 1);
    ptx::tc_fence_after();
    store_item(n_items - 1);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 512);
  }
}

static int fill_params(AttnParams& p, int B, int S, int H, const float* mask, float* lse, void* out,
                       float dropout_p, unsigned seed, const unsigned* seed_dev) {
  DET_REQUIRE(S >= 1 && S <= ATT_S, "fused attention supports sequence length <= 128");
  DET_REQUIRE(B >= 1 && H >= 1, "bad attention shape");
  p.B = B; p.S = S; p.H = H; p.hidden = H * ATT_D;
  p.scale = 0.125f;
  p.mask = mask; p.lse = lse; p.out = reinterpret_cast<__nv_bfloat16*>(out);
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

// qkv: [B*S, 3*H*64] bf16 packed (Q | K | V, each head-major);  ctx: [B*S, H*64] bf16;
// mask: additive fp32 [B, S] or null;  lse: [B, H, S] fp32.
extern "C" int det_attention_fwd(const void* qkv, const float* mask, void* ctx, float* lse, int B,
                                 int S, int H, float dropout_p, unsigned int seed,
                                 const unsigned int* seed_dev, cudaStream_t stream) {
  using namespace det;
  AttnParams p;
  int rc = fill_params(p, B, S, H, mask, lse, ctx, dropout_p, seed, seed_dev);
  if (rc) return rc;
  CUtensorMap tq;
  rc = make_tma_2d_bf16(&tq, qkv, 3ull * p.hidden, (uint64_t)B * S, 3ull * p.hidden, 64, ATT_S);
  if (rc) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  DET_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DET_CHECK_CUDA(cudaFuncSetAttribute(attention_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_FWD_SMEM));
    DET_CHECK_CUDA(cudaFuncSetAttribute(attention_fwd_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_FWD_SMEM));
    attr_set[dev] = true;
  }
  // DET_ATTN_FWD: "persistent" (default: two CTAs per SM walking items with the next item's tiles prefetched), "dense"
  // (four one-item CTAs per SM), "oneshot". Dense and persistent measure the same (19.3-20.0 us at batch 48 x 12 heads,
  // BERT step 18.6 ms either way; profiles/ncu_attention128_r2.csv): with ~4 items per SM the kernel is one item's chain
  // of load -> QK^T -> softmax -> PV -> store, and the softmax's ~1,300 instructions per warp are issue-bound
  // (38 % issue-active) behind a load phase every CTA enters at the same time.
  static const int mode = [] {
    const char* e = getenv("DET_ATTN_FWD");
    if (e != nullptr && e[0] == 'd') return 0;
    if (e != nullptr && e[0] == 'o') return 2;
    return 1;
  }();
  if (mode == 0) {
    static bool dense_attr[64] = {false};
    if (dev < 64 && !dense_attr[dev]) {
      DET_CHECK_CUDA(cudaFuncSetAttribute(attention_fwd_dense_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_FWD_DENSE_SMEM));
      dense_attr[dev] = true;
    }
    attention_fwd_dense_kernel<<<B * H, ATT_FWD_THREADS, ATT_FWD_DENSE_SMEM, stream>>>(tq, p);
    DET_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  static const int one_shot = [] { const char* e = getenv("DET_ATTN_ONESHOT"); return e != nullptr && e[0] == '1' ? 1 : 0; }();
  if (mode != 2 && !one_shot && B * H > 2 * sm_count()) {
    attention_fwd_persistent_kernel<<<2 * sm_count(), ATT_FWD_THREADS, ATT_FWD_SMEM, stream>>>(tq, p);
    DET_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  attention_fwd_kernel<<<B * H, ATT_FWD_THREADS, ATT_FWD_SMEM, stream>>>(tq, p);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// dctx: [B*S, H*64] bf16;  dqkv (out): [B*S, 3*H*64] bf16.
extern "C" int det_attention_bwd(const void* qkv, const void* dctx, const float* mask,
                                 const float* lse, void* dqkv, int B, int S, int H, float dropout_p,
                                 unsigned int seed, const unsigned int* seed_dev, cudaStream_t stream) {
  using namespace det;
  AttnParams p;
  int rc = fill_params(p, B, S, H, mask, const_cast<float*>(lse), dqkv, dropout_p, seed, seed_dev);
  if (rc) return rc;
  CUtensorMap tq, td;
  rc = make_tma_2d_bf16(&tq, qkv, 3ull * p.hidden, (uint64_t)B * S, 3ull * p.hidden, 64, ATT_S);
  if (rc) return rc;
  rc = make_tma_2d_bf16(&td, dctx, (uint64_t)p.hidden, (uint64_t)B * S, (uint64_t)p.hidden, 64, ATT_S);
  if (rc) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  DET_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && !attr_set[dev]) {
    DET_CHECK_CUDA(cudaFuncSetAttribute(attention_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_BWD_SMEM));
    DET_CHECK_CUDA(cudaFuncSetAttribute(attention_bwd_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_BWDP_SMEM));
    attr_set[dev] = true;
  }
  static const int one_shot = [] { const char* e = getenv("DET_ATTN_ONESHOT"); return e != nullptr && e[0] == '1' ? 1 : 0; }();
  if (!one_shot && B * H > sm_count()) {
    attention_bwd_persistent_kernel<<<std::min(B * H, sm_count()), ATT_BWD_THREADS, ATT_BWDP_SMEM, stream>>>(tq, td, p);
    DET_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  attention_bwd_kernel<<<B * H, ATT_BWD_THREADS, ATT_BWD_SMEM, stream>>>(tq, td, p);
  DET_CHECK_CUDA(cudaGetLastError());
  return 0;
}
