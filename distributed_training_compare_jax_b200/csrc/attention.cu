// Causal flash attention forward / backward on tcgen05 tensor cores (op sites K3/K6, reference:
// model/CausalSelfAttention.py:34-42 — dense softmax(q k^T * hd^-0.5 + causal mask) v; the (1,1,T,T)
// additive mask of model/GPTModel.py:50-51 is never materialised, the causal predicate is applied
// to the diagonal tile only and tiles above the diagonal are skipped).
//
// Layout: q, k, v are column slices of one [M = B*T, ld] bf16 buffer (head h occupies columns
// [h*HD, (h+1)*HD) of each slice); o / do / dq / dk / dv use the same convention.
//
// Forward  (flash_attn_fwd_pipe_kernel; CTA = one (batch, head, 128-query block), thread r of the softmax warps owns query row r):
//   S  = Q K_j^T      tcgen05.mma 128x64xHD    -> TMEM, double-buffered; issued one key block ahead by the issue warp
//   online softmax: tcgen05.ld of the thread's row, exp2, P (bf16) -> 128B-swizzled smem (K-major A operand)
//   PV = P V_j        tcgen05.mma 128xHDx64 (V consumed MN-major straight from its TMA tile) -> TMEM, folded one block late
// Backward (head_dim 32: flash_attn_bwd_persist_kernel, see there; head_dim 64: flash_attn_bwd_kernel —
//   CTA = one (batch, head, 128-key block j); loops over query blocks i >= j; thread r owns key row r):
//   St  = K_j Q_i^T ; dPt = V_j dO_i^T                         (TMEM cols [0,128), [128,256))
//   Pt  = exp2(St*c - lse2_i) ; dSt = Pt o (dPt - delta_i)     -> two swizzled smem tiles
//   dV += Pt dO_i ; dK += dSt Q_i     (accumulate in TMEM across i)
//   dQ_i = dS K_j                     (the dSt tile re-read as an MN-major A operand) -> fp32 red.add to dq_accum
#include <stdlib.h>

#include "common.cuh"
#include "tma_host.h"

namespace dtc {

constexpr int ATT_BLK = 128;
constexpr float LOG2E = 1.4426950408889634f;

template <int HD>
struct AttnCfg {
  static constexpr int ROW_BYTES = HD * 2;                       // 64 or 128
  static constexpr uint32_t SWZ = (HD == 64) ? SWZ_128B : SWZ_64B;
  static constexpr int SWZ_BITS = (HD == 64) ? 3 : 2;
  static constexpr int TILE_BYTES = ATT_BLK * ROW_BYTES;         // one [128 x HD] tile
  static constexpr uint32_t SBO = 8 * ROW_BYTES;                 // 8-row group pitch (both majors)
  static constexpr int P_BYTES = ATT_BLK * ATT_BLK * 2;          // [128 x 128] bf16, two 64-col blocks
};

// P / dS tile: [128 rows x 128 cols] bf16 stored as two [128 x 64] K-major 128B-swizzled blocks.
DTC_DEVICE uint32_t p_tile_offset(int row, int col8 /* 16-byte chunk index 0..15 */) {
  const uint32_t blk = col8 >> 3;
  return blk * (ATT_BLK * 128) + swizzle_offset<3>(row * 128 + (col8 & 7) * 16);
}

// Forward, software-pipelined (default).  Same math as the kernels above; what changes is who waits for whom:
//   * warp 4 is a dedicated issue warp (one elected lane): TMA loads of Q / K_j / V_j and every tcgen05.mma;
//   * warps 0..3 are the softmax warps (thread r owns query row r);
//   * S is double-buffered in tensor memory and Q K_{j+1}^T is issued BEFORE the softmax of block j starts, so the
//     tensor-core latency of the next block hides behind the exp2 work of this one;
//   * P is double-buffered in shared memory and the P V product of block j is folded into the register accumulator one
//     block late — O = (O + PV_{j-1}) * alpha_j — so no thread ever waits for the MMA it has just triggered.
// The only per-block synchronisation left is mbarrier traffic between the softmax threads and the issue lane (no
// __syncthreads in the loop).  Softmax inner loop: FMNMX, FFMA, MUFU.EX2, FADD and half a CVT per element; the causal
// predicate is evaluated on the (at most two) diagonal blocks only.
// TMEM: S 2 x 64 columns, PV 2 x HD columns -> 256 allocated -> two CTAs per SM.

template <int HD>
__global__ void __launch_bounds__(160, HD == 32 ? 2 : 1)
flash_attn_fwd_pipe_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                           const __grid_constant__ CUtensorMap tm_v, bf16* __restrict__ out, long long ldo,
                           float* __restrict__ lse2_out, int T, int H, float scale_log2) {
  using C = AttnCfg<HD>;
  constexpr int KVB = 64;
  constexpr int KV_BYTES = KVB * C::ROW_BYTES;
  constexpr int P64_BYTES = ATT_BLK * 128;  // [128 x 64] bf16, one K-major 128B-swizzled block
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sQ = smem;
  uint8_t* sP = sQ + C::TILE_BYTES;      // 2 buffers
  uint8_t* sK = sP + 2 * P64_BYTES;      // 2 stages
  uint8_t* sV = sK + 2 * KV_BYTES;       // 2 stages
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + 2 * KV_BYTES);
  uint64_t* bar_q = bars;
  uint64_t* bar_k = bars + 1;  // [2] K_j landed
  uint64_t* bar_v = bars + 3;  // [2] V_j landed
  uint64_t* bar_s = bars + 5;  // [2] S_j complete in tensor memory
  uint64_t* bar_p = bars + 7;  // [2] P_j written by the 4 softmax warps (and S_j / PV_{j-1} consumed)
  uint64_t* bar_o = bars + 9;  // [2] PV_j complete in tensor memory
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 11);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // Longest-processing-time-first over the WHOLE grid: CTAs are dispatched in linear block order, and only two fit an SM
  // (256 TMEM columns each), so the causal work (1..T/128 key blocks) is ordered globally — every (batch, head) pair's
  // heaviest query block first, then all second-heaviest ones, ... — instead of heavy and light blocks interleaved, which
  // left 4-unit CTAs in the last partial wave.
  const int lin = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  const int n_bh = gridDim.y * gridDim.z;
  const int qblk = gridDim.x - 1 - lin / n_bh;
  const int h = (lin % n_bh) % H, b = (lin % n_bh) / H;
  const int row0 = b * T + qblk * ATT_BLK;
  const int nkv = 2 * (qblk + 1);  // causal: 64-key blocks 0 .. 2*qblk+1

  if (tid == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
    mbar_init(bar_q, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&bar_k[s], 1);
      mbar_init(&bar_v[s], 1);
      mbar_init(&bar_s[s], 1);
      mbar_init(&bar_p[s], 4);  // one arrival per softmax warp
      mbar_init(&bar_o[s], 1);
    }
    fence_barrier_init();
  }
  if (warp == 4) {
    tmem_alloc(tmem_ptr_smem, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_launch_dependents();
  pdl_wait();
  // columns: S0 [0,64) S1 [64,128) PV0 [128,128+HD) PV1 [128+HD,128+2HD)
  constexpr uint32_t idesc_s = make_idesc_bf16(128, KVB, 0, 0);
  constexpr uint32_t idesc_pv = make_idesc_bf16(128, HD, 0, 1);

  if (warp == 4) {
    // ------------------------------ issue warp ------------------------------
    if (elect_one()) {
      auto load_k = [&](int j) {
        mbar_expect_tx(&bar_k[j & 1], KV_BYTES);
        tma_load_2d(sK + (j & 1) * KV_BYTES, &tm_k, &bar_k[j & 1], h * HD, b * T + j * KVB);
      };
      auto load_v = [&](int j) {
        mbar_expect_tx(&bar_v[j & 1], KV_BYTES);
        tma_load_2d(sV + (j & 1) * KV_BYTES, &tm_v, &bar_v[j & 1], h * HD, b * T + j * KVB);
      };
      auto issue_qk = [&](int j) {  // S[j&1] = Q K_j^T
        mbar_wait(&bar_k[j & 1], (j >> 1) & 1);
        tc_fence_after();
        const uint32_t aq = smem_u32(sQ), ak = smem_u32(sK + (j & 1) * KV_BYTES);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_bf16(tmem_base + (j & 1) * KVB, make_smem_desc(aq + k * 32, 16, C::SBO, C::SWZ),
                    make_smem_desc(ak + k * 32, 16, C::SBO, C::SWZ), idesc_s, k > 0);
        umma_commit(&bar_s[j & 1]);
      };
      mbar_expect_tx(bar_q, C::TILE_BYTES);
      tma_load_2d(sQ, &tm_q, bar_q, h * HD, row0);
      load_k(0);
      load_v(0);
      load_k(1);  // nkv >= 2 always
      load_v(1);
      mbar_wait(bar_q, 0);
      issue_qk(0);
      for (int j = 0; j < nkv; ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        if (j + 1 < nkv) issue_qk(j + 1);  // S[(j+1)&1] was consumed before bar_p of block j-1 completed (waited below)
        // K stage `st` is free once Q K_j^T has completed: fetch K_{j+2} into it
        if (j + 2 < nkv) {
          mbar_wait(&bar_s[st], ph);
          load_k(j + 2);
        }
        mbar_wait(&bar_p[st], ph);  // P_j is in shared memory; the softmax warps have also read S_j and PV_{j-1}
        mbar_wait(&bar_v[st], ph);
        tc_fence_after();
        const uint32_t ap = smem_u32(sP + st * P64_BYTES), av = smem_u32(sV + st * KV_BYTES);
#pragma unroll
        for (int k = 0; k < KVB / 16; ++k)
          umma_bf16(tmem_base + 2 * KVB + st * HD, make_smem_desc(ap + k * 32, 16, 1024, SWZ_128B),
                    make_smem_desc(av + k * 16 * C::ROW_BYTES, 16, C::SBO, C::SWZ), idesc_pv, k > 0);
        umma_commit(&bar_o[st]);
        // V stage of block j-1 is free once P V_{j-1} has completed: fetch V_{j+1} into it
        if (j >= 1 && j + 1 < nkv) {
          mbar_wait(&bar_o[st ^ 1], ((j - 1) >> 1) & 1);
          load_v(j + 1);
        }
      }
    }
  } else {
    // ------------------------------ softmax warps (thread = query row) ------------------------------
    const int qrow = qblk * ATT_BLK + tid;  // position of this thread's query inside the sequence
    const uint32_t lane_sel = (uint32_t)(warp * 32) << 16;
    float o_acc[HD];
#pragma unroll
    for (int e = 0; e < HD; ++e) o_acc[e] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;

    for (int j = 0; j < nkv; ++j) {
      const int st = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      mbar_wait(&bar_s[st], ph);
      tc_fence_after();
      uint32_t r[64];
      tmem_ld_32x32b_x32(tmem_base + st * KVB + lane_sel, r);
      tmem_ld_32x32b_x32(tmem_base + st * KVB + lane_sel + 32, r + 32);
      tmem_ld_wait();
      const int lim = qrow - j * KVB;  // key column c of this block is visible iff c <= lim
      if (lim < KVB - 1) {             // diagonal block: mask (uniform per warp except inside the two diagonal warps)
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i > lim) r[i] = 0xff800000u;  // -inf
      }
      // four independent partial maxima / sums: a single 64-long dependent FMNMX (or FADD) chain costs ~250 cycles of pure
      // latency per block, which the two warps per scheduler cannot hide
      float mx4[4] = {__uint_as_float(r[0]), __uint_as_float(r[1]), __uint_as_float(r[2]), __uint_as_float(r[3])};
#pragma unroll
      for (int i = 4; i < 64; ++i) mx4[i & 3] = fmaxf(mx4[i & 3], __uint_as_float(r[i]));
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      // m_run is finite from block 0 on (key 0 is visible to every query); a fully masked block leaves it unchanged
      const float m_new = fmaxf(m_run, mx * scale_log2);
      const float alpha = ex2_ftz(m_run - m_new);
      float l4[4] = {0.f, 0.f, 0.f, 0.f};
      uint8_t* prow = sP + st * P64_BYTES;
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        float p[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          p[i] = ex2_ftz(fmaf(__uint_as_float(r[g * 8 + i]), scale_log2, -m_new));
          l4[i & 3] += p[i];
        }
        *reinterpret_cast<uint4*>(prow + swizzle_offset<3>(tid * 128 + g * 16)) =
            make_uint4(pack_bf16x2(p[0], p[1]), pack_bf16x2(p[2], p[3]), pack_bf16x2(p[4], p[5]), pack_bf16x2(p[6], p[7]));
      }
      l_run = l_run * alpha + ((l4[0] + l4[1]) + (l4[2] + l4[3]));
      m_run = m_new;
      // fold the product of the PREVIOUS block (its MMA was issued a whole softmax ago) before P_j is published:
      // bar_p also tells the issue warp that PV[(j-1)&1] may be overwritten by block j+1
      if (j >= 1) {
        mbar_wait(&bar_o[st ^ 1], ((j - 1) >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < HD / 32; ++c) {
          uint32_t pv[32];
          tmem_ld_32x32b_x32(tmem_base + 2 * KVB + (st ^ 1) * HD + lane_sel + c * 32, pv);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] = (o_acc[c * 32 + i] + __uint_as_float(pv[i])) * alpha;
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_p[st]);
    }
    {  // last block's product
      const int jl = nkv - 1;
      mbar_wait(&bar_o[jl & 1], (jl >> 1) & 1);
      tc_fence_after();
      const float inv_l = 1.f / l_run;
      bf16* orow = out + (long long)(row0 + tid) * ldo + h * HD;
#pragma unroll
      for (int c = 0; c < HD / 32; ++c) {
        uint32_t pv[32];
        tmem_ld_32x32b_x32(tmem_base + 2 * KVB + (jl & 1) * HD + lane_sel + c * 32, pv);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] = (o_acc[c * 32 + i] + __uint_as_float(pv[i])) * inv_l;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int e = c * 32 + g * 8;
          *reinterpret_cast<uint4*>(orow + e) =
              make_uint4(pack_bf16x2(o_acc[e], o_acc[e + 1]), pack_bf16x2(o_acc[e + 2], o_acc[e + 3]),
                         pack_bf16x2(o_acc[e + 4], o_acc[e + 5]), pack_bf16x2(o_acc[e + 6], o_acc[e + 7]));
        }
      }
      lse2_out[((long long)b * H + h) * T + qblk * ATT_BLK + tid] = m_run + log2f(l_run);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

// delta[b,h,t] = sum_d dO[b,t,h,d] * O[b,t,h,d]   (one thread per 8 columns, HD/8 threads per (row, head))
template <int HD>
__global__ void attn_bwd_prep_kernel(const bf16* __restrict__ o, long long ldo, const bf16* __restrict__ d_o,
                                     long long lddo, float* __restrict__ delta, int M, int T, int H) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  constexpr int TPR = HD / 8;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long item = idx / TPR;
  const int part = (int)(idx % TPR);
  const bool ok = item < (long long)M * H;
  float acc = 0.f;
  int m = 0, h = 0;
  if (ok) {
    m = (int)(item / H);
    h = (int)(item % H);
    const uint4 a = *reinterpret_cast<const uint4*>(o + (long long)m * ldo + h * HD + part * 8);
    const uint4 g = *reinterpret_cast<const uint4*>(d_o + (long long)m * lddo + h * HD + part * 8);
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 fa = unpack_bf16x2(aw[j]), fg = unpack_bf16x2(gw[j]);
      acc += fa.x * fg.x + fa.y * fg.y;
    }
  }
#pragma unroll
  for (int off = TPR / 2; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if (ok && part == 0) {
    const int b = m / T, t = m % T;
    delta[((long long)b * H + h) * T + t] = acc;
  }
}

template <int HD>
__global__ void __launch_bounds__(128)
flash_attn_bwd_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                      const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_do,
                      const float* __restrict__ lse2, const float* __restrict__ delta, float* __restrict__ dq_accum,
                      long long lddq, bf16* __restrict__ dk, bf16* __restrict__ dv, long long lddkv, int T, int H,
                      float scale, float scale_log2) {
  using C = AttnCfg<HD>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sK = smem;
  uint8_t* sV = sK + C::TILE_BYTES;
  uint8_t* sQ = sV + C::TILE_BYTES;       // 2 stages
  uint8_t* sDO = sQ + 2 * C::TILE_BYTES;  // 2 stages
  uint8_t* sPt = sDO + 2 * C::TILE_BYTES;
  uint8_t* sDSt = sPt + C::P_BYTES;
  float* s_lse = reinterpret_cast<float*>(sDSt + C::P_BYTES);  // [2][128]
  float* s_delta = s_lse + 2 * ATT_BLK;                        // [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_delta + 2 * ATT_BLK);
  uint64_t* bar_kv = bars;
  uint64_t* bar_qdo = bars + 1;  // [2]
  uint64_t* bar_s = bars + 3;
  uint64_t* bar_d = bars + 4;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 5);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int jblk = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int nblk = T / ATT_BLK;
  const int kv_row0 = b * T + jblk * ATT_BLK;
  const long long stat_base = ((long long)b * H + h) * T;

  if (tid == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
    tma_prefetch_desc(&tm_do);
    mbar_init(bar_kv, 1);
    mbar_init(&bar_qdo[0], 1);
    mbar_init(&bar_qdo[1], 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_d, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_ptr_smem, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_launch_dependents();  // PDL: only after this CTA owns its TMEM (a dependent may not allocate first)
  pdl_wait();               // inputs come from the previous kernel: no global access before it has completed
  const uint32_t tm_st = tmem_base, tm_dpt = tmem_base + 128, tm_dk = tmem_base + 256, tm_dv = tmem_base + 256 + HD,
                 tm_dq = tmem_base + 256 + 2 * HD;
  const uint32_t lane_sel = (uint32_t)(warp * 32) << 16;

  if (tid == 0) {
    mbar_expect_tx(bar_kv, 2 * C::TILE_BYTES);
    tma_load_2d(sK, &tm_k, bar_kv, h * HD, kv_row0);
    tma_load_2d(sV, &tm_v, bar_kv, h * HD, kv_row0);
    mbar_expect_tx(&bar_qdo[0], 2 * C::TILE_BYTES);
    tma_load_2d(sQ, &tm_q, &bar_qdo[0], h * HD, b * T + jblk * ATT_BLK);
    tma_load_2d(sDO, &tm_do, &bar_qdo[0], h * HD, b * T + jblk * ATT_BLK);
  }

  constexpr uint32_t idesc_kk = make_idesc_bf16(128, 128, 0, 0);  // St, dPt : A K-major, B K-major
  constexpr uint32_t idesc_kn = make_idesc_bf16(128, HD, 0, 1);   // dV, dK  : A K-major (Pt/dSt), B MN-major
  constexpr uint32_t idesc_nn = make_idesc_bf16(128, HD, 1, 1);   // dQ      : A MN-major (dSt as dS), B MN-major (K)

  int it = 0;
  for (int i = jblk; i < nblk; ++i, ++it) {
    const int st = it & 1;
    // stage the per-row statistics of query block i
    s_lse[st * ATT_BLK + tid] = lse2[stat_base + i * ATT_BLK + tid];
    s_delta[st * ATT_BLK + tid] = delta[stat_base + i * ATT_BLK + tid];
    if (it == 0) mbar_wait(bar_kv, 0);
    mbar_wait(&bar_qdo[st], (it >> 1) & 1);
    tc_fence_after();
    __syncthreads();
    if (tid == 0) {
      const uint32_t ak = smem_u32(sK), av = smem_u32(sV), aq = smem_u32(sQ + st * C::TILE_BYTES),
                     ado = smem_u32(sDO + st * C::TILE_BYTES);
#pragma unroll
      for (int k = 0; k < HD / 16; ++k)
        umma_bf16(tm_st, make_smem_desc(ak + k * 32, 16, C::SBO, C::SWZ), make_smem_desc(aq + k * 32, 16, C::SBO, C::SWZ),
                  idesc_kk, k > 0);
#pragma unroll
      for (int k = 0; k < HD / 16; ++k)
        umma_bf16(tm_dpt, make_smem_desc(av + k * 32, 16, C::SBO, C::SWZ), make_smem_desc(ado + k * 32, 16, C::SBO, C::SWZ),
                  idesc_kk, k > 0);
      umma_commit(bar_s);
      if (i + 1 < nblk) {  // prefetch Q / dO of the next query block
        const int ns = st ^ 1;
        mbar_expect_tx(&bar_qdo[ns], 2 * C::TILE_BYTES);
        tma_load_2d(sQ + ns * C::TILE_BYTES, &tm_q, &bar_qdo[ns], h * HD, b * T + (i + 1) * ATT_BLK);
        tma_load_2d(sDO + ns * C::TILE_BYTES, &tm_do, &bar_qdo[ns], h * HD, b * T + (i + 1) * ATT_BLK);
      }
    }
    mbar_wait(bar_s, it & 1);
    tc_fence_after();

    const bool diag = (i == jblk);
    const float* lse_i = s_lse + st * ATT_BLK;
    const float* del_i = s_delta + st * ATT_BLK;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t rs[32], rp[32];
      tmem_ld_32x32b_x32(tm_st + lane_sel + c * 32, rs);
      tmem_ld_32x32b_x32(tm_dpt + lane_sel + c * 32, rp);
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float p[8], ds[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int q = c * 32 + g * 8 + e;  // query column inside the block; this thread's key row is `tid`
          const float s = __uint_as_float(rs[g * 8 + e]);
          const float pv = (!diag || q >= tid) ? exp2f(s * scale_log2 - lse_i[q]) : 0.f;
          p[e] = pv;
          ds[e] = pv * (__uint_as_float(rp[g * 8 + e]) - del_i[q]);
        }
        const uint32_t off = p_tile_offset(tid, c * 4 + g);
        *reinterpret_cast<uint4*>(sPt + off) =
            make_uint4(pack_bf16x2(p[0], p[1]), pack_bf16x2(p[2], p[3]), pack_bf16x2(p[4], p[5]), pack_bf16x2(p[6], p[7]));
        *reinterpret_cast<uint4*>(sDSt + off) = make_uint4(pack_bf16x2(ds[0], ds[1]), pack_bf16x2(ds[2], ds[3]),
                                                           pack_bf16x2(ds[4], ds[5]), pack_bf16x2(ds[6], ds[7]));
      }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint32_t apt = smem_u32(sPt), adst = smem_u32(sDSt), aq = smem_u32(sQ + st * C::TILE_BYTES),
                     ado = smem_u32(sDO + st * C::TILE_BYTES), ak = smem_u32(sK);
      const uint32_t acc = it > 0;
#pragma unroll
      for (int k = 0; k < ATT_BLK / 16; ++k) {
        const uint32_t a_off = (k >> 2) * (ATT_BLK * 128) + (k & 3) * 32;  // K-major walk over the 128 q columns
        // dV += Pt[kv, q] * dO[q, :]      (B = dO tile consumed MN-major: rows are the reduction index)
        umma_bf16(tm_dv, make_smem_desc(apt + a_off, 16, 1024, SWZ_128B),
                  make_smem_desc(ado + k * 16 * C::ROW_BYTES, 16, C::SBO, C::SWZ), idesc_kn, acc | (k > 0));
      }
#pragma unroll
      for (int k = 0; k < ATT_BLK / 16; ++k) {
        const uint32_t a_off = (k >> 2) * (ATT_BLK * 128) + (k & 3) * 32;
        // dK += dSt[kv, q] * Q[q, :]
        umma_bf16(tm_dk, make_smem_desc(adst + a_off, 16, 1024, SWZ_128B),
                  make_smem_desc(aq + k * 16 * C::ROW_BYTES, 16, C::SBO, C::SWZ), idesc_kn, acc | (k > 0));
      }
#pragma unroll
      for (int k = 0; k < ATT_BLK / 16; ++k) {
        // dQ = dS[q, kv] * K[kv, :] : the dSt tile read as an MN-major A operand (q contiguous): 64-wide q blocks
        // are 128*128 B apart (LBO), 8-kv-row groups 1024 B apart (SBO); one K step = 16 kv rows = 2048 B.
        umma_bf16(tm_dq, make_smem_desc(adst + k * 2048, ATT_BLK * 128, 1024, SWZ_128B),
                  make_smem_desc(ak + k * 16 * C::ROW_BYTES, 16, C::SBO, C::SWZ), idesc_nn, k > 0);
      }
      umma_commit(bar_d);
    }
    mbar_wait(bar_d, it & 1);
    tc_fence_after();
    {
      float* dq_row = dq_accum + (long long)(b * T + i * ATT_BLK + tid) * lddq + h * HD;
#pragma unroll
      for (int c = 0; c < HD / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tm_dq + lane_sel + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; e += 4)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dq_row + c * 32 + e),
                       "f"(__uint_as_float(r[e]) * scale), "f"(__uint_as_float(r[e + 1]) * scale),
                       "f"(__uint_as_float(r[e + 2]) * scale), "f"(__uint_as_float(r[e + 3]) * scale)
                       : "memory");
      }
    }
    tc_fence_before();
    __syncthreads();
  }

  // write dK (scaled) and dV for this key block
  bf16* dk_row = dk + (long long)(kv_row0 + tid) * lddkv + h * HD;
  bf16* dv_row = dv + (long long)(kv_row0 + tid) * lddkv + h * HD;
#pragma unroll
  for (int c = 0; c < HD / 32; ++c) {
    uint32_t rk[32], rv[32];
    tmem_ld_32x32b_x32(tm_dk + lane_sel + c * 32, rk);
    tmem_ld_32x32b_x32(tm_dv + lane_sel + c * 32, rv);
    tmem_ld_wait();
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int e = g * 8;
      *reinterpret_cast<uint4*>(dk_row + c * 32 + e) = make_uint4(
          pack_bf16x2(__uint_as_float(rk[e]) * scale, __uint_as_float(rk[e + 1]) * scale),
          pack_bf16x2(__uint_as_float(rk[e + 2]) * scale, __uint_as_float(rk[e + 3]) * scale),
          pack_bf16x2(__uint_as_float(rk[e + 4]) * scale, __uint_as_float(rk[e + 5]) * scale),
          pack_bf16x2(__uint_as_float(rk[e + 6]) * scale, __uint_as_float(rk[e + 7]) * scale));
      *reinterpret_cast<uint4*>(dv_row + c * 32 + e) =
          make_uint4(pack_bf16x2(__uint_as_float(rv[e]), __uint_as_float(rv[e + 1])),
                     pack_bf16x2(__uint_as_float(rv[e + 2]), __uint_as_float(rv[e + 3])),
                     pack_bf16x2(__uint_as_float(rv[e + 4]), __uint_as_float(rv[e + 5])),
                     pack_bf16x2(__uint_as_float(rv[e + 6]), __uint_as_float(rv[e + 7])));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

DTC_DEVICE void red_add_v4_f32_attn(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Backward, persistent + software-pipelined (head_dim 32).  One CTA per SM walks a static list of work items
// (batch, head, 128-key block j), heaviest first; an item loops over the query blocks i >= j.  All (item, i) steps of a
// CTA form ONE flattened sequence g = 0, 1, 2 ... and every resource is ring-indexed by g (or by the item ordinal k), so
// the pipeline never drains between items: the issue warp loads K / V / Q / dO of the next item and issues its score
// MMAs while the softmax warps are still finishing the current one, and the dK / dV read-out of item k overlaps the
// first block of item k+1.  A launch therefore pays ONE prologue and one tail per SM instead of one per (b, h, j).
//   * scores in the S = Q_i K_j^T orientation (thread r owns QUERY row r): lse2 / delta are two per-thread scalars;
//   * 8 softmax warps: warps 0..3 take key columns [0,64), warps 4..7 take [64,128) of the same rows;
//   * warp 8 is the issue warp (one elected lane): all TMA loads and tcgen05.mma; S_{g+1} / dP_{g+1} are issued as soon as
//     every softmax warp has pulled S_g / dP_g into registers, i.e. they run under the exp2 work of step g;
//   * P / dS tiles double-buffered in shared memory: the products of step g (dV += P^T dO_i, dK += dS^T Q_i,
//     dQ_i = dS K_j) overlap the element-wise work of step g+1;
//   * dK / dV accumulate in tensor memory across the steps of an item (two accumulator sets, by item parity); dQ_i is
//     double-buffered in tensor memory and pushed to global (fp32 red.add) one step late.
// TMEM columns: S [0,128) dP [128,256) {dV,dK} x 2 [256,384) dQ x 2 [384,448) -> 512 allocated, 1 CTA / SM.
struct BwdCursor {
  int w, k, it, n_it, b, h, jblk;
  bool valid;
};
DTC_DEVICE void bwd_cursor_decode(BwdCursor& c, int n_items, int BH, int H, int nblk) {
  c.valid = c.w < n_items;
  if (!c.valid) return;
  c.jblk = c.w / BH;  // items are ordered by key block: block 0 (most query blocks) first
  const int rest = c.w - c.jblk * BH;
  c.b = rest / H;
  c.h = rest - c.b * H;
  c.n_it = nblk - c.jblk;
  c.it = 0;
}
DTC_DEVICE void bwd_cursor_next(BwdCursor& c, int stride, int n_items, int BH, int H, int nblk) {
  if (++c.it < c.n_it) return;
  c.w += stride;
  ++c.k;
  bwd_cursor_decode(c, n_items, BH, H, nblk);
}

template <int HD, int NG>
__global__ void __launch_bounds__(NG * 128 + 32, 1)
flash_attn_bwd_persist_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                              const __grid_constant__ CUtensorMap tm_v, const __grid_constant__ CUtensorMap tm_do,
                              const float* __restrict__ lse2, const float* __restrict__ delta,
                              float* __restrict__ dq_accum,
                              long long lddq, bf16* __restrict__ dk, bf16* __restrict__ dv, long long lddkv, int T, int H,
                              int B, float scale, float scale_log2) {
  using C = AttnCfg<HD>;
  static_assert(HD == 32, "the pipelined backward is laid out for head_dim 32");
  static_assert(NG == 2 || NG == 4, "2 or 4 column groups of softmax warps");
  constexpr int CW = ATT_BLK / NG;   // key columns of S / dP handled by one softmax warp
  constexpr int QC = HD / NG;        // columns of dQ / dK / dV read out by one softmax warp
  constexpr int ISSUE_WARP = 4 * NG;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sP = smem;                      // 2 buffers of [128 q x 128 k] bf16
  uint8_t* sDS = sP + 2 * C::P_BYTES;      // 2 buffers
  uint8_t* sK = sDS + 2 * C::P_BYTES;      // 2 buffers (item parity)
  uint8_t* sV = sK + 2 * C::TILE_BYTES;    // 2 buffers
  uint8_t* sQ = sV + 2 * C::TILE_BYTES;    // 2 stages (step parity)
  uint8_t* sDO = sQ + 2 * C::TILE_BYTES;   // 2 stages
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDO + 2 * C::TILE_BYTES);
  uint64_t* bar_kv = bars;      // [2] K_j / V_j of item k landed
  uint64_t* bar_ld = bars + 2;  // [2] Q_i and dO_i of step g landed
  uint64_t* bar_s = bars + 4;   //     S_g / dP_g complete
  uint64_t* bar_c = bars + 5;   //     S_g / dP_g pulled into registers by all 8 softmax warps
  uint64_t* bar_p = bars + 6;   // [2] P_g / dS_g written (and the deferred read-outs of step g-1 done)
  uint64_t* bar_d = bars + 8;   // [2] products of step g complete
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 10);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nblk = T / ATT_BLK;
  const int BH = B * H;
  const int n_items = BH * nblk;
  const int stride = gridDim.x;

  if (tid == 0) {
    tma_prefetch_desc(&tm_q);
    tma_prefetch_desc(&tm_k);
    tma_prefetch_desc(&tm_v);
    tma_prefetch_desc(&tm_do);
    mbar_init(bar_s, 1);
    mbar_init(bar_c, 4 * NG);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&bar_kv[s], 1);
      mbar_init(&bar_ld[s], 1);
      mbar_init(&bar_p[s], 4 * NG);
      mbar_init(&bar_d[s], 1);
    }
    fence_barrier_init();
  }
  if (warp == ISSUE_WARP) {
    tmem_alloc(tmem_ptr_smem, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_launch_dependents();
  pdl_wait();
  const uint32_t tm_s = tmem_base, tm_dp = tmem_base + 128;
  const uint32_t tm_acc = tmem_base + 256;        // + (k & 1) * 2 * HD : dV, then dK at + HD
  const uint32_t tm_dq = tmem_base + 256 + 4 * HD;  // + (g & 1) * HD

  BwdCursor first;
  first.w = blockIdx.x;
  first.k = 0;
  bwd_cursor_decode(first, n_items, BH, H, nblk);

  if (warp == ISSUE_WARP) {
    // ------------------------------ issue warp ------------------------------
    if (elect_one() && first.valid) {
      constexpr uint32_t idesc_kk = make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_kn = make_idesc_bf16(128, HD, 0, 1);
      constexpr uint32_t idesc_nn = make_idesc_bf16(128, HD, 1, 1);
      auto load_kv = [&](const BwdCursor& c) {
        const int kb = c.k & 1;
        const int kv_row0 = c.b * T + c.jblk * ATT_BLK;
        mbar_expect_tx(&bar_kv[kb], 2 * C::TILE_BYTES);
        tma_load_2d(sK + kb * C::TILE_BYTES, &tm_k, &bar_kv[kb], c.h * HD, kv_row0);
        tma_load_2d(sV + kb * C::TILE_BYTES, &tm_v, &bar_kv[kb], c.h * HD, kv_row0);
      };
      auto load_qdo = [&](const BwdCursor& c, int g) {
        const int st = g & 1;
        const int row = c.b * T + (c.jblk + c.it) * ATT_BLK;
        mbar_expect_tx(&bar_ld[st], 2 * C::TILE_BYTES);
        tma_load_2d(sQ + st * C::TILE_BYTES, &tm_q, &bar_ld[st], c.h * HD, row);
        tma_load_2d(sDO + st * C::TILE_BYTES, &tm_do, &bar_ld[st], c.h * HD, row);
      };
      auto issue_scores = [&](const BwdCursor& c, int g) {  // S = Q_i K_j^T, dP = dO_i V_j^T
        const int st = g & 1, kb = c.k & 1;
        if (c.it == 0) mbar_wait(&bar_kv[kb], (c.k >> 1) & 1);
        mbar_wait(&bar_ld[st], (g >> 1) & 1);
        tc_fence_after();
        const uint32_t ak = smem_u32(sK + kb * C::TILE_BYTES), av = smem_u32(sV + kb * C::TILE_BYTES),
                       aq = smem_u32(sQ + st * C::TILE_BYTES), ado = smem_u32(sDO + st * C::TILE_BYTES);
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk)
          umma_bf16(tm_s, make_smem_desc(aq + kk * 32, 16, C::SBO, C::SWZ), make_smem_desc(ak + kk * 32, 16, C::SBO, C::SWZ),
                    idesc_kk, kk > 0);
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk)
          umma_bf16(tm_dp, make_smem_desc(ado + kk * 32, 16, C::SBO, C::SWZ), make_smem_desc(av + kk * 32, 16, C::SBO, C::SWZ),
                    idesc_kk, kk > 0);
        umma_commit(bar_s);
      };
      BwdCursor cur = first, ld = first;
      load_kv(ld);
      load_qdo(ld, 0);
      bwd_cursor_next(ld, stride, n_items, BH, H, nblk);
      if (ld.valid) {
        if (ld.it == 0) load_kv(ld);
        load_qdo(ld, 1);
        bwd_cursor_next(ld, stride, n_items, BH, H, nblk);
      }
      issue_scores(cur, 0);
      for (int g = 0; cur.valid; ++g) {
        const int st = g & 1;
        const uint32_t ph = (g >> 1) & 1;
        // S_g / dP_g are in registers everywhere: the tensor cores may start on step g+1 under the exp2 work
        mbar_wait(bar_c, g & 1);
        BwdCursor nxt = cur;
        bwd_cursor_next(nxt, stride, n_items, BH, H, nblk);
        if (nxt.valid) issue_scores(nxt, g + 1);
        // P_g / dS_g are in shared memory (dQ buffer st and, for a new item, its accumulator set have been read out)
        mbar_wait(&bar_p[st], ph);
        tc_fence_after();
        const int kb = cur.k & 1;
        const uint32_t apt = smem_u32(sP + st * C::P_BYTES), adst = smem_u32(sDS + st * C::P_BYTES),
                       aq = smem_u32(sQ + st * C::TILE_BYTES), ado = smem_u32(sDO + st * C::TILE_BYTES),
                       ak = smem_u32(sK + kb * C::TILE_BYTES);
        const uint32_t tm_dv = tm_acc + kb * 2 * HD, tm_dk = tm_dv + HD;
        // dV += P^T dO_i ; dK += dS^T Q_i : A = the [q x k] tile read MN-major (M = keys contiguous, 16 query rows per step)
#pragma unroll
        for (int kk = 0; kk < ATT_BLK / 16; ++kk)
          umma_bf16(tm_dv, make_smem_desc(apt + kk * 2048, ATT_BLK * 128, 1024, SWZ_128B),
                    make_smem_desc(ado + kk * 16 * C::ROW_BYTES, 16, C::SBO, C::SWZ), idesc_nn,
                    (cur.it > 0 || kk > 0) ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < ATT_BLK / 16; ++kk)
          umma_bf16(tm_dk, make_smem_desc(adst + kk * 2048, ATT_BLK * 128, 1024, SWZ_128B),
                    make_smem_desc(aq + kk * 16 * C::ROW_BYTES, 16, C::SBO, C::SWZ), idesc_nn,
                    (cur.it > 0 || kk > 0) ? 1u : 0u);
        // dQ_i = dS K_j : A = the same tile K-major (two 64-key blocks), B = K_j MN-major
#pragma unroll
        for (int kk = 0; kk < ATT_BLK / 16; ++kk) {
          const uint32_t a_off = (kk >> 2) * (ATT_BLK * 128) + (kk & 3) * 32;
          umma_bf16(tm_dq + st * HD, make_smem_desc(adst + a_off, 16, 1024, SWZ_128B),
                    make_smem_desc(ak + kk * 16 * C::ROW_BYTES, 16, C::SBO, C::SWZ), idesc_kn, kk > 0);
        }
        umma_commit(&bar_d[st]);
        // operands of step g (Q / dO stage st; K / V of an item whose last step this was) are free once its products
        // complete: fetch step g+2 — and the K / V of its item when it opens one — behind them
        if (ld.valid) {
          mbar_wait(&bar_d[st], ph);
          if (ld.it == 0) load_kv(ld);
          load_qdo(ld, g + 2);
          bwd_cursor_next(ld, stride, n_items, BH, H, nblk);
        }
        cur = nxt;
      }
    }
  } else {
    // ------------------------------ softmax warps ------------------------------
    const int row = (warp & 3) * 32 + lane;  // query row inside the block (TMEM lane)
    const int gq = warp >> 2;                // key-column group handled by this warp
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
    // read-outs owed for the previous step: its dQ block and, when it closed an item, that item's dK / dV
    auto flush = [&](const BwdCursor& p, int g) {
      const int st = g & 1;
      mbar_wait(&bar_d[st], (g >> 1) & 1);
      tc_fence_after();
      uint32_t rq[QC];
      tmem_ld_cols<QC>(tm_dq + st * HD + lane_sel + gq * QC, rq);
      const bool last = p.it == p.n_it - 1;
      uint32_t rk[QC], rv[QC];
      if (last) {
        tmem_ld_cols<QC>(tm_acc + (p.k & 1) * 2 * HD + lane_sel + gq * QC, rv);
        tmem_ld_cols<QC>(tm_acc + (p.k & 1) * 2 * HD + HD + lane_sel + gq * QC, rk);
      }
      tmem_ld_wait();
      float* dq_row = dq_accum + (long long)(p.b * T + (p.jblk + p.it) * ATT_BLK + row) * lddq + p.h * HD + gq * QC;
#pragma unroll
      for (int e = 0; e < QC; e += 4)
        red_add_v4_f32_attn(dq_row + e, __uint_as_float(rq[e]) * scale, __uint_as_float(rq[e + 1]) * scale,
                            __uint_as_float(rq[e + 2]) * scale, __uint_as_float(rq[e + 3]) * scale);
      if (last) {
        const long long kv_row = (long long)(p.b * T + p.jblk * ATT_BLK + row);
        bf16* dk_row = dk + kv_row * lddkv + p.h * HD + gq * QC;
        bf16* dv_row = dv + kv_row * lddkv + p.h * HD + gq * QC;
#pragma unroll
        for (int gg = 0; gg < QC / 8; ++gg) {
          const int e = gg * 8;
          float fk[8], fv[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            fk[u] = __uint_as_float(rk[e + u]) * scale;
            fv[u] = __uint_as_float(rv[e + u]);
          }
          *reinterpret_cast<uint4*>(dk_row + e) = make_uint4(pack_bf16x2(fk[0], fk[1]), pack_bf16x2(fk[2], fk[3]),
                                                             pack_bf16x2(fk[4], fk[5]), pack_bf16x2(fk[6], fk[7]));
          *reinterpret_cast<uint4*>(dv_row + e) = make_uint4(pack_bf16x2(fv[0], fv[1]), pack_bf16x2(fv[2], fv[3]),
                                                             pack_bf16x2(fv[4], fv[5]), pack_bf16x2(fv[6], fv[7]));
        }
      }
    };
    BwdCursor cur = first, prev = first;
    int g = 0;
    for (; cur.valid; ++g) {
      const int st = g & 1;
      const int i = cur.jblk + cur.it;
      const long long stat = ((long long)cur.b * H + cur.h) * T + i * ATT_BLK + row;
      const float lse_r = lse2[stat];
      // delta = rowsum(dO * O) comes from a pre-pass kernel: computing it in here was measured twice (global row loads:
      // 45 -> 66 us; from TMA-loaded O / dO tiles in shared memory: 45 -> 50 us) and lost both times
      const float delta_r = delta[stat];
      mbar_wait(bar_s, g & 1);
      tc_fence_after();
      uint32_t rs[CW], rp[CW];
#pragma unroll
      for (int c = 0; c < CW / 32; ++c) {
        tmem_ld_32x32b_x32(tm_s + lane_sel + gq * CW + c * 32, rs + c * 32);
        tmem_ld_32x32b_x32(tm_dp + lane_sel + gq * CW + c * 32, rp + c * 32);
      }
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_c);
      if (cur.it == 0) {  // diagonal block (i == j): key column c is visible to query row r iff c <= r
        const int lim = row - gq * CW;
#pragma unroll
        for (int e = 0; e < CW; ++e)
          if (e > lim) rs[e] = 0xff800000u;  // -inf -> P = 0
      }
      uint8_t* pbuf = sP + st * C::P_BYTES;
      uint8_t* dsbuf = sDS + st * C::P_BYTES;
#pragma unroll
      for (int gg = 0; gg < CW / 8; ++gg) {
        float p[8], ds[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          p[e] = ex2_ftz(fmaf(__uint_as_float(rs[gg * 8 + e]), scale_log2, -lse_r));
          ds[e] = p[e] * (__uint_as_float(rp[gg * 8 + e]) - delta_r);
        }
        const uint32_t off = p_tile_offset(row, gq * (CW / 8) + gg);
        *reinterpret_cast<uint4*>(pbuf + off) =
            make_uint4(pack_bf16x2(p[0], p[1]), pack_bf16x2(p[2], p[3]), pack_bf16x2(p[4], p[5]), pack_bf16x2(p[6], p[7]));
        *reinterpret_cast<uint4*>(dsbuf + off) = make_uint4(pack_bf16x2(ds[0], ds[1]), pack_bf16x2(ds[2], ds[3]),
                                                            pack_bf16x2(ds[4], ds[5]), pack_bf16x2(ds[6], ds[7]));
      }
      if (g >= 1) flush(prev, g - 1);  // issued a whole element-wise pass ago: complete by now
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_p[st]);
      prev = cur;
      bwd_cursor_next(cur, stride, n_items, BH, H, nblk);
    }
    if (g >= 1) flush(prev, g - 1);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == ISSUE_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

template <int HD>
static int attn_tmap(CUtensorMap* m, const void* p, long long rows, int H, long long ld) {
  return make_tmap_2d_bf16(m, p, rows, (uint64_t)H * HD, ld, ATT_BLK, HD,
                           HD == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B);
}

template <int HD>
static int launch_fwd(const void* q, const void* k, const void* v, long long ld, void* out, long long ldo, float* lse2,
                      int B, int H, int T, cudaStream_t stream) {
  using C = AttnCfg<HD>;
  CUtensorMap tq, tk64, tv64;
  int rc;
  const long long M = (long long)B * T;
  if ((rc = attn_tmap<HD>(&tq, q, M, H, ld))) return rc;
  const CUtensorMapSwizzle swz = HD == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  if ((rc = make_tmap_2d_bf16(&tk64, k, M, (uint64_t)H * HD, ld, 64, HD, swz))) return rc;  // 64-key blocks
  if ((rc = make_tmap_2d_bf16(&tv64, v, M, (uint64_t)H * HD, ld, 64, HD, swz))) return rc;
  const float scale = 1.f / sqrtf((float)HD);
  dim3 grid(T / ATT_BLK, H, B);
  const int smem_p = C::TILE_BYTES + 2 * ATT_BLK * 128 + 4 * 64 * C::ROW_BYTES + 1024 + 256;
  auto kern_p = flash_attn_fwd_pipe_kernel<HD>;
  static dtc::SmemOptIn configured_p;  // per device
  DTC_CHECK_CUDA(dtc::ensure_dyn_smem(kern_p, smem_p, configured_p));
  DTC_CHECK_CUDA(dtc::launch(kern_p, grid, 160, smem_p, stream, tq, tk64, tv64, reinterpret_cast<bf16*>(out), ldo, lse2, T, H,
                             scale * LOG2E));
  DTC_CHECK_LAUNCH();
  return 0;
}

// delta pre-pass, then head_dim 32: the persistent pipelined kernel; head_dim 64: the TMEM-accumulator kernel.
template <int HD>
static int launch_bwd(const void* q, const void* k, const void* v, long long ld, const void* o, long long ldo,
                      const void* d_o, long long lddo, const float* lse2, float* delta, float* dq_accum, long long lddq,
                      void* dk, void* dv, long long lddkv, int B, int H, int T, cudaStream_t stream) {
  using C = AttnCfg<HD>;
  CUtensorMap tq, tk, tv, tdo;
  int rc;
  const long long M = (long long)B * T;
  if ((rc = attn_tmap<HD>(&tq, q, M, H, ld))) return rc;
  if ((rc = attn_tmap<HD>(&tk, k, M, H, ld))) return rc;
  if ((rc = attn_tmap<HD>(&tv, v, M, H, ld))) return rc;
  if ((rc = attn_tmap<HD>(&tdo, d_o, M, H, lddo))) return rc;
  const float scale = 1.f / sqrtf((float)HD);
  {  // delta[b, h, t] = rowsum(dO * O)
    const long long threads = M * H * (HD / 8);
    DTC_CHECK_CUDA(dtc::launch(attn_bwd_prep_kernel<HD>, (int)((threads + 255) / 256), 256, 0, stream,
        reinterpret_cast<const bf16*>(o), ldo, reinterpret_cast<const bf16*>(d_o), lddo, delta, (int)M, T, H));
    DTC_CHECK_LAUNCH();
  }
  if constexpr (HD == 32) {
    static const bool ng2 = getenv("DTC_ATTN_BWD_NG2") != nullptr;  // A/B: 8 softmax warps per CTA instead of 16
    const int smem4 = 4 * C::P_BYTES + 8 * C::TILE_BYTES + 1024 + 256;
    auto kern4 = ng2 ? flash_attn_bwd_persist_kernel<HD, 2> : flash_attn_bwd_persist_kernel<HD, 4>;
    static dtc::SmemOptIn configured4[2];  // per instantiation, per device
    DTC_CHECK_CUDA(dtc::ensure_dyn_smem(kern4, smem4, configured4[ng2 ? 0 : 1]));
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int n_items = B * H * (T / ATT_BLK);
    const int grid4 = n_items < sms ? n_items : sms;
    DTC_CHECK_CUDA(dtc::launch(kern4, grid4, ng2 ? 288 : 544, smem4, stream, tq, tk, tv, tdo, lse2, delta, dq_accum, lddq,
                               reinterpret_cast<bf16*>(dk), reinterpret_cast<bf16*>(dv), lddkv, T, H, B, scale,
                               scale * LOG2E));
    DTC_CHECK_LAUNCH();
    return 0;
  } else {
    const int smem = 6 * C::TILE_BYTES + 2 * C::P_BYTES + 4 * ATT_BLK * 4 + 1024 + 128;
    auto kern = flash_attn_bwd_kernel<HD>;
    static dtc::SmemOptIn configured;  // per device
    DTC_CHECK_CUDA(dtc::ensure_dyn_smem(kern, smem, configured));
    dim3 grid(T / ATT_BLK, H, B);
    DTC_CHECK_CUDA(dtc::launch(kern, grid, 128, smem, stream, tq, tk, tv, tdo, lse2, delta, dq_accum, lddq,
                               reinterpret_cast<bf16*>(dk), reinterpret_cast<bf16*>(dv), lddkv, T, H, scale, scale * LOG2E));
    DTC_CHECK_LAUNCH();
    return 0;
  }
}

}  // namespace dtc

using namespace dtc;

// q/k/v: bf16 pointers to the first column of each slice inside a [B*T, ld] buffer; out [B*T, ldo] bf16;
// lse2 [B,H,T] fp32 (log2-domain log-sum-exp, consumed by the backward).  T % 128 == 0, head_dim 32 or 64.
extern "C" int dtc_flash_attn_fwd(const void* q, const void* k, const void* v, long long ld, void* out, long long ldo,
                                  float* lse2, int B, int H, int T, int head_dim, cudaStream_t stream) {
  DTC_CHECK_ARG(B > 0 && H > 0 && T > 0 && T % ATT_BLK == 0, "flash_attn_fwd: T=%d must be a multiple of %d", T, ATT_BLK);
  DTC_CHECK_ARG(head_dim == 32 || head_dim == 64, "flash_attn_fwd: head_dim %d not supported (32, 64)", head_dim);
  DTC_CHECK_ARG(ld % 8 == 0 && ldo % 8 == 0, "flash_attn_fwd: leading dimensions must be multiples of 8");
  DTC_CHECK_ARG((((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)out) & 15) == 0, "flash_attn_fwd: alignment");
  if (head_dim == 32) return launch_fwd<32>(q, k, v, ld, out, ldo, lse2, B, H, T, stream);
  return launch_fwd<64>(q, k, v, ld, out, ldo, lse2, B, H, T, stream);
}

// dq_accum [B*T, lddq] fp32 must be ZERO on entry (partial dQ tiles are reduce-added);
// dk/dv [B*T, lddkv] bf16 are overwritten; delta [B,H,T] fp32 is scratch written by the prep kernel.
extern "C" int dtc_flash_attn_bwd(const void* q, const void* k, const void* v, long long ld, const void* o, long long ldo,
                                  const void* d_o, long long lddo, const float* lse2, float* delta, float* dq_accum,
                                  long long lddq, void* dk, void* dv, long long lddkv, int B, int H, int T, int head_dim,
                                  cudaStream_t stream) {
  DTC_CHECK_ARG(B > 0 && H > 0 && T > 0 && T % ATT_BLK == 0, "flash_attn_bwd: T=%d must be a multiple of %d", T, ATT_BLK);
  DTC_CHECK_ARG(head_dim == 32 || head_dim == 64, "flash_attn_bwd: head_dim %d not supported (32, 64)", head_dim);
  DTC_CHECK_ARG(ld % 8 == 0 && ldo % 8 == 0 && lddo % 8 == 0 && lddq % 4 == 0 && lddkv % 8 == 0,
                "flash_attn_bwd: leading dimension alignment");
  if (head_dim == 32)
    return launch_bwd<32>(q, k, v, ld, o, ldo, d_o, lddo, lse2, delta, dq_accum, lddq, dk, dv, lddkv, B, H, T, stream);
  return launch_bwd<64>(q, k, v, ld, o, ldo, d_o, lddo, lse2, delta, dq_accum, lddq, dk, dv, lddkv, B, H, T, stream);
}

// ---------------------------------------------------------------------------------------------------------
// Allocation-free attention backward.
//
// dtc_flash_attn_bwd needs a [B,H,T] fp32 scratch buffer (`delta`) that the caller has to allocate.  Some
// callers cannot do that: plugins and custom calls hosted by a runtime that owns every device allocation and
// forbids cudaMalloc from foreign code, step loops that cannot afford the device-wide synchronisation a
// cudaMalloc / cudaFree implies, and jobs whose device-memory budget is fixed before training starts.  For them
// this variant keeps the scratch in a fixed array that is part of the library image.  Every GPU that loads the library holds these
// DTC_ATTN_STATIC_DELTA_MAX * 4 bytes (1 MiB) from load until the process exits.
//
// The array name below is only meaningful to device code; on the host it is a stand-in symbol, not a device
// address.  The host therefore asks the runtime for the real address (cudaGetSymbolAddress) once per device,
// with that device current, and hands that pointer to the kernels.
// ---------------------------------------------------------------------------------------------------------
#define DTC_ATTN_STATIC_DELTA_MAX (1 << 18) /* floats: B*H*T may not exceed this (reference config uses 65536) */
#define DTC_MAX_DEVICES 64

static __device__ __align__(256) float g_attn_delta_scratch[DTC_ATTN_STATIC_DELTA_MAX];
static float* g_attn_delta_devptr[DTC_MAX_DEVICES];  // per-device resolved address, filled on first use

static int attn_static_delta_ptr(float** out) {
  int dev = -1;
  DTC_CHECK_CUDA(cudaGetDevice(&dev));
  DTC_CHECK_ARG(dev >= 0 && dev < DTC_MAX_DEVICES, "flash_attn_bwd_static: device index %d out of range", dev);
  if (g_attn_delta_devptr[dev] == nullptr) {
    void* p = nullptr;
    DTC_CHECK_CUDA(cudaGetSymbolAddress(&p, g_attn_delta_scratch));
    DTC_CHECK_ARG(p != nullptr, "flash_attn_bwd_static: cudaGetSymbolAddress returned a null address");
    g_attn_delta_devptr[dev] = static_cast<float*>(p);
  }
  *out = g_attn_delta_devptr[dev];
  return 0;
}

// Same contract as dtc_flash_attn_bwd (dq_accum fp32 must be zero on entry and receives dQ; dk / dv are
// overwritten) except that no scratch pointer is passed: the per-row delta = rowsum(dO * O) lives in the
// library's static device array.
//  * Limit: B*H*T <= DTC_ATTN_STATIC_DELTA_MAX (262144); larger problems are rejected with an error.
//  * Calls no cudaMalloc / cudaFree / allocating library routine.
//  * The first call on each device resolves the array's device address; make that first call outside of any
//    stream capture (e.g. during warm-up), with the device current.
//  * All calls on one GPU share the array and it keeps the previous call's contents.  Nothing in it has to
//    start at zero — the first kernel overwrites every element the second one reads — but two calls that
//    overlap on the GPU (different streams, or graph branches running concurrently) race on it.
extern "C" int dtc_flash_attn_bwd_static(const void* q, const void* k, const void* v, long long ld, const void* o,
                                         long long ldo, const void* d_o, long long lddo, const float* lse2,
                                         float* dq_accum, long long lddq, void* dk, void* dv, long long lddkv, int B,
                                         int H, int T, int head_dim, cudaStream_t stream) {
  DTC_CHECK_ARG(B > 0 && H > 0 && T > 0 && T % ATT_BLK == 0, "flash_attn_bwd_static: T=%d must be a multiple of %d", T,
                ATT_BLK);
  DTC_CHECK_ARG(head_dim == 32 || head_dim == 64, "flash_attn_bwd_static: head_dim %d not supported (32, 64)", head_dim);
  DTC_CHECK_ARG(ld % 8 == 0 && ldo % 8 == 0 && lddo % 8 == 0 && lddq % 4 == 0 && lddkv % 8 == 0,
                "flash_attn_bwd_static: leading dimension alignment");
  const long long need = (long long)B * H * T;
  DTC_CHECK_ARG(need <= DTC_ATTN_STATIC_DELTA_MAX,
                "flash_attn_bwd_static: B*H*T = %lld exceeds the static scratch capacity of %d floats; "
                "use dtc_flash_attn_bwd with a caller-provided delta buffer",
                need, DTC_ATTN_STATIC_DELTA_MAX);
  float* delta = nullptr;
  const int rc = attn_static_delta_ptr(&delta);
  if (rc) return rc;
  if (head_dim == 32)
    return launch_bwd<32>(q, k, v, ld, o, ldo, d_o, lddo, lse2, delta, dq_accum, lddq, dk, dv, lddkv, B, H, T, stream);
  return launch_bwd<64>(q, k, v, ld, o, ldo, d_o, lddo, lse2, delta, dq_accum, lddq, dk, dv, lddkv, B, H, T, stream);
}
