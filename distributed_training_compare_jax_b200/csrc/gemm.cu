// tcgen05 / TMEM / TMA GEMM for sm_100a with fused epilogues.
//
//   D[M,N] = A[M,K] * B[N,K]^T      bf16 operands, fp32 accumulation in tensor memory
//
// Either operand may be stored K-major (reduction dim contiguous) or MN-major (row dim contiguous),
// which covers all three GEMMs of a Linear layer without materialising a transpose:
//   forward   Y  = X  W          A = X  (K-major)   B = W  [K,N] (MN-major)
//   dgrad     dX = dY W^T        A = dY (K-major)   B = W  [K,N] read as [N_out=K, red=N] (K-major)
//   wgrad     dW = X^T dY        A = X  [M,K] read as [K, red=M] (MN-major)   B = dY (MN-major)
// Covers the op sites K5/K7/K8/K9/K10/K11/K16 of SURVEY.md §2.5 (reference: model/CausalSelfAttention.py:18-20,46,
// model/MLP.py:13-24, model/GPTModel.py:72 and their autodiff transposes).
//
// Structure (one 128 x BN output tile per CTA, 192 threads):
//   warp 0      TMA producer  : cp.async.bulk.tensor -> 128B-swizzled smem ring, mbarrier complete_tx
//   warp 1      MMA issuer    : one elected lane issues tcgen05.mma (UMMA 128 x BN x 16), tcgen05.commit frees stages
//   warps 2..5  epilogue      : tcgen05.ld (one accumulator row per thread) -> bias / GELU / residual /
//                               GELU-backward / fp32 reduce-add (split-K) -> 16-byte global stores
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "spsync.cuh"
#include "tma_host.h"

namespace dtc {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one swizzle row
constexpr int MAX_STAGES = 8;
constexpr int GEMM_THREADS = 192;

enum GemmEpiMode : int {
  EPI_STORE_BF16 = 0,     // out_bf16 = alpha*acc + bias
  EPI_BIAS_GELU = 1,      // out_bf16 = acc + bias ; out2_bf16 = gelu_tanh(acc + bias)
  EPI_BIAS_RESIDUAL = 2,  // out_bf16 = acc + bias + aux
  EPI_DGELU = 3,          // out_bf16 = acc * gelu_tanh'(aux)
  EPI_ATOMIC_F32 = 4,     // out_f32 += alpha*acc      (split-K / gradient accumulation)
  EPI_STORE_F32 = 5,      // out_f32 = alpha*acc + bias
  // (6 was the GEMM -> multimem.red all-reduce epilogue of round 1: slower than GEMM + NCCL at every rank count and
  // superseded by the reduce-scatter epilogue below; removed)
  // Fused GEMM -> reduce-scatter for sequence-parallel TP: output rows [o*sp_rows, (o+1)*sp_rows) belong to rank o; the
  // bf16 tile is stored over NVLink into THIS rank's partial-sum slot of rank o's symmetric buffer
  // (peer_out[o] + (sp_rank*sp_rows + row - o*sp_rows)*ldo), tile by tile while later tiles are still in the tensor
  // cores.  The owner sums the n slots in fp32 inside its next kernel (csrc/seqpar.cu).  Persistent kernel only.
  EPI_SCATTER_BF16 = 7,
  // out_bf16 = alpha*acc + bias, and per-row softmax statistics of what was stored: for every (row, 128-column part of a
  // tile) one float2 {max, sum exp(x - max)} over the part's VALID columns (n < n_valid; vocab padding excluded), written
  // to stats[row * n_parts + 2*n_blk + half] (`out2` carries the float2 array, `ldo2` the number of valid columns).  The
  // lm_head GEMM computes them while the accumulator is in registers anyway, which saves the cross-entropy kernel one full
  // pass over the 412 MB of logits.  CTA-pair kernel only.
  EPI_STORE_STATS = 8,
};

struct GemmEpi {
  int mode;
  float alpha;
  const bf16* bias;  // [N] bf16 (compute copy of the parameter) or null
  const bf16* aux;
  long long ld_aux;
  void* out;
  long long ldo;
  bf16* out2;
  long long ldo2;
  int coop_store;             // 1: epilogue warps copy the staged tile to global themselves (no TMA store)
  unsigned long long* trace;  // debug: per-k-block clock64 timestamps of CTA 0 (null in production)
};

// Parameter block of the SCATTER instantiations only (the other kernels keep the plain GemmEpi parameter).
struct GemmEpiScatter : GemmEpi {
  bf16* peer_out[8];     // base of every rank's partial-sum buffer ([n_ranks * sp_rows, ldo] bf16)
  int sp_rank, sp_rows;  // this rank; rows owned per rank
};
// ... plus the folded-barrier signal block (experimental SIGNAL instantiations: the last CTA publishes the scattered tiles
// to every rank itself instead of a barrier kernel after the GEMM)
struct GemmEpiScatterSync : GemmEpiScatter {
  SpSync sync;
};
template <bool SCATTER, bool SIGNAL = false>
struct EpiArgT { using type = GemmEpi; };
template <>
struct EpiArgT<true, false> { using type = GemmEpiScatter; };
template <>
struct EpiArgT<true, true> { using type = GemmEpiScatterSync; };

DTC_DEVICE void scatter_signal(const GemmEpi&, unsigned) {}
DTC_DEVICE void scatter_signal(const GemmEpiScatterSync& epi, unsigned n_ctas) { sp_signal(epi.sync, n_ctas); }

// destination of the output tile whose first row is m0: (base pointer, row shift)
DTC_DEVICE void scatter_dest(const GemmEpi& epi, int, bf16*& base, long long& rshift) {
  base = reinterpret_cast<bf16*>(epi.out);
  rshift = 0;
}
DTC_DEVICE void scatter_dest(const GemmEpiScatter& epi, int m0, bf16*& base, long long& rshift) {
  const int owner = m0 / epi.sp_rows;  // the rows of this tile belong to rank `owner`: my slot of ITS partial-sum buffer
  base = epi.peer_out[0];
#pragma unroll
  for (int i = 1; i < 8; ++i)
    if (i == owner) base = epi.peer_out[i];
  rshift = (long long)(epi.sp_rank - owner) * epi.sp_rows;
}

template <int BN>
struct GemmSmem {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int bytes(int stages) { return stages * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/; }
};

DTC_DEVICE void red_add_v4_f32(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <bool A_MN, bool B_MN, int BN>
__global__ void __launch_bounds__(GEMM_THREADS)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, int M, int N,
                    int K, int m_blocks, int kb_per_split, int num_stages, GemmEpi epi) {
  using S = GemmSmem<BN>;
  extern __shared__ uint8_t smem_raw[];
  // 128B swizzle atoms need a 1024-byte aligned base
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + num_stages * S::STAGE_BYTES);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + MAX_STAGES;
  uint64_t* tmem_full_bar = bars + 2 * MAX_STAGES;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * MAX_STAGES + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m_blk = blockIdx.x % m_blocks;  // m fastest: CTAs that share a B tile are adjacent
  const int n_blk = blockIdx.x / m_blocks;
  const int total_kb = (K + BK - 1) / BK;
  const int kb_begin = blockIdx.y * kb_per_split;
  const int kb_end = min(total_kb, kb_begin + kb_per_split);
  const int nkb = kb_end - kb_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int s = 0; s < num_stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr_smem, BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_launch_dependents();  // PDL: only after this CTA owns its TMEM (a dependent may not allocate first)
  pdl_wait();               // inputs come from the previous kernel: no global access before it has completed

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (elect_one()) {
      for (int i = 0; i < nkb; ++i) {
        const int s = i % num_stages;
        const uint32_t ph = (i / num_stages) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1);
        uint8_t* sa = smem + s * S::STAGE_BYTES;
        uint8_t* sb = sa + S::A_BYTES;
        mbar_expect_tx(&full_bar[s], S::STAGE_BYTES);
        const int k0 = (kb_begin + i) * BK;
        if (A_MN) {
          // global [K, M] (M contiguous): one 64(m) x 64(k) box per 64 rows of the tile
#pragma unroll
          for (int j = 0; j < BM / 64; ++j) tma_load_2d(sa + j * (BK * 128), &tma_a, &full_bar[s], m_blk * BM + j * 64, k0);
        } else {
          tma_load_2d(sa, &tma_a, &full_bar[s], k0, m_blk * BM);
        }
        if (B_MN) {
#pragma unroll
          for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * (BK * 128), &tma_b, &full_bar[s], n_blk * BN + j * 64, k0);
        } else {
          tma_load_2d(sb, &tma_b, &full_bar[s], k0, n_blk * BN);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer ------------------------------
    constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
    // K-major, 128B swizzle: 8-row groups 1024 B apart (SBO); LBO unused (encoded 1).  One UMMA_K=16 step = 32 B.
    // MN-major, 128B swizzle: 64-element column groups BK*128 B apart (LBO), 8-k groups 1024 B apart (SBO);
    // one UMMA_K=16 step = 16 rows of 128 B.
    constexpr uint32_t A_LBO = A_MN ? BK * 128 : 16, A_STEP = A_MN ? 2048 : 32;
    constexpr uint32_t B_LBO = B_MN ? BK * 128 : 16, B_STEP = B_MN ? 2048 : 32;
    for (int i = 0; i < nkb; ++i) {
      const int s = i % num_stages;
      const uint32_t ph = (i / num_stages) & 1;
      mbar_wait(&full_bar[s], ph);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sa = smem_u32(smem + s * S::STAGE_BYTES);
        const uint32_t sb = sa + S::A_BYTES;
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          const uint64_t da = make_smem_desc(sa + k * A_STEP, A_LBO, 1024, SWZ_128B);
          const uint64_t db = make_smem_desc(sb + k * B_STEP, B_LBO, 1024, SWZ_128B);
          umma_bf16(tmem_base, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);                   // frees this smem stage once the MMAs have read it
        if (i == nkb - 1) umma_commit(tmem_full_bar);  // accumulator complete
      }
      __syncwarp();
    }
  } else {
    // ------------------------------ epilogue ------------------------------
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    const int row = q * 32 + lane;
    const long long gm = (long long)m_blk * BM + row;
    const bool row_ok = gm < M;
    if (nkb > 0) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
    const int mode = epi.mode;
    const float alpha = epi.alpha;
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      uint32_t r[32];
      if (nkb > 0) {
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + c * 32, r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = 0;
      }
      const int n0 = n_blk * BN + c * 32;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const int n = n0 + g * 8;
        if (!row_ok || n >= N) continue;  // N % 8 == 0, so an 8-group is all in or all out
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[g * 8 + j]) * alpha;
        if (epi.bias != nullptr && mode != EPI_ATOMIC_F32 && mode != EPI_DGELU) {
          const uint4 bb = *reinterpret_cast<const uint4*>(epi.bias + n);
          const uint32_t bw[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_bf16x2(bw[j]);
            v[2 * j] += f.x;
            v[2 * j + 1] += f.y;
          }
        }
        if (mode == EPI_ATOMIC_F32) {
          float* o = reinterpret_cast<float*>(epi.out) + gm * epi.ldo + n;
          red_add_v4_f32(o, v[0], v[1], v[2], v[3]);
          red_add_v4_f32(o + 4, v[4], v[5], v[6], v[7]);
          continue;
        }
        if (mode == EPI_STORE_F32) {
          float* o = reinterpret_cast<float*>(epi.out) + gm * epi.ldo + n;
          *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
          *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
          continue;
        }
        if (mode == EPI_BIAS_RESIDUAL || mode == EPI_DGELU) {
          const uint4 a = *reinterpret_cast<const uint4*>(epi.aux + gm * epi.ld_aux + n);
          const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_bf16x2(aw[j]);
            if (mode == EPI_BIAS_RESIDUAL) {
              v[2 * j] += f.x;
              v[2 * j + 1] += f.y;
            } else {
              v[2 * j] *= gelu_tanh_grad(f.x);
              v[2 * j + 1] *= gelu_tanh_grad(f.y);
            }
          }
        }
        bf16* o = reinterpret_cast<bf16*>(epi.out) + gm * epi.ldo + n;
        *reinterpret_cast<uint4*>(o) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                                  pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
        if (mode == EPI_BIAS_GELU) {
          bf16* o2 = epi.out2 + gm * epi.ldo2 + n;
          *reinterpret_cast<uint4*>(o2) =
              make_uint4(pack_bf16x2(gelu_tanh(v[0]), gelu_tanh(v[1])), pack_bf16x2(gelu_tanh(v[2]), gelu_tanh(v[3])),
                         pack_bf16x2(gelu_tanh(v[4]), gelu_tanh(v[5])), pack_bf16x2(gelu_tanh(v[6]), gelu_tanh(v[7])));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BN);
  }
}

template <bool A_MN, bool B_MN, int BN>
static int launch_gemm(const void* A, long long lda, const void* B, long long ldb, int M, int N, int K, int splits,
                       int stages, const GemmEpi& epi, cudaStream_t stream) {
  using S = GemmSmem<BN>;
  CUtensorMap ta, tb;
  int rc;
  if (A_MN) rc = make_tmap_2d_bf16(&ta, A, K, M, lda, BK, 64, CU_TENSOR_MAP_SWIZZLE_128B);
  else rc = make_tmap_2d_bf16(&ta, A, M, K, lda, BM, BK, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  if (B_MN) rc = make_tmap_2d_bf16(&tb, B, K, N, ldb, BK, 64, CU_TENSOR_MAP_SWIZZLE_128B);
  else rc = make_tmap_2d_bf16(&tb, B, N, K, ldb, BN, BK, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;

  const int total_kb = (K + BK - 1) / BK;
  if (splits < 1) splits = 1;
  if (splits > total_kb) splits = total_kb > 0 ? total_kb : 1;
  const int kb_per_split = (total_kb + splits - 1) / splits;
  splits = kb_per_split > 0 ? (total_kb + kb_per_split - 1) / kb_per_split : 1;
  if (splits < 1) splits = 1;
  if (stages < 2) stages = 2;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  while (S::bytes(stages) > 227 * 1024) --stages;
  const int smem_bytes = S::bytes(stages);
  auto kern = gemm_tcgen05_kernel<A_MN, B_MN, BN>;
  static dtc::SmemOptIn configured;  // per instantiation, per device
  DTC_CHECK_CUDA(dtc::ensure_dyn_smem(kern, 227 * 1024, configured));
  const int m_blocks = (M + BM - 1) / BM, n_blocks = (N + BN - 1) / BN;
  dim3 grid(m_blocks * n_blocks, splits, 1);
  DTC_CHECK_CUDA(dtc::launch(kern, grid, GEMM_THREADS, smem_bytes, stream, ta, tb, M, N, K, m_blocks, kb_per_split, stages, epi));
  DTC_CHECK_LAUNCH();
  return 0;
}


// =========================================================================================================
// Persistent variant (default): one CTA per SM walks a static round-robin tile list.
//   * the accumulator is double-buffered in TMEM (2 x BN columns): the MMA warp starts tile i+1 while the
//     epilogue warps drain tile i;
//   * bf16 outputs leave through a 128B-swizzled smem staging tile and TMA stores (coalesced, clipped at the
//     matrix edge by the tensor map); the residual / pre-activation tile needed by the epilogue is TMA-loaded
//     into a second staging tile by the producer warp while the mainloop runs;
//   * fp32 outputs (split-K reduce-add, fp32 store) keep the register path.
// =========================================================================================================
constexpr int PG_STAGE_MAX = 6;

template <int BN>
struct PgSmem {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int HALF_BYTES = BM * 128;                   // 128 rows x 64 bf16
  static constexpr int TILE_BYTES = (BN / 64) * HALF_BYTES;     // staging tile [128 x BN] bf16
  // staging tiles: 2 (out + aux/out2), 1 in the DUAL variant, none for the fp32 register-path epilogues of DUAL
  static constexpr int bytes(int stages, bool dual = false, bool staging = true) {
    return stages * STAGE_BYTES + (dual ? (staging ? 1 : 0) : 2) * TILE_BYTES + 1024 + 256 + 1024 /* bias tiles */;
  }
};

constexpr int PG_THREADS = 64 + 256;  // TMA warp, MMA warp, 8 epilogue warps (two per TMEM lane quadrant)
DTC_DEVICE void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// DUAL = two CTAs per SM.  Clock traces (tools/gemm_trace.py) show that ONE CTA gets its TMA tile loads served at a
// fixed ~48 B/cycle (one 32 KB stage per 676 cycles, against 256 cycles of MMA work), whatever the number of stages
// or issuing warps, while two co-resident CTAs are each served at that rate.  The DUAL variant therefore keeps the
// CTA under 113 KB of shared memory (2-3 stages, ONE staging tile: the residual / pre-activation tile is TMA-loaded
// into the output staging tile by the epilogue leader and overwritten in place) and 256 TMEM columns.
template <bool A_MN, bool B_MN, int BN, bool DUAL, bool SCATTER = false, bool SIGNAL = false>
__global__ void __launch_bounds__(PG_THREADS, DUAL ? 2 : 1)
gemm_persistent_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                       const __grid_constant__ CUtensorMap tma_out, const __grid_constant__ CUtensorMap tma_out2,
                       const __grid_constant__ CUtensorMap tma_aux, int M, int N, int K, int m_blocks, int n_blocks,
                       int splits, int kb_per_split, int num_stages, const typename EpiArgT<SCATTER, SIGNAL>::type epi) {
  using S = PgSmem<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* stage_out = smem + num_stages * S::STAGE_BYTES;  // out tile (bf16)
  uint8_t* stage_aux = DUAL ? stage_out : stage_out + S::TILE_BYTES;  // aux tile (input) or out2 tile (GELU output)
  uint64_t* bars =
      reinterpret_cast<uint64_t*>(stage_out + (DUAL ? ((epi.mode <= EPI_DGELU || SCATTER) ? 1 : 0) : 2) * S::TILE_BYTES);
  uint64_t* full_bar = bars;                       // [PG_STAGE_MAX]
  uint64_t* empty_bar = bars + PG_STAGE_MAX;       // [PG_STAGE_MAX]
  uint64_t* tmem_full = bars + 2 * PG_STAGE_MAX;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;            // [2]
  uint64_t* aux_full = tmem_empty + 2;
  uint64_t* aux_empty = aux_full + 1;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(aux_empty + 1);
  bf16* s_bias = reinterpret_cast<bf16*>(bars + 32);  // [2][BN] bias slice of the current / next tile (256 B past the barriers)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_kb = (K + BK - 1) / BK;
  const int tiles_mn = m_blocks * n_blocks;
  const int total_tiles = tiles_mn * splits;
  const int mode = epi.mode;
  const bool bf16_out = mode <= EPI_DGELU || SCATTER;  // outputs that leave through the smem staging tile
  const bool use_aux = (mode == EPI_BIAS_RESIDUAL || mode == EPI_DGELU);
  const bool coop_store = SCATTER || epi.coop_store;   // staged tile written out by the epilogue warps (no TMA store)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    if (bf16_out && !SCATTER) tma_prefetch_desc(&tma_out);
    for (int s = 0; s < num_stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full[b], 1);
      mbar_init(&tmem_empty[b], 8);  // one arrival per epilogue warp
    }
    mbar_init(aux_full, 1);
    mbar_init(aux_empty, 8);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr_smem, 2 * BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_launch_dependents();  // PDL: only after this CTA owns its TMEM (a dependent may not allocate first)
  pdl_wait();               // inputs come from the previous kernel: no global access before it has completed

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (elect_one()) {
      uint32_t kit = 0;  // running k-block counter across tiles -> stage / phase
      int it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int split = tile / tiles_mn, mn = tile % tiles_mn;
        const int m_blk = mn % m_blocks, n_blk = mn / m_blocks;
        const int kb_begin = split * kb_per_split;
        const int nkb = min(total_kb, kb_begin + kb_per_split) - kb_begin;
        if (use_aux && !DUAL) {  // stage the residual / pre-activation tile for this output tile
          mbar_wait(aux_empty, (it & 1) ^ 1);
          mbar_expect_tx(aux_full, S::TILE_BYTES);
#pragma unroll
          for (int hf = 0; hf < BN / 64; ++hf)
            tma_load_2d(stage_aux + hf * S::HALF_BYTES, &tma_aux, aux_full, n_blk * BN + hf * 64, m_blk * BM);
        }
        for (int i = 0; i < nkb; ++i, ++kit) {
          const int s = kit % num_stages;
          const uint32_t ph = (kit / num_stages) & 1;
          mbar_wait(&empty_bar[s], ph ^ 1);
          if (epi.trace && blockIdx.x == 0 && kit < 256) epi.trace[kit * 4 + 0] = clock64();
          uint8_t* sa = smem + s * S::STAGE_BYTES;
          uint8_t* sb = sa + S::A_BYTES;
          mbar_expect_tx(&full_bar[s], S::STAGE_BYTES);
          const int k0 = (kb_begin + i) * BK;
          if (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) tma_load_2d(sa + j * (BK * 128), &tma_a, &full_bar[s], m_blk * BM + j * 64, k0);
          } else {
            tma_load_2d(sa, &tma_a, &full_bar[s], k0, m_blk * BM);
          }
          if (B_MN) {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * (BK * 128), &tma_b, &full_bar[s], n_blk * BN + j * 64, k0);
          } else {
            tma_load_2d(sb, &tma_b, &full_bar[s], k0, n_blk * BN);
          }
          if (epi.trace && blockIdx.x == 0 && kit < 256) epi.trace[kit * 4 + 1] = clock64();
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer ------------------------------
    constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
    constexpr uint32_t A_LBO = A_MN ? BK * 128 : 16, A_STEP = A_MN ? 2048 : 32;
    constexpr uint32_t B_LBO = B_MN ? BK * 128 : 16, B_STEP = B_MN ? 2048 : 32;
    uint32_t kit = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int split = tile / tiles_mn;
      const int kb_begin = split * kb_per_split;
      const int nkb = min(total_kb, kb_begin + kb_per_split) - kb_begin;
      const int buf = it & 1;
      mbar_wait(&tmem_empty[buf], ((it >> 1) & 1) ^ 1);  // epilogue has drained this accumulator buffer
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + buf * BN;
      for (int i = 0; i < nkb; ++i, ++kit) {
        const int s = kit % num_stages;
        const uint32_t ph = (kit / num_stages) & 1;
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        if (elect_one()) {
          if (epi.trace && blockIdx.x == 0 && kit < 256) epi.trace[kit * 4 + 2] = clock64();
          const uint32_t sa = smem_u32(smem + s * S::STAGE_BYTES);
          const uint32_t sb = sa + S::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = make_smem_desc(sa + k * A_STEP, A_LBO, 1024, SWZ_128B);
            const uint64_t db = make_smem_desc(sb + k * B_STEP, B_LBO, 1024, SWZ_128B);
            umma_bf16(tmem_d, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);
          if (i == nkb - 1) umma_commit(&tmem_full[buf]);
          if (epi.trace && blockIdx.x == 0 && kit < 256) epi.trace[kit * 4 + 3] = clock64();
        }
        __syncwarp();
      }
    }
  } else {
    // ------------------------------ epilogue (warps 2..9, 256 threads) ------------------------------
    // warp w reads TMEM lane quadrant w % 4; the two warps of a quadrant split the BN columns in halves
    const int q = warp & 3;
    const int chalf = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const bool store_leader = (threadIdx.x == 64);  // first epilogue thread issues the TMA stores
    const float alpha = epi.alpha;
    int it = 0;
    if (DUAL && use_aux && store_leader && blockIdx.x < total_tiles) {  // aux tile of the first output tile
      const int mn = blockIdx.x % tiles_mn;
      mbar_expect_tx(aux_full, S::TILE_BYTES);
#pragma unroll
      for (int hf = 0; hf < BN / 64; ++hf)
        tma_load_2d(stage_aux + hf * S::HALF_BYTES, &tma_aux, aux_full, (mn / m_blocks) * BN + hf * 64, (mn % m_blocks) * BM);
    }
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int mn = tile % tiles_mn;
      const int m_blk = mn % m_blocks, n_blk = mn / m_blocks;
      const int buf = it & 1;
      const long long gm = (long long)m_blk * BM + row;
      const bool row_ok = gm < M;
      if (epi.trace && blockIdx.x == 0 && threadIdx.x == 64 && it < 32) epi.trace[1024 + it * 4 + 0] = clock64();
      // bias slice of this tile -> smem while the mainloop is still running (a global load per 8-column group inside
      // the accumulator loop costs ~400 cycles each, serialised: 3500 cycles per tile in the clock traces)
      const bool has_bias = epi.bias != nullptr && mode != EPI_ATOMIC_F32 && mode != EPI_DGELU;
      if (has_bias) {
        const int et = threadIdx.x - 64;
        if (et < BN / 8) {
          const int n = n_blk * BN + et * 8;
          uint4 bv = make_uint4(0, 0, 0, 0);
          if (n < N) bv = *reinterpret_cast<const uint4*>(epi.bias + n);
          *reinterpret_cast<uint4*>(s_bias + buf * BN + et * 8) = bv;
        }
      }
      mbar_wait(&tmem_full[buf], (it >> 1) & 1);
      tc_fence_after();
      if (has_bias) epi_bar_sync();
      if (epi.trace && blockIdx.x == 0 && threadIdx.x == 64 && it < 32) epi.trace[1024 + it * 4 + 1] = clock64();
      if (use_aux) mbar_wait(aux_full, it & 1);
      if (bf16_out && !(DUAL && use_aux) && !coop_store) {
        // the previous tile's TMA stores must have finished READING the staging tiles before they are rewritten
        // (DUAL with an aux tile: the leader already waited for that before it fetched this tile's aux tile)
        if (store_leader) tma_store_wait_read();
        epi_bar_sync();
      }
      const uint32_t tmem_d = tmem_base + buf * BN + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c = chalf * (BN / 64); c < (chalf + 1) * (BN / 64); ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem_d + c * 32, r);
        tmem_ld_wait();
        const int n0 = n_blk * BN + c * 32;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int n = n0 + g * 8;
          float v[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[g * 8 + j]) * alpha;
          if (has_bias) {
            const uint4 bb = *reinterpret_cast<const uint4*>(s_bias + buf * BN + c * 32 + g * 8);
            const uint32_t bw[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float2 f = unpack_bf16x2(bw[j]);
              v[2 * j] += f.x;
              v[2 * j + 1] += f.y;
            }
          }
          if (!bf16_out) {
            if (!row_ok || n >= N) continue;
            float* o = reinterpret_cast<float*>(epi.out) + gm * epi.ldo + n;
            if (mode == EPI_ATOMIC_F32) {
              red_add_v4_f32(o, v[0], v[1], v[2], v[3]);
              red_add_v4_f32(o + 4, v[4], v[5], v[6], v[7]);
            } else {
              *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
              *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
            }
            continue;
          }
          // staging offset of this 16-byte chunk: 64-column half, 128-byte rows, 128B swizzle (matches the TMA maps)
          const uint32_t off = (uint32_t)(c >> 1) * S::HALF_BYTES + swizzle_offset<3>(row * 128 + ((c & 1) * 4 + g) * 16);
          if (use_aux) {
            const uint4 a = *reinterpret_cast<const uint4*>(stage_aux + off);
            const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float2 f = unpack_bf16x2(aw[j]);
              if (mode == EPI_BIAS_RESIDUAL) {
                v[2 * j] += f.x;
                v[2 * j + 1] += f.y;
              } else {
                v[2 * j] *= gelu_tanh_grad(f.x);
                v[2 * j + 1] *= gelu_tanh_grad(f.y);
              }
            }
          }
          *reinterpret_cast<uint4*>(stage_out + off) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                                                 pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
          if (mode == EPI_BIAS_GELU) {
            *reinterpret_cast<uint4*>(stage_aux + off) = make_uint4(
                pack_bf16x2(gelu_tanh(v[0]), gelu_tanh(v[1])), pack_bf16x2(gelu_tanh(v[2]), gelu_tanh(v[3])),
                pack_bf16x2(gelu_tanh(v[4]), gelu_tanh(v[5])), pack_bf16x2(gelu_tanh(v[6]), gelu_tanh(v[7])));
          }
        }
      }
      // accumulator buffer (and the aux tile) may be reused by the MMA / producer warps
      if (epi.trace && blockIdx.x == 0 && threadIdx.x == 64 && it < 32) epi.trace[1024 + it * 4 + 2] = clock64();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&tmem_empty[buf]);
        if (use_aux && !DUAL) mbar_arrive(aux_empty);
      }
      if (bf16_out && coop_store) {
        // The TMA queue of a CTA is served in order at ~48 B/cycle and is already saturated by the operand loads: the
        // epilogue warps therefore write the staged tile out themselves (8 threads per 128-byte row: coalesced).
        epi_bar_sync();
        const int et = threadIdx.x - 64, ch = et & 7, r0 = et >> 3;
        bf16* obase;
        long long rshift;  // destination row = output row + rshift
        scatter_dest(epi, m_blk * BM, obase, rshift);
#pragma unroll
        for (int hf = 0; hf < BN / 64; ++hf) {
          const int ncol = n_blk * BN + hf * 64 + ch * 8;
          if (ncol < N) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int r = r0 + 32 * i;
              const long long grow = (long long)m_blk * BM + r;
              if (grow < M) {
                const uint32_t so = (uint32_t)hf * S::HALF_BYTES + swizzle_offset<3>(r * 128 + ch * 16);
                *reinterpret_cast<uint4*>(obase + (grow + rshift) * epi.ldo + ncol) =
                    *reinterpret_cast<const uint4*>(stage_out + so);
                if (mode == EPI_BIAS_GELU)
                  *reinterpret_cast<uint4*>(epi.out2 + grow * epi.ldo2 + ncol) = *reinterpret_cast<const uint4*>(stage_aux + so);
              }
            }
          }
        }
        epi_bar_sync();  // staging tile(s) free again
        if (epi.trace && blockIdx.x == 0 && threadIdx.x == 64 && it < 32) epi.trace[1024 + it * 4 + 3] = clock64();
        if (DUAL && use_aux && store_leader && tile + (int)gridDim.x < total_tiles) {
          const int nmn = (tile + (int)gridDim.x) % tiles_mn;
          fence_proxy_async_smem();
          mbar_expect_tx(aux_full, S::TILE_BYTES);
#pragma unroll
          for (int hf = 0; hf < BN / 64; ++hf)
            tma_load_2d(stage_aux + hf * S::HALF_BYTES, &tma_aux, aux_full, (nmn / m_blocks) * BN + hf * 64,
                        (nmn % m_blocks) * BM);
        }
      } else if (bf16_out) {
        fence_proxy_async_smem();
        epi_bar_sync();
        if (store_leader) {
#pragma unroll
          for (int hf = 0; hf < BN / 64; ++hf) {
            tma_store_2d(&tma_out, stage_out + hf * S::HALF_BYTES, n_blk * BN + hf * 64, m_blk * BM);
            if (mode == EPI_BIAS_GELU)
              tma_store_2d(&tma_out2, stage_aux + hf * S::HALF_BYTES, n_blk * BN + hf * 64, m_blk * BM);
          }
          tma_store_commit();
          if (DUAL && use_aux && tile + (int)gridDim.x < total_tiles) {
            // one staging tile: once the store has read it, fetch the NEXT tile's aux tile into it
            tma_store_wait_read();
            const int nmn = (tile + (int)gridDim.x) % tiles_mn;
            mbar_expect_tx(aux_full, S::TILE_BYTES);
#pragma unroll
            for (int hf = 0; hf < BN / 64; ++hf)
              tma_load_2d(stage_aux + hf * S::HALF_BYTES, &tma_aux, aux_full, (nmn / m_blocks) * BN + hf * 64,
                          (nmn % m_blocks) * BM);
          }
        }
      }
    }
    if (bf16_out && store_leader) tma_store_wait_all();
    if (SIGNAL) {  // every tile of this CTA has been stored to its owner: fence, then the last CTA of the grid signals all ranks
      __threadfence_system();
      epi_bar_sync();
      if (store_leader) scatter_signal(epi, gridDim.x);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

static int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        n <= 0)
      n = 148;
  }
  return n;
}

template <bool A_MN, bool B_MN, int BN, bool DUAL, bool SCATTER = false, bool SIGNAL = false>
static int launch_gemm_persistent(const void* A, long long lda, const void* B, long long ldb, int M, int N, int K,
                                  int splits, int stages, const typename EpiArgT<SCATTER, SIGNAL>::type& epi,
                                  cudaStream_t stream) {
  using S = PgSmem<BN>;
  CUtensorMap ta, tb, tout, tout2, taux;
  int rc;
  if (A_MN) rc = make_tmap_2d_bf16(&ta, A, K, M, lda, BK, 64, CU_TENSOR_MAP_SWIZZLE_128B);
  else rc = make_tmap_2d_bf16(&ta, A, M, K, lda, BM, BK, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  if (B_MN) rc = make_tmap_2d_bf16(&tb, B, K, N, ldb, BK, 64, CU_TENSOR_MAP_SWIZZLE_128B);
  else rc = make_tmap_2d_bf16(&tb, B, N, K, ldb, BN, BK, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  tout = ta;
  tout2 = ta;
  taux = ta;  // placeholders: never dereferenced when the mode does not use them
  if (epi.mode <= EPI_DGELU) {
    if ((rc = make_tmap_2d_bf16(&tout, epi.out, M, N, epi.ldo, BM, 64, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    if (epi.mode == EPI_BIAS_GELU &&
        (rc = make_tmap_2d_bf16(&tout2, epi.out2, M, N, epi.ldo2, BM, 64, CU_TENSOR_MAP_SWIZZLE_128B)))
      return rc;
    if ((epi.mode == EPI_BIAS_RESIDUAL || epi.mode == EPI_DGELU) &&
        (rc = make_tmap_2d_bf16(&taux, epi.aux, M, N, epi.ld_aux, BM, 64, CU_TENSOR_MAP_SWIZZLE_128B)))
      return rc;
  }
  const int total_kb = (K + BK - 1) / BK;
  if (splits < 1) splits = 1;
  if (splits > total_kb) splits = total_kb;
  const int kb_per_split = (total_kb + splits - 1) / splits;
  splits = (total_kb + kb_per_split - 1) / kb_per_split;
  if (stages < 2) stages = 2;
  if (stages > PG_STAGE_MAX) stages = PG_STAGE_MAX;
  const int smem_cap = DUAL ? 112 * 1024 : 227 * 1024;  // DUAL: two CTAs (+1 KB reserved each) per 228 KB SM
  const bool staging = epi.mode <= EPI_DGELU || SCATTER;
  while (stages > 2 && S::bytes(stages, DUAL, staging) > smem_cap) --stages;
  const int smem_bytes = S::bytes(stages, DUAL, staging);
  auto kern = gemm_persistent_kernel<A_MN, B_MN, BN, DUAL, SCATTER, SIGNAL>;
  static dtc::SmemOptIn configured;  // per device
  DTC_CHECK_CUDA(dtc::ensure_dyn_smem(kern, DUAL ? 113 * 1024 : 227 * 1024, configured));
  const int m_blocks = (M + BM - 1) / BM, n_blocks = (N + BN - 1) / BN;
  const int total_tiles = m_blocks * n_blocks * splits;
  const int max_ctas = (DUAL ? 2 : 1) * num_sms();
  const int grid = total_tiles < max_ctas ? total_tiles : max_ctas;
  DTC_CHECK_CUDA(dtc::launch(kern, grid, PG_THREADS, smem_bytes, stream, ta, tb, tout, tout2, taux, M, N, K, m_blocks, n_blocks, splits,
                                                   kb_per_split, stages, epi));
  DTC_CHECK_LAUNCH();
  return 0;
}


// =========================================================================================================
// Grouped weight-gradient GEMM: up to GG_MAX problems dW_p[M_p, N_p] += X_pᵀ · dY_p that share the reduction length K
// (the tokens of the microbatch) run as ONE persistent launch — the four weight gradients of a transformer layer
// (q/k/v, out_proj, fc1, fc2).  Against four split-K launches: one prologue / tail instead of four, no split-K (every
// tile runs the whole K = B·T mainloop, so the fp32 reduce-add traffic drops by the split factor) and 192 tiles in one
// wave-packed grid instead of 4 × (16 … 64).  Same pipeline as gemm_persistent_kernel<true, true, BN, DUAL> (both
// operands MN-major: X is [K, M_p], dY is [K, N_p] row-major), register-path fp32 reduce-add epilogue, 2 CTAs per SM.
// The tensor maps travel in the parameter block (__grid_constant__), indexed by the problem a tile belongs to.
// =========================================================================================================
constexpr int GG_MAX = 4;

struct GemmGroup {
  CUtensorMap a[GG_MAX];
  CUtensorMap b[GG_MAX];
  CUtensorMap o[GG_MAX];  // fp32 outputs in boxes of [128 rows x 32 columns] (CTA-pair kernel: TMA reduce-add epilogue)
  float* out[GG_MAX];
  long long ldo[GG_MAX];
  int M[GG_MAX];
  int N[GG_MAX];
  int m_blocks[GG_MAX];
  int tile_start[GG_MAX + 1];  // prefix sums of m_blocks * n_blocks
  int count;
};

DTC_DEVICE void group_decode(const GemmGroup& g, int t, int& p, int& m_blk, int& n_blk) {
  p = 0;
#pragma unroll
  for (int i = 1; i < GG_MAX; ++i)
    if (i < g.count && t >= g.tile_start[i]) p = i;
  const int local = t - g.tile_start[p];
  m_blk = local % g.m_blocks[p];
  n_blk = local / g.m_blocks[p];
}

template <int BN>
__global__ void __launch_bounds__(PG_THREADS, 2)
gemm_grouped_wgrad_kernel(const __grid_constant__ GemmGroup grp, int K, int splits, int kb_per_split, int num_stages,
                          float alpha) {
  using S = PgSmem<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + num_stages * S::STAGE_BYTES);
  uint64_t* full_bar = bars;                       // [PG_STAGE_MAX]
  uint64_t* empty_bar = bars + PG_STAGE_MAX;       // [PG_STAGE_MAX]
  uint64_t* tmem_full = bars + 2 * PG_STAGE_MAX;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;            // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_kb = (K + BK - 1) / BK;
  const int tiles_mn = grp.tile_start[grp.count];
  const int total_tiles = tiles_mn * splits;

  if (warp == 0 && lane == 0) {
    for (int p = 0; p < grp.count; ++p) {
      tma_prefetch_desc(&grp.a[p]);
      tma_prefetch_desc(&grp.b[p]);
    }
    for (int s = 0; s < num_stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full[b], 1);
      mbar_init(&tmem_empty[b], 8);  // one arrival per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr_smem, 2 * BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_launch_dependents();
  pdl_wait();

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (elect_one()) {
      uint32_t kit = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int split = tile / tiles_mn;
        int p, m_blk, n_blk;
        group_decode(grp, tile % tiles_mn, p, m_blk, n_blk);
        const CUtensorMap* ta = &grp.a[p];
        const CUtensorMap* tb = &grp.b[p];
        const int kb_begin = split * kb_per_split;
        const int nkb = min(total_kb, kb_begin + kb_per_split) - kb_begin;
        for (int i = 0; i < nkb; ++i, ++kit) {
          const int s = kit % num_stages;
          const uint32_t ph = (kit / num_stages) & 1;
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* sa = smem + s * S::STAGE_BYTES;
          uint8_t* sb = sa + S::A_BYTES;
          mbar_expect_tx(&full_bar[s], S::STAGE_BYTES);
          const int k0 = (kb_begin + i) * BK;
#pragma unroll
          for (int j = 0; j < BM / 64; ++j) tma_load_2d(sa + j * (BK * 128), ta, &full_bar[s], m_blk * BM + j * 64, k0);
#pragma unroll
          for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * (BK * 128), tb, &full_bar[s], n_blk * BN + j * 64, k0);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer ------------------------------
    constexpr uint32_t idesc = make_idesc_bf16(BM, BN, 1, 1);
    constexpr uint32_t LBO = BK * 128, STEP = 2048;  // MN-major operands (see gemm_persistent_kernel)
    uint32_t kit = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int split = tile / tiles_mn;
      const int kb_begin = split * kb_per_split;
      const int nkb = min(total_kb, kb_begin + kb_per_split) - kb_begin;
      const int buf = it & 1;
      mbar_wait(&tmem_empty[buf], ((it >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + buf * BN;
      for (int i = 0; i < nkb; ++i, ++kit) {
        const int s = kit % num_stages;
        const uint32_t ph = (kit / num_stages) & 1;
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + s * S::STAGE_BYTES);
          const uint32_t sb = sa + S::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = make_smem_desc(sa + k * STEP, LBO, 1024, SWZ_128B);
            const uint64_t db = make_smem_desc(sb + k * STEP, LBO, 1024, SWZ_128B);
            umma_bf16(tmem_d, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);
          if (i == nkb - 1) umma_commit(&tmem_full[buf]);
        }
        __syncwarp();
      }
    }
  } else {
    // ------------------------------ epilogue (warps 2..9): fp32 reduce-add of the accumulator ------------------------------
    const int q = warp & 3;
    const int chalf = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      int p, m_blk, n_blk;
      group_decode(grp, tile % tiles_mn, p, m_blk, n_blk);
      const int buf = it & 1;
      const long long gm = (long long)m_blk * BM + row;
      const bool row_ok = gm < grp.M[p];
      const int Np = grp.N[p];
      float* obase = grp.out[p] + gm * grp.ldo[p];
      mbar_wait(&tmem_full[buf], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + buf * BN + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c = chalf * (BN / 64); c < (chalf + 1) * (BN / 64); ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem_d + c * 32, r);
        tmem_ld_wait();
        const int n0 = n_blk * BN + c * 32;
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          const int n = n0 + g * 4;
          if (row_ok && n < Np)
            red_add_v4_f32(obase + n, __uint_as_float(r[g * 4]) * alpha, __uint_as_float(r[g * 4 + 1]) * alpha,
                           __uint_as_float(r[g * 4 + 2]) * alpha, __uint_as_float(r[g * 4 + 3]) * alpha);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[buf]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * BN);
  }
}

// =========================================================================================================
// Grouped weight gradients on CTA pairs: the same problem list as gemm_grouped_wgrad_kernel, but every work item is a
// 256 x 256 output tile computed by two CTAs (tcgen05.mma.cta_group::2, UMMA M = 256) and a K slice (split-K).  Per
// k-block a CTA loads 128 x 64 of Xᵀ and 128 x 64 of dY (32 KB) for 2 x 128 x 256 x 64 FLOP — twice the arithmetic
// intensity of the 128 x 128 tiles, which is what the 48 B/clk L2 -> SM port of an SM asks for (the 1-CTA grouped kernel
// runs two CTAs per SM that share that port: 39 us per layer, profiles/r2/step_budget_ncu.md).  Both operands are
// MN-major (X is [K, M_p], dY is [K, N_p] row-major); fp32 reduce-add epilogue straight from registers; six 32 KB stages.
// =========================================================================================================
// epilogue helpers of the CTA-pair kernels: bounded wait on outstanding TMA stores / reduces, 128-thread group barrier
template <int N>
DTC_DEVICE void tma_store_wait_read_n() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
DTC_DEVICE void epi_group_sync(int g) { asm volatile("bar.sync %0, 128;" ::"r"(g + 1) : "memory"); }

constexpr int GW_SLAB_BYTES = BM * 128;  // [128 rows x 32 fp32], one TMA box

template <int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(PG_THREADS, 1)
gemm_grouped_wgrad_2cta_kernel(const __grid_constant__ GemmGroup grp, int K, int splits, int kb_per_split, int num_stages,
                               float alpha) {
  constexpr int A_BYTES = BM * BK * 2, B_BYTES = (BN / 2) * BK * 2, STAGE_BYTES = A_BYTES + B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* slabs = smem + num_stages * STAGE_BYTES;  // 4 x [128 rows x 32 fp32] (two per epilogue group)
  uint64_t* bars = reinterpret_cast<uint64_t*>(slabs + 4 * GW_SLAB_BYTES);
  uint64_t* full_bar = bars;                       // [PG_STAGE_MAX]  (leader's copy counts both CTAs' bytes)
  uint64_t* empty_bar = bars + PG_STAGE_MAX;       // [PG_STAGE_MAX]
  uint64_t* tmem_full = bars + 2 * PG_STAGE_MAX;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;            // [2]  (leader's copy: 16 arrivals)
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int total_kb = (K + BK - 1) / BK;
  const int tiles_mn = grp.tile_start[grp.count];
  const int total_tiles = tiles_mn * splits;

  if (warp == 0 && lane == 0) {
    for (int p = 0; p < grp.count; ++p) {
      tma_prefetch_desc(&grp.a[p]);
      tma_prefetch_desc(&grp.b[p]);
      tma_prefetch_desc(&grp.o[p]);
    }
    for (int s = 0; s < num_stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full[b], 1);
      mbar_init(&tmem_empty[b], 16);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc_2sm(tmem_ptr_smem, 2 * BN);
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_relaxed();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_launch_dependents();
  pdl_wait();

  if (warp == 0) {
    // ------------------------------ TMA producer (both CTAs) ------------------------------
    if (elect_one()) {
      uint32_t kit = 0;
      for (int tile = pair; tile < total_tiles; tile += num_pairs) {
        const int split = tile / tiles_mn;
        int p, m_blk, n_blk;
        group_decode(grp, tile % tiles_mn, p, m_blk, n_blk);
        const CUtensorMap* ta = &grp.a[p];
        const CUtensorMap* tb = &grp.b[p];
        const int row0 = m_blk * 256 + (int)cta * BM;        // this CTA's 128 rows of the output (columns of X)
        const int bcol0 = n_blk * BN + (int)cta * (BN / 2);  // this CTA's half of the dY columns
        const int kb_begin = split * kb_per_split;
        const int nkb = min(total_kb, kb_begin + kb_per_split) - kb_begin;
        for (int i = 0; i < nkb; ++i, ++kit) {
          const int s = kit % num_stages;
          const uint32_t ph = (kit / num_stages) & 1;
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* sa = smem + s * STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          const uint32_t full_leader = mapa_shared(smem_u32(&full_bar[s]), 0);
          if (cta == 0) mbar_expect_tx(&full_bar[s], 2 * STAGE_BYTES);
          const int k0 = (kb_begin + i) * BK;
#pragma unroll
          for (int j = 0; j < BM / 64; ++j) tma_load_2d_2sm(sa + j * (BK * 128), ta, full_leader, row0 + j * 64, k0);
#pragma unroll
          for (int j = 0; j < BN / 128; ++j) tma_load_2d_2sm(sb + j * (BK * 128), tb, full_leader, bcol0 + j * 64, k0);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer (leader CTA only) ------------------------------
    if (cta == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(256, BN, 1, 1);
      uint32_t kit = 0;
      int it = 0;
      for (int tile = pair; tile < total_tiles; tile += num_pairs, ++it) {
        const int split = tile / tiles_mn;
        const int kb_begin = split * kb_per_split;
        const int nkb = min(total_kb, kb_begin + kb_per_split) - kb_begin;
        const int buf = it & 1;
        mbar_wait(&tmem_empty[buf], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * BN;
        for (int i = 0; i < nkb; ++i, ++kit) {
          const int s = kit % num_stages;
          const uint32_t ph = (kit / num_stages) & 1;
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t sa = smem_u32(smem + s * STAGE_BYTES);
            const uint32_t sb = sa + A_BYTES;
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = make_smem_desc(sa + k * 2048, BK * 128, 1024, SWZ_128B);
              const uint64_t db = make_smem_desc(sb + k * 2048, BK * 128, 1024, SWZ_128B);
              umma_bf16_2sm(tmem_d, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
            }
            umma_commit_2sm(&empty_bar[s], 0x3);
            if (i == nkb - 1) umma_commit_2sm(&tmem_full[buf], 0x3);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ------------------------------ epilogue (warps 2..9 of both CTAs): fp32 reduce-add ------------------------------
    // Two groups of four warps (one warp per TMEM lane quadrant) take the two column halves of the tile in slabs of
    // [128 rows x 32 fp32] = one 128B-swizzled TMA box, added to the gradient in L2 by TMA reduce: one bulk request per slab
    // where each thread used to issue eight 16-byte red.global.add into eight different 128-byte lines.
    const int q = warp & 3;
    const int g = (warp - 2) >> 2;      // column half of the tile
    const int row = q * 32 + lane;
    const int gcol = g * (BN / 2);
    const bool leader = (int)threadIdx.x - 64 - g * 128 == 0;
    uint8_t* gslab = slabs + g * 2 * GW_SLAB_BYTES;
    uint32_t jc = 0;
    int it = 0;
    for (int tile = pair; tile < total_tiles; tile += num_pairs, ++it) {
      int p, m_blk, n_blk;
      group_decode(grp, tile % tiles_mn, p, m_blk, n_blk);
      const int buf = it & 1;
      const int row0 = m_blk * 256 + (int)cta * BM;
      mbar_wait(&tmem_full[buf], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + buf * BN + ((uint32_t)(q * 32) << 16) + gcol;
#pragma unroll 1
      for (int c = 0; c < BN / 64; ++c, ++jc) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem_d + c * 32, r);
        tmem_ld_wait();
        if (c == BN / 64 - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty[buf]), 0));
        }
        uint8_t* sbuf = gslab + (jc & 1) * GW_SLAB_BYTES;
        const uint32_t sbuf_u32 = smem_u32(sbuf);
        if (leader) tma_store_wait_read_n<1>();  // the reduce that last used this buffer has read it
        epi_group_sync(g);
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4)
          sts_v4(sbuf_u32 + swizzle_offset<3>(row * 128 + c4 * 16), __float_as_uint(__uint_as_float(r[c4 * 4]) * alpha),
                 __float_as_uint(__uint_as_float(r[c4 * 4 + 1]) * alpha), __float_as_uint(__uint_as_float(r[c4 * 4 + 2]) * alpha),
                 __float_as_uint(__uint_as_float(r[c4 * 4 + 3]) * alpha));
        fence_proxy_async_smem();
        epi_group_sync(g);
        if (leader) {
          tma_reduce_add_2d(&grp.o[p], sbuf, n_blk * BN + gcol + c * 32, row0);  // rows / columns past M_p / N_p are clipped
          tma_store_commit();
        }
      }
    }
    if (leader) tma_store_wait_read();
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_relaxed();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, 2 * BN);
  }
}

// =========================================================================================================
// CTA-pair variant (cta_group::2): two CTAs on the two SMs of a TPC share one 256 x BN output tile.
// Each CTA TMA-loads its own 128 rows of A and HALF of the B tile; the leader CTA's elected thread issues
// tcgen05.mma.cta_group::2 (UMMA M = 256), which reads both CTAs' shared memory and writes 128 accumulator rows
// into each CTA's tensor memory.  Per CTA and per MMA cycle only half the B bytes cross L2 -> SM and shared
// memory, which is what lets a 4..6-stage ring cover the L2 latency at tensor-core rate (the 1-CTA kernel is
// bound by bytes-in-flight: 34 % tensor-pipe activity on the lm_head GEMM, profiles/r1/ncu_gemm_persistent.md).
//   full[s]       lives in the leader; both CTAs' TMA loads complete_tx on it (expect_tx = bytes of both CTAs)
//   empty[s]      one per CTA; tcgen05.commit multicasts the arrival to both
//   tmem_full[b]  one per CTA (commit multicast); tmem_empty[b] lives in the leader, 16 arrivals (8 epilogue warps x 2 CTAs)
// Epilogue: two groups of four warps drain the two column halves in 64-column slab jobs (see P2Smem).
// =========================================================================================================
// Staging of the pair kernel's epilogue: NSLAB slab buffers of [128 rows x 64 bf16] (16 KB, 128B-swizzled = one TMA box).
// The 8 epilogue warps form two groups of 4 (one warp per TMEM lane quadrant); group g drains the column half
// [g*BN/2, (g+1)*BN/2) of the accumulator in 64-column slabs.  Every slab written to global is one "job": TMEM -> registers
// -> bias / GELU / residual -> swizzled st.shared into a slab buffer -> ONE TMA store.  A group cycles through NSLAB/2
// buffers, so the TMA store of job j drains while the group converts job j+1 (and the other group works on its own
// half): the read-out of a tile is a pipeline of small stores instead of one long serial drain.
template <int BN, int NSLAB>
struct P2Smem {
  static constexpr int A_BYTES = BM * BK * 2;             // this CTA's 128 rows of A
  static constexpr int B_BYTES = (BN / 2) * BK * 2;       // this CTA's half of the B tile
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int SLAB_BYTES = BM * 128;             // 128 rows x 64 bf16 (one TMA box)
  static constexpr int TAIL_BYTES = 512 /* barriers */ + 2 * BN * 2 /* bias, two tiles */;
  static constexpr int bytes(int stages) { return stages * STAGE_BYTES + NSLAB * SLAB_BYTES + 1024 + TAIL_BYTES; }
};

// Debug trace of the pair kernel (compiled in with -DDTC_GEMM_TRACE only; tools/gemm_pair_trace.py): the leader CTA of the first
// and of the last pair of every launch stamp clock64 at the phase boundaries of their first tile into a record of 16 words.
#ifdef DTC_GEMM_TRACE
__device__ unsigned long long* g_pair_trace = nullptr;  // [0] launch counter, records from word 16
#define GT_STAMP(i)                                    \
  do {                                                 \
    if (s_trace_rec) s_trace_rec[i] = clock64();       \
  } while (0)
DTC_DEVICE unsigned long long gt_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#else
#define GT_STAMP(i) do { } while (0)
#endif


// MODE >= 0: the epilogue mode is a compile-time constant (the launch's epi.mode must equal it); MODE < 0: read at run time.
// A run-time mode makes the compiler if-convert the mode branches: every slab then ISSUES the GELU / residual / statistics
// instructions of all modes predicated off — 700 instructions per warp and slab for a plain store, and with two epilogue
// warps per scheduler that issue time was the 1.5 us per slab the clock trace shows (profiles/r2/gemm_pair_trace.md).
template <bool A_MN, bool B_MN, int BN, int NSLAB, int MODE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(PG_THREADS, NSLAB == 2 ? 2 : 1)
gemm_2cta_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                 const __grid_constant__ CUtensorMap tma_out, const __grid_constant__ CUtensorMap tma_out2,
                 const __grid_constant__ CUtensorMap tma_aux, int M, int N, int K, int m_blocks /* of 256 rows */,
                 int n_blocks, int splits, int kb_per_split, int num_stages, GemmEpi epi) {
  using S = P2Smem<BN, NSLAB>;
  constexpr int GB = NSLAB / 2;  // slab buffers per epilogue group
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* slabs = smem + num_stages * S::STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(slabs + NSLAB * S::SLAB_BYTES);
  uint64_t* full_bar = bars;                      // [PG_STAGE_MAX]
  uint64_t* empty_bar = bars + PG_STAGE_MAX;      // [PG_STAGE_MAX]
  uint64_t* tmem_full = bars + 2 * PG_STAGE_MAX;  // [2]
  uint64_t* tmem_empty = tmem_full + 2;           // [2]
  uint64_t* aux_full = tmem_empty + 2;            // [NSLAB <= 4]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(aux_full + 4);
  bf16* s_bias = reinterpret_cast<bf16*>(bars + 64);  // [2][BN], 512 B past the barriers

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();  // 0 = leader
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int total_kb = (K + BK - 1) / BK;
  const int tiles_mn = m_blocks * n_blocks;
  const int total_tiles = tiles_mn * splits;
  const int mode = MODE >= 0 ? MODE : epi.mode;
  const bool bf16_out = mode <= EPI_DGELU || mode == EPI_STORE_STATS;
  const bool use_aux = (mode == EPI_BIAS_RESIDUAL || mode == EPI_DGELU);
#ifdef DTC_GEMM_TRACE
  unsigned long long*& s_trace_rec = *reinterpret_cast<unsigned long long**>(bars + 62);  // (spare word of the barrier block)
  if (threadIdx.x == 0) {
    unsigned long long* rec = nullptr;
    if (g_pair_trace != nullptr && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 2)) {
      const unsigned long long slot = atomicAdd(g_pair_trace, 1ull);
      if (slot < 4096) {
        rec = g_pair_trace + 16 + slot * 16;
        rec[0] = gt_globaltimer();
        rec[1] = clock64();
        rec[15] = ((unsigned long long)gridDim.x << 32) | blockIdx.x;
      }
    }
    s_trace_rec = rec;
  }
#endif

  // This CTA's share of a tile, decoded for the FIRST tile ahead of the prologue barriers: the integer divisions and the
  // cold kernel-parameter loads cost ~0.5 us when they sat between the cluster barrier and the first TMA issue (clock trace).
  auto decode_tile = [&](int tile, int& row0, int& bcol0, int& kb_begin, int& nkb) {
    const int split = tile / tiles_mn, mn = tile % tiles_mn;
    const int m_blk = mn % m_blocks, n_blk = mn / m_blocks;
    row0 = m_blk * 256 + (int)cta * BM;        // this CTA's rows of A / of the output
    bcol0 = n_blk * BN + (int)cta * (BN / 2);  // this CTA's half of the B tile
    kb_begin = split * kb_per_split;
    nkb = min(total_kb, kb_begin + kb_per_split) - kb_begin;
  };
  int f_row0 = 0, f_bcol0 = 0, f_kb_begin = 0, f_nkb = 0;
  if (pair < total_tiles) decode_tile(pair, f_row0, f_bcol0, f_kb_begin, f_nkb);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    if (bf16_out || MODE == EPI_ATOMIC_F32) tma_prefetch_desc(&tma_out);
    if (use_aux) tma_prefetch_desc(&tma_aux);
    if (mode == EPI_BIAS_GELU) tma_prefetch_desc(&tma_out2);
    for (int s = 0; s < num_stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full[b], 1);
      mbar_init(&tmem_empty[b], 16);  // 8 epilogue warps of each CTA (used in the leader only)
    }
    for (int b = 0; b < 4; ++b) mbar_init(&aux_full[b], 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc_2sm(tmem_ptr_smem, 2 * BN);
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_relaxed();  // the peer's barriers are initialised (and fenced: fence_barrier_init) before anything may arrive on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_launch_dependents();  // PDL: only after this CTA owns its TMEM (a dependent may not allocate first)
  pdl_wait();               // inputs come from the previous kernel: no global access before it has completed
  if (threadIdx.x == 0) GT_STAMP(2);

  if (warp == 0) {
    // ------------------------------ TMA producer (both CTAs) ------------------------------
    if (elect_one()) {
      uint32_t kit = 0;
      int s = 0;
      uint32_t ph = 0;  // ring position: stage s, lap parity ph
      for (int tile = pair; tile < total_tiles; tile += num_pairs) {
        int row0 = f_row0, bcol0 = f_bcol0, kb_begin = f_kb_begin, nkb = f_nkb;
        if (tile != pair) decode_tile(tile, row0, bcol0, kb_begin, nkb);
        for (int i = 0; i < nkb; ++i, ++kit) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* sa = smem + s * S::STAGE_BYTES;
          uint8_t* sb = sa + S::A_BYTES;
          const uint32_t full_leader = mapa_shared(smem_u32(&full_bar[s]), 0);
          if (cta == 0) mbar_expect_tx(&full_bar[s], 2 * S::STAGE_BYTES);  // bytes of both CTAs
          const int k0 = (kb_begin + i) * BK;
          if (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) tma_load_2d_2sm(sa + j * (BK * 128), &tma_a, full_leader, row0 + j * 64, k0);
          } else {
            tma_load_2d_2sm(sa, &tma_a, full_leader, k0, row0);
          }
          if (B_MN) {
#pragma unroll
            for (int j = 0; j < BN / 128; ++j) tma_load_2d_2sm(sb + j * (BK * 128), &tma_b, full_leader, bcol0 + j * 64, k0);
          } else {
            tma_load_2d_2sm(sb, &tma_b, full_leader, k0, bcol0);
          }
          if (kit == 0) GT_STAMP(3);
          if (++s == num_stages) {
            s = 0;
            ph ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer (leader CTA only) ------------------------------
    if (cta == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(256, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      constexpr uint32_t A_LBO = A_MN ? BK * 128 : 16, A_STEP = A_MN ? 2048 : 32;
      constexpr uint32_t B_LBO = B_MN ? BK * 128 : 16, B_STEP = B_MN ? 2048 : 32;
      uint32_t kit = 0;
      int s = 0;
      uint32_t ph = 0;
      int it = 0;
      for (int tile = pair; tile < total_tiles; tile += num_pairs, ++it) {
        int nkb = f_nkb;
        if (tile != pair) {
          const int kb_begin = (tile / tiles_mn) * kb_per_split;
          nkb = min(total_kb, kb_begin + kb_per_split) - kb_begin;
        }
        const int buf = it & 1;
        mbar_wait(&tmem_empty[buf], ((it >> 1) & 1) ^ 1);  // both CTAs' epilogues have drained this buffer
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * BN;
        for (int i = 0; i < nkb; ++i, ++kit) {
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          if (elect_one()) {
            if (kit == 0) GT_STAMP(4);
            const uint32_t sa = smem_u32(smem + s * S::STAGE_BYTES);
            const uint32_t sb = sa + S::A_BYTES;
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = make_smem_desc(sa + k * A_STEP, A_LBO, 1024, SWZ_128B);
              const uint64_t db = make_smem_desc(sb + k * B_STEP, B_LBO, 1024, SWZ_128B);
              umma_bf16_2sm(tmem_d, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
            }
            umma_commit_2sm(&empty_bar[s], 0x3);
            if (i == nkb - 1) umma_commit_2sm(&tmem_full[buf], 0x3);
            if (i == nkb - 1 && it == 0) GT_STAMP(5);
          }
          __syncwarp();
          if (++s == num_stages) {
            s = 0;
            ph ^= 1;
          }
        }
      }
    }
  } else {
    // ------------------------------ epilogue (warps 2..9 of both CTAs) ------------------------------
    const int q = warp & 3;                      // TMEM lane quadrant this warp may read
    const int g = (warp - 2) >> 2;               // epilogue group = column half of the tile
    const int row = q * 32 + lane;
    const int gt = (int)threadIdx.x - 64 - g * 128;  // thread index inside the group
    const bool leader = gt == 0;                 // issues this group's TMA loads / stores
    const float alpha = epi.alpha;
    const bool has_bias = epi.bias != nullptr && mode != EPI_ATOMIC_F32 && mode != EPI_DGELU;
    constexpr int SPG = BN / 128;                // 64-column slabs per group and tile
    const int gcol = g * (BN / 2);               // first tile column of this group
    uint8_t* gslab = slabs + g * GB * S::SLAB_BYTES;
    uint64_t* gaux = aux_full + g * GB;
    int it = 0;

    if (MODE == EPI_ATOMIC_F32) {
      // fp32 reduce-add (weight gradients, split-K) with the mode compiled in: the group's column half leaves in slabs of
      // [128 rows x 32 fp32] = one 128B-swizzled TMA box each, added to global memory by TMA reduce (tma_out is an fp32 map).
      // One bulk request per slab replaces 1024 16-byte red.global.add, each of which touched its own 128-byte line.
      uint32_t jc = 0;
      for (int tile = pair; tile < total_tiles; tile += num_pairs, ++it) {
        const int mn = tile % tiles_mn;
        const int m_blk = mn % m_blocks, n_blk = mn / m_blocks;
        const int buf = it & 1;
        const int row0 = m_blk * 256 + (int)cta * BM;
        mbar_wait(&tmem_full[buf], (it >> 1) & 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * BN + ((uint32_t)(q * 32) << 16) + gcol;
#pragma unroll 1
        for (int c = 0; c < BN / 64; ++c, ++jc) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(tmem_d + c * 32, r);
          tmem_ld_wait();
          if (c == BN / 64 - 1) {  // accumulator buffer drained by this warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty[buf]), 0));
          }
          const uint32_t sbuf_u32 = smem_u32(gslab + (jc % GB) * S::SLAB_BYTES);
          if (leader) tma_store_wait_read_n<GB - 1>();  // the reduce that last used this buffer has read it
          epi_group_sync(g);
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4)
            sts_v4(sbuf_u32 + swizzle_offset<3>(row * 128 + c4 * 16), __float_as_uint(__uint_as_float(r[c4 * 4]) * alpha),
                   __float_as_uint(__uint_as_float(r[c4 * 4 + 1]) * alpha),
                   __float_as_uint(__uint_as_float(r[c4 * 4 + 2]) * alpha),
                   __float_as_uint(__uint_as_float(r[c4 * 4 + 3]) * alpha));
          fence_proxy_async_smem();
          epi_group_sync(g);
          if (leader) {
            tma_reduce_add_2d(&tma_out, gslab + (jc % GB) * S::SLAB_BYTES, n_blk * BN + gcol + c * 32, row0);
            tma_store_commit();
          }
        }
      }
      if (leader) tma_store_wait_read();
    } else if (!bf16_out) {
      // fp32 outputs (fp32 store, or reduce-add into a buffer TMA cannot address): straight from registers, one row per thread
      for (int tile = pair; tile < total_tiles; tile += num_pairs, ++it) {
        const int mn = tile % tiles_mn;
        const int m_blk = mn % m_blocks, n_blk = mn / m_blocks;
        const int buf = it & 1;
        const long long gm = (long long)m_blk * 256 + (int)cta * BM + row;
        const bool row_ok = gm < M;
        mbar_wait(&tmem_full[buf], (it >> 1) & 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + buf * BN + ((uint32_t)(q * 32) << 16) + gcol;
#pragma unroll 1
        for (int c = 0; c < BN / 64; ++c) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(tmem_d + c * 32, r);
          tmem_ld_wait();
          if (c == BN / 64 - 1) {  // accumulator buffer drained by this warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty[buf]), 0));
          }
          const int n0 = n_blk * BN + gcol + c * 32;
#pragma unroll
          for (int c8 = 0; c8 < 4; ++c8) {
            const int n = n0 + c8 * 8;
            if (!row_ok || n >= N) continue;
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[c8 * 8 + j]) * alpha;
            if (has_bias) {
              const uint4 bb = *reinterpret_cast<const uint4*>(epi.bias + n);
              const uint32_t bw[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_bf16x2(bw[j]);
                v[2 * j] += f.x;
                v[2 * j + 1] += f.y;
              }
            }
            float* o = reinterpret_cast<float*>(epi.out) + gm * epi.ldo + n;
            if (mode == EPI_ATOMIC_F32) {
              red_add_v4_f32(o, v[0], v[1], v[2], v[3]);
              red_add_v4_f32(o + 4, v[4], v[5], v[6], v[7]);
            } else {
              *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
              *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
            }
          }
        }
      }
    } else {
      // bf16 outputs: pipelined slab jobs (see P2Smem)
      uint32_t jc = 0;  // jobs issued by this group so far (uniform over its threads): buffer = jc % GB
      if (use_aux && leader && pair < total_tiles) {  // residual / pre-activation slab of the first job
        const int mn = pair % tiles_mn;
        mbar_expect_tx(&gaux[0], S::SLAB_BYTES);
        tma_load_2d(gslab, &tma_aux, &gaux[0], (mn / m_blocks) * BN + gcol, (mn % m_blocks) * 256 + (int)cta * BM);
      }
      const int n_out = (mode == EPI_BIAS_GELU) ? 2 : 1;
      for (int tile = pair; tile < total_tiles; tile += num_pairs, ++it) {
        const int mn = tile % tiles_mn;
        const int m_blk = mn % m_blocks, n_blk = mn / m_blocks;
        const int buf = it & 1;
        const int row0 = m_blk * 256 + (int)cta * BM;
        bf16* tbias = s_bias + (it & 1) * BN + gcol;
        if (has_bias) {  // this group's bias slice -> smem while the mainloop of the tile is still running
          if (gt < BN / 16) {
            const int n = n_blk * BN + gcol + gt * 8;
            uint4 bv = make_uint4(0, 0, 0, 0);
            if (n < N) bv = *reinterpret_cast<const uint4*>(epi.bias + n);
            *reinterpret_cast<uint4*>(tbias + gt * 8) = bv;
          }
          epi_group_sync(g);
        }
        mbar_wait(&tmem_full[buf], (it >> 1) & 1);
        tc_fence_after();
        if (it == 0 && threadIdx.x == 64) GT_STAMP(6);
        const uint32_t tmem_d = tmem_base + buf * BN + ((uint32_t)(q * 32) << 16) + gcol;
        float run_m = -INFINITY, run_s = 0.f;  // EPI_STORE_STATS: this row's (max, sum exp) over the group's columns of the tile
#pragma unroll 1
        for (int sl = 0; sl < SPG; ++sl) {
          uint32_t r[64];
          tmem_ld_32x32b_x32(tmem_d + sl * 64, r);
          tmem_ld_32x32b_x32(tmem_d + sl * 64 + 32, r + 32);
          tmem_ld_wait();
          if (it == 0 && sl == 0 && threadIdx.x == 64) GT_STAMP(12);
          if (sl == SPG - 1) {  // this warp has read its whole share of the accumulator: hand the buffer back early
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&tmem_empty[buf]), 0));
          }
          // v = alpha * acc + bias, kept in r as float bits
          if (has_bias) {
            const uint32_t bias_u32 = smem_u32(tbias + sl * 64);
#pragma unroll
            for (int c8 = 0; c8 < 8; ++c8) {
              const uint4 bb = lds_v4(bias_u32 + c8 * 16);
              const uint32_t bw[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float2 f = unpack_bf16x2(bw[j]);
                r[c8 * 8 + 2 * j] = __float_as_uint(fmaf(__uint_as_float(r[c8 * 8 + 2 * j]), alpha, f.x));
                r[c8 * 8 + 2 * j + 1] = __float_as_uint(fmaf(__uint_as_float(r[c8 * 8 + 2 * j + 1]), alpha, f.y));
              }
            }
          } else if (alpha != 1.f) {
#pragma unroll
            for (int j = 0; j < 64; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) * alpha);
          }
          const int col0 = n_blk * BN + gcol + sl * 64;
          if (mode == EPI_STORE_STATS) {
            // statistics of the ROUNDED values (the loss kernel re-reads the bf16 logits), pad columns masked out
            const int n_valid = (int)epi.ldo2;
            // (two passes over the registers, rounding twice, instead of keeping 64 rounded values alive)
            auto rounded = [&](int c2, float& x, float& y) {
              const float2 f = unpack_bf16x2(pack_bf16x2(__uint_as_float(r[2 * c2]), __uint_as_float(r[2 * c2 + 1])));
              x = (col0 + 2 * c2 < n_valid) ? f.x : -INFINITY;
              y = (col0 + 2 * c2 + 1 < n_valid) ? f.y : -INFINITY;
            };
            float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
            for (int c2 = 0; c2 < 32; ++c2) {
              float x, y;
              rounded(c2, x, y);
              m4[c2 & 3] = fmaxf(m4[c2 & 3], fmaxf(x, y));
            }
            const float m_slab = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
            if (m_slab > -INFINITY) {
              constexpr float L2E = 1.4426950408889634f;
              if (m_slab > run_m) {
                run_s *= ex2_ftz((run_m - m_slab) * L2E);  // run_m = -inf: 0 * 0
                run_m = m_slab;
              }
              const float mb = run_m * L2E;
              float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
              for (int c2 = 0; c2 < 32; ++c2) {
                float x, y;
                rounded(c2, x, y);
                s4[c2 & 3] += ex2_ftz(fmaf(x, L2E, -mb)) + ex2_ftz(fmaf(y, L2E, -mb));  // exp2(-inf) = 0 for masked columns
              }
              run_s += (s4[0] + s4[1]) + (s4[2] + s4[3]);
            }
            if (sl == SPG - 1 && (long long)row0 + row < M) {
              const int n_parts = 2 * n_blocks;
              reinterpret_cast<float2*>(epi.out2)[((long long)row0 + row) * n_parts + 2 * n_blk + g] = make_float2(run_m, run_s);
            }
          }
#pragma unroll 1
          for (int o = 0; o < n_out; ++o, ++jc) {
            const int b = jc % GB;
            uint8_t* sbuf = gslab + b * S::SLAB_BYTES;
            const uint32_t sbuf_u32 = smem_u32(sbuf);
            if (use_aux) {
              mbar_wait(&gaux[b], (jc / GB) & 1);  // the aux slab has landed (it was requested once the buffer was free)
            } else {
              if (leader) tma_store_wait_read_n<GB - 1>();  // the store that last used this buffer has read it
              epi_group_sync(g);
            }
            if (it == 0 && sl == 0 && o == 0 && threadIdx.x == 64) GT_STAMP(13);
#pragma unroll
            for (int c8 = 0; c8 < 8; ++c8) {
              const uint32_t off = swizzle_offset<3>(row * 128 + c8 * 16);
              float v[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[c8 * 8 + j]);
              if (use_aux) {
                const uint4 a = lds_v4(sbuf_u32 + off);
                const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const float2 f = unpack_bf16x2(aw[j]);
                  if (mode == EPI_BIAS_RESIDUAL) {
                    v[2 * j] += f.x;
                    v[2 * j + 1] += f.y;
                  } else {
                    v[2 * j] *= gelu_tanh_grad(f.x);
                    v[2 * j + 1] *= gelu_tanh_grad(f.y);
                  }
                }
              }
              if (o == 1) {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = gelu_tanh(v[j]);
              }
              sts_v4(sbuf_u32 + off, pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                     pack_bf16x2(v[6], v[7]));
            }
            fence_proxy_async_smem();
            if (it == 0 && sl == 0 && o == 0 && threadIdx.x == 64) GT_STAMP(14);
            epi_group_sync(g);
            if (leader) {
              tma_store_2d(o == 0 ? &tma_out : &tma_out2, sbuf, col0, row0);
              tma_store_commit();
              if (it == 0 && g == 0 && sl == SPG - 1 && o == n_out - 1) GT_STAMP(7);
              if (use_aux) {
                // request the aux slab of the group's NEXT job (same tile, next slab — or the next tile of this pair)
                int ntile = tile, nsl = sl + 1;
                if (nsl == SPG) {
                  ntile = tile + num_pairs;
                  nsl = 0;
                }
                if (ntile < total_tiles) {
                  const int nb = (jc + 1) % GB;
                  const int nmn = ntile % tiles_mn;
                  tma_store_wait_read_n<GB - 1>();  // the store that last used buffer nb has read it
                  mbar_expect_tx(&gaux[nb], S::SLAB_BYTES);
                  tma_load_2d(gslab + nb * S::SLAB_BYTES, &tma_aux, &gaux[nb], (nmn / m_blocks) * BN + gcol + nsl * 64,
                              (nmn % m_blocks) * 256 + (int)cta * BM);
                }
              }
            }
          }
        }
      }
      // shared memory must outlive the stores' READS only; their writes are complete (and visible to the next kernel) when
      // the grid is — waiting for the writes here kept every CTA ~0.5 us longer than necessary
      if (leader) tma_store_wait_read();
      if (leader && g == 0) GT_STAMP(8);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) GT_STAMP(9);
  cluster_sync_relaxed();  // neither CTA leaves (or frees TMEM) while its peer may still touch its smem / barriers
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_2sm(tmem_base, 2 * BN);
#ifdef DTC_GEMM_TRACE
    if (lane == 0 && s_trace_rec) {
      s_trace_rec[10] = clock64();
      s_trace_rec[11] = gt_globaltimer();
    }
#endif
  }
}

template <bool A_MN, bool B_MN, int BN, int NSLAB, int MODE>
static int launch_gemm_2cta_mode(const void* A, long long lda, const void* B, long long ldb, int M, int N, int K, int splits,
                            int stages, const GemmEpi& epi, cudaStream_t stream) {
  using S = P2Smem<BN, NSLAB>;
  CUtensorMap ta, tb, tout, tout2, taux;
  int rc;
  if (A_MN) rc = make_tmap_2d_bf16(&ta, A, K, M, lda, BK, 64, CU_TENSOR_MAP_SWIZZLE_128B);
  else rc = make_tmap_2d_bf16(&ta, A, M, K, lda, BM, BK, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  if (B_MN) rc = make_tmap_2d_bf16(&tb, B, K, N, ldb, BK, 64, CU_TENSOR_MAP_SWIZZLE_128B);
  else rc = make_tmap_2d_bf16(&tb, B, N, K, ldb, BN / 2, BK, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  tout = ta;
  tout2 = ta;
  taux = ta;
  if (epi.mode <= EPI_DGELU || epi.mode == EPI_STORE_STATS) {
    if ((rc = make_tmap_2d_bf16(&tout, epi.out, M, N, epi.ldo, BM, 64, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    if (epi.mode == EPI_BIAS_GELU &&
        (rc = make_tmap_2d_bf16(&tout2, epi.out2, M, N, epi.ldo2, BM, 64, CU_TENSOR_MAP_SWIZZLE_128B)))
      return rc;
    if ((epi.mode == EPI_BIAS_RESIDUAL || epi.mode == EPI_DGELU) &&
        (rc = make_tmap_2d_bf16(&taux, epi.aux, M, N, epi.ld_aux, BM, 64, CU_TENSOR_MAP_SWIZZLE_128B)))
      return rc;
  } else if (MODE == EPI_ATOMIC_F32) {
    if ((rc = make_tmap_2d_f32(&tout, epi.out, M, N, epi.ldo, BM, 32, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  }
  const int total_kb = (K + BK - 1) / BK;
  if (splits < 1) splits = 1;
  if (splits > total_kb) splits = total_kb;
  const int kb_per_split = (total_kb + splits - 1) / splits;
  splits = (total_kb + kb_per_split - 1) / kb_per_split;
  if (stages < 2) stages = 2;
  if (stages > PG_STAGE_MAX) stages = PG_STAGE_MAX;
  const int smem_cap = NSLAB == 2 ? 112 * 1024 : 227 * 1024;  // light configuration: two CTAs (+1 KB reserved each) per SM
  while (stages > 2 && S::bytes(stages) > smem_cap) --stages;
  const int smem_bytes = S::bytes(stages);
  auto kern = gemm_2cta_kernel<A_MN, B_MN, BN, NSLAB, MODE>;
  static dtc::SmemOptIn configured;  // per device (and per instantiation)
  DTC_CHECK_CUDA(dtc::ensure_dyn_smem(kern, NSLAB == 2 ? 113 * 1024 : 227 * 1024, configured));
  const int m_blocks = (M + 255) / 256, n_blocks = (N + BN - 1) / BN;
  const int total_tiles = m_blocks * n_blocks * splits;
  const int max_pairs = (NSLAB == 2 ? 2 : 1) * (num_sms() / 2);
  const int pairs = total_tiles < max_pairs ? total_tiles : max_pairs;
  DTC_CHECK_CUDA(dtc::launch(kern, 2 * pairs, PG_THREADS, smem_bytes, stream, ta, tb, tout, tout2, taux, M, N, K, m_blocks, n_blocks, splits,
                                                      kb_per_split, stages, epi));
  DTC_CHECK_LAUNCH();
  return 0;
}

// The epilogue modes the model's GEMMs use get their own instantiation (see the kernel's MODE parameter); the rest share
// the run-time-mode one.  Transposed-A problems are weight gradients: fp32 reduce-add, or the run-time-mode kernel.
template <bool A_MN, bool B_MN, int BN, int NSLAB>
static int launch_gemm_2cta(const void* A, long long lda, const void* B, long long ldb, int M, int N, int K, int splits,
                            int stages, const GemmEpi& epi, cudaStream_t stream) {
#define DTC_2CTA_MODE(MODE) \
  return launch_gemm_2cta_mode<A_MN, B_MN, BN, NSLAB, MODE>(A, lda, B, ldb, M, N, K, splits, stages, epi, stream)
  // (the TMA reduce epilogue needs a 16-byte aligned fp32 output with 16-byte row pitch; anything else takes the register path)
  if (epi.mode == EPI_ATOMIC_F32 && epi.ldo % 4 == 0 && ((uintptr_t)epi.out & 15) == 0) DTC_2CTA_MODE(EPI_ATOMIC_F32);
  if constexpr (!A_MN) {
    switch (epi.mode) {
      case EPI_STORE_BF16: DTC_2CTA_MODE(EPI_STORE_BF16);
      case EPI_BIAS_GELU: DTC_2CTA_MODE(EPI_BIAS_GELU);
      case EPI_BIAS_RESIDUAL: DTC_2CTA_MODE(EPI_BIAS_RESIDUAL);
      case EPI_DGELU: DTC_2CTA_MODE(EPI_DGELU);
      case EPI_STORE_STATS: DTC_2CTA_MODE(EPI_STORE_STATS);
      default: break;
    }
  }
  DTC_2CTA_MODE(-1);
#undef DTC_2CTA_MODE
}

}  // namespace dtc

using namespace dtc;

static int g_gemm_dual = 1;  // bit 0: allow the two-CTAs-per-SM persistent variant; bit 1: 64-wide tiles for small grids
static int g_gemm_bn64 = 1;
static int g_gemm_coop_store = 0;  // bit 2 of dtc_gemm_set_dual: epilogue warps store the tile (no TMA store)
extern "C" int dtc_gemm_set_dual(int on) {
  g_gemm_dual = on & 1;
  g_gemm_bn64 = (on >> 1) & 1;
  g_gemm_coop_store = (on >> 2) & 1;
  return 0;
}
static unsigned long long* g_gemm_trace = nullptr;  // debug timestamps (tools/gemm_trace.py)
extern "C" int dtc_gemm_set_trace(unsigned long long* buf) {
  g_gemm_trace = buf;
  return 0;
}

static int g_gemm_impl = 0;  // 0 = persistent 1-CTA kernel, 1 = one-tile-per-CTA kernel, 2 = CTA-pair 256x128, 3 = CTA-pair, 256x256 where it fits
// Process-wide default variant of dtc_gemm_bf16 (tools / C callers); the Python layer passes the variant per call (dtc_gemm_bf16_ex).
extern "C" int dtc_gemm_set_impl(int impl) {
  g_gemm_impl = impl;
  return 0;
}

// C ABI -----------------------------------------------------------------------------------------
// A: a_mn_major == 0 -> [M, K] row-major (lda elements between rows); == 1 -> [K, M] row-major.
// B: b_mn_major == 0 -> [N, K] row-major;                             == 1 -> [K, N] row-major.
// Requirements: all leading dimensions and base pointers 16-byte aligned; N % 8 == 0.
// block_n: 64 / 128 / 256 (0 = choose); splits > 1 requires epi_mode == EPI_ATOMIC_F32.
static int gemm_bf16_dispatch(const void* A, long long lda, int a_mn_major, const void* B, long long ldb, int b_mn_major,
                             int M, int N, int K, int epi_mode, float alpha, const void* bias, const void* aux,
                             long long ld_aux, void* out, long long ldo, void* out2, long long ldo2, int splits,
                             int stages, int block_n, int impl, cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && N > 0 && K > 0, "gemm: empty problem M=%d N=%d K=%d", M, N, K);
  DTC_CHECK_ARG(N % 8 == 0, "gemm: N=%d must be a multiple of 8", N);
  DTC_CHECK_ARG(lda % 8 == 0 && ldb % 8 == 0, "gemm: lda=%lld ldb=%lld must be multiples of 8 elements", lda, ldb);
  DTC_CHECK_ARG(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0 && ((uintptr_t)out & 15) == 0,
                "gemm: operand pointers must be 16-byte aligned");
  DTC_CHECK_ARG((epi_mode >= 0 && epi_mode <= 5) || epi_mode == EPI_STORE_STATS, "gemm: bad epilogue mode %d (the scatter epilogue has its own entry point)", epi_mode);
  DTC_CHECK_ARG(splits <= 1 || epi_mode == EPI_ATOMIC_F32, "gemm: split-K needs the fp32 reduce-add epilogue");
  DTC_CHECK_ARG(!(epi_mode == EPI_BIAS_RESIDUAL || epi_mode == EPI_DGELU) || aux != nullptr, "gemm: aux missing");
  DTC_CHECK_ARG(epi_mode != EPI_BIAS_GELU || out2 != nullptr, "gemm: out2 missing");
  DTC_CHECK_ARG(ldo % 8 == 0 && (aux == nullptr || ld_aux % 8 == 0) &&
                    (out2 == nullptr || ldo2 % 8 == 0 || epi_mode == EPI_STORE_STATS),
                "gemm: output/aux leading dimensions must be multiples of 8");
  DTC_CHECK_ARG(epi_mode != EPI_STORE_STATS || (out2 != nullptr && M >= 256 && N >= 128 && !a_mn_major && splits <= 1 &&
                                                ldo2 > 0 && ldo2 <= N && (block_n == 128 || block_n == 256)),
                "gemm: the statistics epilogue needs out2 = float2 stats, ldo2 = valid columns (1..N), an explicit block_n of "
                "128 or 256 (it fixes the number of parts per row: 2 * ceil(N / block_n)), M >= 256, N >= 128, no split-K");
  GemmEpi epi;
  epi.mode = epi_mode;
  epi.alpha = alpha;
  epi.bias = reinterpret_cast<const bf16*>(bias);
  epi.aux = reinterpret_cast<const bf16*>(aux);
  epi.ld_aux = ld_aux;
  epi.out = out;
  epi.ldo = ldo;
  epi.out2 = reinterpret_cast<bf16*>(out2);
  epi.ldo2 = ldo2;
  epi.trace = g_gemm_trace;
  epi.coop_store = g_gemm_coop_store;
  const int block_n_arg = block_n;  // 0 = let the dispatcher choose (the pair kernel decides on its own width below)
  if (block_n == 0) {
    block_n = (N <= 64) ? 64 : 128;
    // few 128-wide tiles (N = 512 layers: 128 tiles for 148 SMs): 64-wide tiles double the CTA count so that every SM
    // runs two CTAs, each served by the TMA path at its own ~48 B/cycle
    const long long t128 = (long long)((M + 127) / 128) * ((N + 127) / 128) * (splits > 1 ? splits : 1);
    if (g_gemm_dual && g_gemm_bn64 && t128 <= num_sms() && N % 64 == 0 && epi_mode != EPI_BIAS_GELU)
      block_n = 64;
  }
  static const bool env_v1 = getenv("DTC_GEMM_V1") != nullptr;  // A/B switch: one-tile-per-CTA kernel
  const bool use_v1 = env_v1 || impl == 1;
  // CTA-pair kernel: needs >= 256 rows; 256-wide tiles when that still leaves enough pair-tiles to fill the GPU.
  // Default policy (impl 0): every GEMM with >= 256 rows and >= 128 columns runs on CTA pairs.  With the epilogue mode
  // compiled in, relaxed cluster barriers and the TMA reduce-add epilogue the pair kernel beats the 1-CTA persistent kernel
  // on every shape of the model, at 4096 tokens per GPU and at the 512 / 1024 / 2048 of the strong-scaling runs alike
  // (profiles/r2/gemm_microbench_mode_templated.txt, gemm_microbench_small_m.txt: e.g. M = 512: qkv 7.6 -> 6.3, fc2 17.2 ->
  // 12.2, fc1 wgrad 17.0 -> 11.1, lm_head wgrad 99 -> 54 us).  Narrower problems (tensor-parallel shards below 128 columns,
  // fewer than 256 tokens) stay on the 1-CTA kernel with its 64-wide tiles.  DTC_GEMM_PAIR_LAYERS=0 restores the old split
  // (pairs only for the lm_head-sized GEMMs).
  static const bool pair_layers = !(getenv("DTC_GEMM_PAIR_LAYERS") && getenv("DTC_GEMM_PAIR_LAYERS")[0] == '0');
  const double flops = 2.0 * M * N * (double)K;
  const bool big = flops >= 1e11 || (flops >= 2e10 && splits > 1);  // (split-K: the lm_head dgrad of a small local batch)
  const bool mid = pair_layers && epi_mode <= EPI_STORE_F32;
  const bool want_2cta = ((impl == 2 || impl == 3 || (impl == 0 && (big || mid) && !env_v1)) &&
                          M >= 256 && N >= 128) || epi_mode == EPI_STORE_STATS;  // impl 4 forces the 1-CTA persistent kernel
  if (want_2cta) {
    const long long tiles256 = (long long)((M + 255) / 256) * ((N + 255) / 256) * (splits > 1 ? splits : 1);
    // 256-wide pair tiles (half the operand bytes per MMA cycle again) whenever they still fill all 74 CTA pairs;
    // ragged N is fine: TMA zero-fills the loads and clips the stores
    const bool wide = (block_n_arg == 256) || ((impl == 3 || impl == 0) && block_n_arg == 0 && tiles256 >= 56);  // >= 3/4 of the 74 CTA pairs busy
    if (stages == 0) stages = PG_STAGE_MAX;
#define DTC_2CTA_DISPATCH(AMN, BMN)                                                                               \
  if (wide) return launch_gemm_2cta<AMN, BMN, 256, 4>(A, lda, B, ldb, M, N, K, splits, stages, epi, stream);      \
  return launch_gemm_2cta<AMN, BMN, 128, 4>(A, lda, B, ldb, M, N, K, splits, stages, epi, stream);
    if (!a_mn_major && !b_mn_major) { DTC_2CTA_DISPATCH(false, false) }
    if (!a_mn_major && b_mn_major) { DTC_2CTA_DISPATCH(false, true) }
    if (a_mn_major && !b_mn_major) { DTC_2CTA_DISPATCH(true, false) }
    DTC_2CTA_DISPATCH(true, true)
#undef DTC_2CTA_DISPATCH
  }
  if (!use_v1 && block_n != 256) {
    if (stages == 0) stages = PG_STAGE_MAX;
    // two CTAs per SM whenever there is more than one tile per SM to hand out (not for the two-output GELU epilogue
    // and the multicast-reduce epilogue, which need the second staging tile)
    const long long tiles128 = (long long)((M + 127) / 128) * ((N + block_n - 1) / block_n) * (splits > 1 ? splits : 1);
    // (always: the light configuration also lets a GEMM from the side stream share the SM)
    const bool dual = g_gemm_dual && epi_mode != EPI_BIAS_GELU;
    (void)tiles128;
#define DTC_PGEMM_DISPATCH(AMN, BMN)                                                                                      \
  if (dual) {                                                                                                             \
    if (block_n == 64) return launch_gemm_persistent<AMN, BMN, 64, true>(A, lda, B, ldb, M, N, K, splits, stages, epi, stream);  \
    return launch_gemm_persistent<AMN, BMN, 128, true>(A, lda, B, ldb, M, N, K, splits, stages, epi, stream);            \
  }                                                                                                                       \
  if (block_n == 64) return launch_gemm_persistent<AMN, BMN, 64, false>(A, lda, B, ldb, M, N, K, splits, stages, epi, stream);   \
  return launch_gemm_persistent<AMN, BMN, 128, false>(A, lda, B, ldb, M, N, K, splits, stages, epi, stream);
    if (!a_mn_major && !b_mn_major) { DTC_PGEMM_DISPATCH(false, false) }
    if (!a_mn_major && b_mn_major) { DTC_PGEMM_DISPATCH(false, true) }
    if (a_mn_major && !b_mn_major) { DTC_PGEMM_DISPATCH(true, false) }
    DTC_PGEMM_DISPATCH(true, true)
#undef DTC_PGEMM_DISPATCH
  }
  if (stages == 0) stages = (block_n == 256) ? 4 : (block_n == 128 ? 3 : 4);
#define DTC_GEMM_DISPATCH(AMN, BMN)                                                                          \
  switch (block_n) {                                                                                         \
    case 64: return launch_gemm<AMN, BMN, 64>(A, lda, B, ldb, M, N, K, splits, stages, epi, stream);         \
    case 128: return launch_gemm<AMN, BMN, 128>(A, lda, B, ldb, M, N, K, splits, stages, epi, stream);       \
    case 256: return launch_gemm<AMN, BMN, 256>(A, lda, B, ldb, M, N, K, splits, stages, epi, stream);       \
    default: dtc_set_error("gemm: block_n must be 64, 128 or 256 (got %d)", block_n); return -1;             \
  }
  if (!a_mn_major && !b_mn_major) { DTC_GEMM_DISPATCH(false, false) }
  if (!a_mn_major && b_mn_major) { DTC_GEMM_DISPATCH(false, true) }
  if (a_mn_major && !b_mn_major) { DTC_GEMM_DISPATCH(true, false) }
  DTC_GEMM_DISPATCH(true, true)
#undef DTC_GEMM_DISPATCH
}

extern "C" int dtc_gemm_bf16(const void* A, long long lda, int a_mn_major, const void* B, long long ldb, int b_mn_major,
                             int M, int N, int K, int epi_mode, float alpha, const void* bias, const void* aux,
                             long long ld_aux, void* out, long long ldo, void* out2, long long ldo2, int splits,
                             int stages, int block_n, cudaStream_t stream) {
  return gemm_bf16_dispatch(A, lda, a_mn_major, B, ldb, b_mn_major, M, N, K, epi_mode, alpha, bias, aux, ld_aux, out, ldo, out2,
                            ldo2, splits, stages, block_n, g_gemm_impl, stream);
}

// Same with the kernel variant chosen PER CALL (no process-global state: safe with several host threads / engines that
// want different policies).  impl: 0 default policy, 1 one-tile-per-CTA, 2 CTA pair 256x128, 3 CTA pair 256x256 where it
// fits, 4 the 1-CTA persistent kernel.
extern "C" int dtc_gemm_bf16_ex(const void* A, long long lda, int a_mn_major, const void* B, long long ldb, int b_mn_major,
                                int M, int N, int K, int epi_mode, float alpha, const void* bias, const void* aux,
                                long long ld_aux, void* out, long long ldo, void* out2, long long ldo2, int splits,
                                int stages, int block_n, int impl, cudaStream_t stream) {
  return gemm_bf16_dispatch(A, lda, a_mn_major, B, ldb, b_mn_major, M, N, K, epi_mode, alpha, bias, aux, ld_aux, out, ldo, out2,
                            ldo2, splits, stages, block_n, impl, stream);
}

// Fused GEMM -> reduce-scatter (sequence-parallel tensor parallelism; SURVEY.md §8 row R4, collectives X3/X4/X5/X8).
// D[M,N] = A·Bᵀ as dtc_gemm_bf16 (A [M,K] row-major; B [N,K] row-major, or [K,N] when b_mn_major), but the output never
// lands in one local matrix: rows [o*M/n, (o+1)*M/n) are stored into slot `rank` of rank o's partial-sum buffer
//   peer_out[o][(rank*M/n + row - o*M/n) * ldo + col]                (bf16; P2P stores over NVLink, or local for o == rank)
// `peer_out` is a HOST array of n_ranks device pointers (the symmetric buffer of every rank, each [M, ldo] bf16).
// The caller separates this kernel from the owner's reduction with a cross-rank barrier (dtc_symm_barrier).
template <bool SIGNAL>
static int gemm_scatter_impl(const void* A, long long lda, const void* B, long long ldb, int b_mn_major, int M, int N, int K,
                             const unsigned long long* peer_out, int n_ranks, int rank, long long ldo, int stages,
                             int block_n, const unsigned long long* sig_ptrs, int* counters, cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && N > 0 && K > 0, "gemm_scatter: empty problem M=%d N=%d K=%d", M, N, K);
  DTC_CHECK_ARG(n_ranks >= 1 && n_ranks <= 8 && rank >= 0 && rank < n_ranks, "gemm_scatter: bad rank %d of %d", rank, n_ranks);
  DTC_CHECK_ARG(M % (n_ranks * BM) == 0, "gemm_scatter: M=%d must be a multiple of n_ranks*128 (no tile may straddle owners)", M);
  DTC_CHECK_ARG(N % 64 == 0 && lda % 8 == 0 && ldb % 8 == 0 && ldo % 8 == 0, "gemm_scatter: N %% 64, ld %% 8 required");
  DTC_CHECK_ARG(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "gemm_scatter: operand pointers must be 16-byte aligned");
  typename EpiArgT<true, SIGNAL>::type epi;
  epi.mode = EPI_SCATTER_BF16;
  epi.alpha = 1.f;
  epi.bias = nullptr;
  epi.aux = nullptr;
  epi.ld_aux = 0;
  epi.out = nullptr;
  epi.ldo = ldo;
  epi.out2 = nullptr;
  epi.ldo2 = 0;
  epi.coop_store = 1;
  epi.trace = nullptr;
  for (int i = 0; i < 8; ++i) {
    epi.peer_out[i] = i < n_ranks ? reinterpret_cast<bf16*>(peer_out[i]) : nullptr;
    DTC_CHECK_ARG(i >= n_ranks || (peer_out[i] != 0 && (peer_out[i] & 15) == 0), "gemm_scatter: peer pointer %d null/unaligned", i);
  }
  epi.sp_rank = rank;
  epi.sp_rows = M / n_ranks;
  if constexpr (SIGNAL) {
    DTC_CHECK_ARG(sig_ptrs != nullptr && counters != nullptr, "gemm_scatter_signal: synchronisation block missing");
    for (int i = 0; i < SP_MAX_RANKS; ++i) epi.sync.sig[i] = i < n_ranks ? reinterpret_cast<int*>(sig_ptrs[i]) : nullptr;
    epi.sync.counters = counters;
    epi.sync.rank = rank;
    epi.sync.world = n_ranks;
  }
  if (block_n == 0) {
    const long long t128 = (long long)(M / 128) * ((N + 127) / 128);
    block_n = (N <= 64 || t128 <= num_sms()) ? 64 : 128;
  }
  DTC_CHECK_ARG(block_n == 64 || block_n == 128, "gemm_scatter: block_n must be 64 or 128 (got %d)", block_n);
  if (stages == 0) stages = PG_STAGE_MAX;
  if (b_mn_major) {
    if (block_n == 64)
      return launch_gemm_persistent<false, true, 64, true, true, SIGNAL>(A, lda, B, ldb, M, N, K, 1, stages, epi, stream);
    return launch_gemm_persistent<false, true, 128, true, true, SIGNAL>(A, lda, B, ldb, M, N, K, 1, stages, epi, stream);
  }
  if (block_n == 64)
    return launch_gemm_persistent<false, false, 64, true, true, SIGNAL>(A, lda, B, ldb, M, N, K, 1, stages, epi, stream);
  return launch_gemm_persistent<false, false, 128, true, true, SIGNAL>(A, lda, B, ldb, M, N, K, 1, stages, epi, stream);
}

extern "C" int dtc_gemm_bf16_scatter(const void* A, long long lda, const void* B, long long ldb, int b_mn_major, int M, int N,
                                     int K, const unsigned long long* peer_out, int n_ranks, int rank, long long ldo,
                                     int stages, int block_n, cudaStream_t stream) {
  return gemm_scatter_impl<false>(A, lda, B, ldb, b_mn_major, M, N, K, peer_out, n_ranks, rank, ldo, stages, block_n, nullptr,
                                  nullptr, stream);
}

// Folded-barrier variant (experimental, `DTC_SP_FOLD=1`): the last CTA of the grid publishes the scattered tiles to every
// rank (sequence number in sig[rank] on every peer) — no barrier kernel between this GEMM and dtc_sp_ln_*_sync(wait_in = 1).
// sig_ptrs / counters: see dtc_sp_ln_fwd_sync.
extern "C" int dtc_gemm_bf16_scatter_signal(const void* A, long long lda, const void* B, long long ldb, int b_mn_major, int M,
                                            int N, int K, const unsigned long long* peer_out, int n_ranks, int rank,
                                            long long ldo, int stages, int block_n, const unsigned long long* sig_ptrs,
                                            int* counters, cudaStream_t stream) {
  return gemm_scatter_impl<true>(A, lda, B, ldb, b_mn_major, M, N, K, peer_out, n_ranks, rank, ldo, stages, block_n, sig_ptrs,
                                 counters, stream);
}

// Grouped weight gradients (see gemm_grouped_wgrad_kernel): for p < count
//   out[p][M_p, N_p] (fp32, leading dimension ldo[p]) += alpha * A[p]ᵀ · B[p],   A[p] = [K, M_p], B[p] = [K, N_p] bf16 row-major.
// All arrays are HOST arrays of `count` (<= 4) entries.  M_p, N_p multiples of 8; N_p % 64 == 0.
extern "C" int dtc_gemm_wgrad_grouped(int count, const unsigned long long* A, const long long* lda,
                                      const unsigned long long* B, const long long* ldb, const int* Ms, const int* Ns, int K,
                                      const unsigned long long* out, const long long* ldo, float alpha, int splits,
                                      int stages, cudaStream_t stream) {
  DTC_CHECK_ARG(count >= 1 && count <= GG_MAX && K > 0, "wgrad_grouped: count=%d (1..%d), K=%d", count, GG_MAX, K);
  constexpr int BN = 128;
  using S = PgSmem<BN>;
  GemmGroup grp;
  memset(&grp, 0, sizeof(grp));
  grp.count = count;
  int tiles = 0;
  for (int p = 0; p < count; ++p) {
    DTC_CHECK_ARG(Ms[p] > 0 && Ns[p] > 0 && Ms[p] % 8 == 0 && Ns[p] % 8 == 0 && lda[p] % 8 == 0 && ldb[p] % 8 == 0 &&
                      ldo[p] % 4 == 0,
                  "wgrad_grouped: problem %d: M=%d N=%d lda=%lld ldb=%lld ldo=%lld", p, Ms[p], Ns[p], lda[p], ldb[p], ldo[p]);
    DTC_CHECK_ARG((A[p] & 15) == 0 && (B[p] & 15) == 0 && (out[p] & 15) == 0 && A[p] && B[p] && out[p],
                  "wgrad_grouped: problem %d: pointers must be non-null and 16-byte aligned", p);
    int rc;
    if ((rc = make_tmap_2d_bf16(&grp.a[p], reinterpret_cast<const void*>(A[p]), K, Ms[p], lda[p], BK, 64,
                                CU_TENSOR_MAP_SWIZZLE_128B)))
      return rc;
    if ((rc = make_tmap_2d_bf16(&grp.b[p], reinterpret_cast<const void*>(B[p]), K, Ns[p], ldb[p], BK, 64,
                                CU_TENSOR_MAP_SWIZZLE_128B)))
      return rc;
    grp.out[p] = reinterpret_cast<float*>(out[p]);
    grp.ldo[p] = ldo[p];
    grp.M[p] = Ms[p];
    grp.N[p] = Ns[p];
    grp.m_blocks[p] = (Ms[p] + BM - 1) / BM;
    grp.tile_start[p] = tiles;
    tiles += grp.m_blocks[p] * ((Ns[p] + BN - 1) / BN);
  }
  for (int p = count; p <= GG_MAX; ++p) grp.tile_start[p] = tiles;
  const int total_kb = (K + BK - 1) / BK;
  if (splits < 1) splits = 1;
  if (splits > total_kb) splits = total_kb;
  const int kb_per_split = (total_kb + splits - 1) / splits;
  splits = (total_kb + kb_per_split - 1) / kb_per_split;
  if (stages < 2) stages = 3;
  if (stages > PG_STAGE_MAX) stages = PG_STAGE_MAX;
  const int smem_cap = 112 * 1024;  // two CTAs per SM
  auto smem_of = [](int st) { return st * S::STAGE_BYTES + 1024 + 256; };
  while (stages > 2 && smem_of(stages) > smem_cap) --stages;
  auto kern = gemm_grouped_wgrad_kernel<BN>;
  static dtc::SmemOptIn configured;  // per device
  DTC_CHECK_CUDA(dtc::ensure_dyn_smem(kern, 113 * 1024, configured));
  const int total_tiles = tiles * splits;
  const int max_ctas = 2 * num_sms();
  const int grid = total_tiles < max_ctas ? total_tiles : max_ctas;
  DTC_CHECK_CUDA(dtc::launch(kern, grid, PG_THREADS, smem_of(stages), stream, grp, K, splits, kb_per_split, stages, alpha));
  DTC_CHECK_LAUNCH();
  return 0;
}

// Same contract as dtc_gemm_wgrad_grouped, computed by CTA pairs on 256 x 256 tiles (gemm_grouped_wgrad_2cta_kernel).
// `splits` slices the reduction so that (number of 256 x 256 tiles) x splits fills the 74 CTA pairs about twice.
extern "C" int dtc_gemm_wgrad_grouped_pair(int count, const unsigned long long* A, const long long* lda,
                                           const unsigned long long* B, const long long* ldb, const int* Ms, const int* Ns,
                                           int K, const unsigned long long* out, const long long* ldo, float alpha,
                                           int splits, int stages, cudaStream_t stream) {
  DTC_CHECK_ARG(count >= 1 && count <= GG_MAX && K > 0, "wgrad_grouped_pair: count=%d (1..%d), K=%d", count, GG_MAX, K);
  constexpr int BN = 256;
  GemmGroup grp;
  memset(&grp, 0, sizeof(grp));
  grp.count = count;
  int tiles = 0;
  for (int p = 0; p < count; ++p) {
    DTC_CHECK_ARG(Ms[p] > 0 && Ns[p] > 0 && Ms[p] % 8 == 0 && Ns[p] % 8 == 0 && lda[p] % 8 == 0 && ldb[p] % 8 == 0 &&
                      ldo[p] % 4 == 0,
                  "wgrad_grouped_pair: problem %d: M=%d N=%d lda=%lld ldb=%lld ldo=%lld", p, Ms[p], Ns[p], lda[p], ldb[p],
                  ldo[p]);
    DTC_CHECK_ARG((A[p] & 15) == 0 && (B[p] & 15) == 0 && (out[p] & 15) == 0 && A[p] && B[p] && out[p],
                  "wgrad_grouped_pair: problem %d: pointers must be non-null and 16-byte aligned", p);
    int rc;
    if ((rc = make_tmap_2d_bf16(&grp.a[p], reinterpret_cast<const void*>(A[p]), K, Ms[p], lda[p], BK, 64,
                                CU_TENSOR_MAP_SWIZZLE_128B)))
      return rc;
    if ((rc = make_tmap_2d_bf16(&grp.b[p], reinterpret_cast<const void*>(B[p]), K, Ns[p], ldb[p], BK, 64,
                                CU_TENSOR_MAP_SWIZZLE_128B)))
      return rc;
    // (box rows clamped to the tensor: a tensor-parallel shard can have fewer than 128 rows; the slab's first rows are the
    // valid ones, and the second CTA of the pair then addresses a box that lies entirely outside and is dropped)
    if ((rc = make_tmap_2d_f32(&grp.o[p], reinterpret_cast<const void*>(out[p]), Ms[p], Ns[p], ldo[p],
                               Ms[p] < BM ? Ms[p] : BM, 32, CU_TENSOR_MAP_SWIZZLE_128B)))
      return rc;
    grp.out[p] = reinterpret_cast<float*>(out[p]);
    grp.ldo[p] = ldo[p];
    grp.M[p] = Ms[p];
    grp.N[p] = Ns[p];
    grp.m_blocks[p] = (Ms[p] + 255) / 256;
    grp.tile_start[p] = tiles;
    tiles += grp.m_blocks[p] * ((Ns[p] + BN - 1) / BN);
  }
  for (int p = count; p <= GG_MAX; ++p) grp.tile_start[p] = tiles;
  const int total_kb = (K + BK - 1) / BK;
  if (splits < 1) splits = 1;
  if (splits > total_kb) splits = total_kb;
  const int kb_per_split = (total_kb + splits - 1) / splits;
  splits = (total_kb + kb_per_split - 1) / kb_per_split;
  if (stages < 2 || stages > PG_STAGE_MAX) stages = PG_STAGE_MAX;
  auto smem_of = [](int st) { return st * (BM * BK * 2 + (BN / 2) * BK * 2) + 4 * GW_SLAB_BYTES + 1024 + 512; };
  while (stages > 2 && smem_of(stages) > 227 * 1024) --stages;
  auto kern = gemm_grouped_wgrad_2cta_kernel<BN>;
  static dtc::SmemOptIn configured;  // per device
  DTC_CHECK_CUDA(dtc::ensure_dyn_smem(kern, 227 * 1024, configured));
  const int total_tiles = tiles * splits;
  const int max_pairs = num_sms() / 2;
  const int pairs = total_tiles < max_pairs ? total_tiles : max_pairs;
  DTC_CHECK_CUDA(dtc::launch(kern, 2 * pairs, PG_THREADS, smem_of(stages), stream, grp, K, splits, kb_per_split, stages, alpha));
  DTC_CHECK_LAUNCH();
  return 0;
}

#ifdef DTC_GEMM_TRACE
// debug builds only: trace buffer of the pair kernel (>= 16 + 16 * 4096 zeroed 64-bit words), nullptr switches it off
extern "C" int dtc_gemm_pair_set_trace(unsigned long long* buf) {
  DTC_CHECK_CUDA(cudaMemcpyToSymbol(dtc::g_pair_trace, &buf, sizeof(buf)));
  return 0;
}
#endif
