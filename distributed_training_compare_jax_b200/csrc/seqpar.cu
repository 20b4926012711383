// Sequence-parallel tensor parallelism: the LayerNorm kernels that END a reduce-scatter and START an all-gather
// (collectives X3 / X4 / X5 / X8 of §2.5 — the all-reduces GSPMD derives from the
// reference's replicated-activation constraints, model/CausalSelfAttention.py:49-50 and model/MLP.py:23-24, and their
// transposes in the backward pass).
//
// With n tensor-parallel ranks every rank owns Ms = M/n rows of the residual stream.  A row-parallel GEMM (out_proj,
// fc2) or the dgrad of a column-parallel one (q/k/v, fc1, lm_head) produces a PARTIAL [M, D] result on every rank; its
// epilogue (EPI_SCATTER_BF16, gemm.cu) stores each row block straight into the owner's partial-sum slot over NVLink.
// After one flag barrier the owner runs ONE kernel from this file on its Ms rows:
//
//   dtc_sp_ln_fwd   x = residual + bias + sum of the n partial slots (fp32)  -> new residual rows (bf16, local)
//                   y = LayerNorm(x)                                          -> pushed into EVERY rank's [M, D] operand
//                                                                                buffer (multimem.st through the switch:
//                                                                                the all-gather in front of the next
//                                                                                column-parallel GEMM)
//   dtc_sp_ln_bwd   dy = sum of the n partial slots (fp32); dx = LayerNorm'(dy) + residual gradient
//                   -> pushed into every rank's [M, D] gradient buffer (all-gather in front of the next dgrad / wgrad
//                   GEMMs); dgamma / dbeta accumulate over the OWNED rows only (summed over the TP group once per step)
//
// so the LayerNorm / residual / bias work is divided by n, nothing is reduced twice, and no collective is a separate
// pass over the data: reduce-scatter = GEMM epilogue stores + the loads of this kernel, all-gather = its stores.
// Without a multicast address (no NVLS, or the single-GPU unit test) the push is a loop of peer stores.
#include "common.cuh"
#include "spsync.cuh"

namespace dtc {

constexpr int SP_WARPS = 8;      // backward: rows per CTA (its column partials are reduced per CTA)
constexpr int SPF_WARPS = 4;     // forward: rows per CTA — at TP = 8 a rank owns 512 rows: 128 CTAs instead of 64

struct SpPeers {
  bf16* p[SP_MAX_RANKS];
};

DTC_DEVICE void sp_push16(bf16* mc, const SpPeers& peers, int n_out, long long off, uint4 v) {
  if (mc != nullptr) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + off), "f"(__uint_as_float(v.x)),
                 "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                 : "memory");
  } else {
#pragma unroll
    for (int i = 0; i < SP_MAX_RANKS; ++i)
      if (i < n_out) *reinterpret_cast<uint4*>(peers.p[i] + off) = v;
  }
}

DTC_DEVICE void add_bf16x8(float* acc, uint4 u) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 f = unpack_bf16x2(w[j]);
    acc[2 * j] += f.x;
    acc[2 * j + 1] += f.y;
  }
}

template <bool L2_COHERENT>
DTC_DEVICE uint4 sp_load16(const bf16* p) {
  // partial-sum slots written by peers while THIS kernel may already be resident must not come from a non-coherent path
  if (L2_COHERENT) return __ldcg(reinterpret_cast<const uint4*>(p));
  return *reinterpret_cast<const uint4*>(p);
}

// acc += sum over the n_part (<= SP_MAX_RANKS) partial-sum slots of one 16-byte chunk.  All loads are issued before the first
// add: a load -> add loop over a run-time slot count serialises n_part DRAM / NVLink round trips per chunk (ncu: 60 % of the
// stall samples of sp_ln_fwd were long-scoreboard waits on exactly that loop at TP = 8).
template <bool L2_COHERENT>
DTC_DEVICE void add_slots_bf16x8(float* acc, const bf16* part, int n_part, long long part_stride, long long off) {
  uint4 v[SP_MAX_RANKS];
#pragma unroll
  for (int p = 0; p < SP_MAX_RANKS; ++p)
    if (p < n_part) v[p] = sp_load16<L2_COHERENT>(part + p * part_stride + off);
#pragma unroll
  for (int p = 0; p < SP_MAX_RANKS; ++p)
    if (p < n_part) add_bf16x8(acc, v[p]);
}

// one warp per owned row; D = NCH * 256, the row lives in registers (shared by both kernel variants)
template <int NCH, bool L2_COHERENT>
DTC_DEVICE void sp_ln_fwd_row(const bf16* __restrict__ res, const bf16* __restrict__ bias, const bf16* __restrict__ part,
                              int n_part, long long part_stride, const bf16* __restrict__ gamma,
                              const bf16* __restrict__ beta, bf16* __restrict__ hsum, float* __restrict__ mean_out,
                              float* __restrict__ rstd_out, bf16* mc_out, const SpPeers& peers, int n_out,
                              long long out_row0, float eps, long long row, int lane) {
  constexpr int D = NCH * 256;
  float x[NCH * 8];
  float s = 0.f, ss = 0.f;
#pragma unroll
  for (int k = 0; k < NCH; ++k) {
    const int c = k * 256 + lane * 8;
    const long long off = row * D + c;
    float* xk = x + k * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) xk[j] = 0.f;
    add_bf16x8(xk, *reinterpret_cast<const uint4*>(res + off));
    if (bias != nullptr) add_bf16x8(xk, *reinterpret_cast<const uint4*>(bias + c));
    add_slots_bf16x8<L2_COHERENT>(xk, part, n_part, part_stride, off);
    // the residual stream is stored in bf16: normalise exactly what the backward pass will read back
    const uint4 packed = make_uint4(pack_bf16x2(xk[0], xk[1]), pack_bf16x2(xk[2], xk[3]), pack_bf16x2(xk[4], xk[5]),
                                    pack_bf16x2(xk[6], xk[7]));
    if (hsum != nullptr) *reinterpret_cast<uint4*>(hsum + off) = packed;
    const uint32_t w[4] = {packed.x, packed.y, packed.z, packed.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(w[j]);
      xk[2 * j] = f.x;
      xk[2 * j + 1] = f.y;
      s += f.x + f.y;
      ss += f.x * f.x + f.y * f.y;
    }
  }
  s = warp_sum(s);
  ss = warp_sum(ss);
  const float mean = s / D;
  const float var = fmaxf(ss / D - mean * mean, 0.f);
  const float rstd = rsqrtf(var + eps);
  if (lane == 0) {
    mean_out[row] = mean;
    rstd_out[row] = rstd;
  }
#pragma unroll
  for (int k = 0; k < NCH; ++k) {
    const int c = k * 256 + lane * 8;
    const uint4 g = *reinterpret_cast<const uint4*>(gamma + c);
    const uint4 b = *reinterpret_cast<const uint4*>(beta + c);
    const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, bw[4] = {b.x, b.y, b.z, b.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 gg = unpack_bf16x2(gw[j]), bb = unpack_bf16x2(bw[j]);
      o[j] = pack_bf16x2((x[k * 8 + 2 * j] - mean) * rstd * gg.x + bb.x, (x[k * 8 + 2 * j + 1] - mean) * rstd * gg.y + bb.y);
    }
    sp_push16(mc_out, peers, n_out, (out_row0 + row) * D + c, make_uint4(o[0], o[1], o[2], o[3]));
  }
}

// Barrier-kernel variant (default): the partial-sum slots were completed before this kernel started (flag barrier in
// front of it), so the plain read-only load path is safe.
template <int NCH>
__global__ void __launch_bounds__(SPF_WARPS * 32)
sp_ln_fwd_kernel(const bf16* __restrict__ res, const bf16* __restrict__ bias, const bf16* __restrict__ part, int n_part,
                 long long part_stride, const bf16* __restrict__ gamma, const bf16* __restrict__ beta,
                 bf16* __restrict__ hsum, float* __restrict__ mean_out, float* __restrict__ rstd_out, bf16* mc_out,
                 SpPeers peers, int n_out, long long out_row0, int Ms, float eps) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * SPF_WARPS + warp;
  if (row >= Ms) return;
  sp_ln_fwd_row<NCH, false>(res, bias, part, n_part, part_stride, gamma, beta, hsum, mean_out, rstd_out, mc_out, peers,
                            n_out, out_row0, eps, row, lane);
}

// Folded-barrier variant: optionally waits for the partial-sum slots of every rank in its prologue (wait_in) and
// publishes its own pushes to every rank from its last CTA (signal_out) — no separate barrier kernels around it.
template <int NCH>
__global__ void __launch_bounds__(SPF_WARPS * 32)
sp_ln_fwd_sync_kernel(const bf16* __restrict__ res, const bf16* __restrict__ bias, const bf16* __restrict__ part, int n_part,
                      long long part_stride, const bf16* __restrict__ gamma, const bf16* __restrict__ beta,
                      bf16* __restrict__ hsum, float* __restrict__ mean_out, float* __restrict__ rstd_out, bf16* mc_out,
                      SpPeers peers, int n_out, long long out_row0, int Ms, float eps, SpSync sync, int wait_in,
                      int signal_out) {
  pdl_launch_dependents();
  pdl_wait();
  if (wait_in) {
    sp_wait_signals(sync, threadIdx.x);
    __syncthreads();
    if (threadIdx.x == 0) sp_wait_done(sync, gridDim.x);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * SPF_WARPS + warp;
  if (row < Ms)
    sp_ln_fwd_row<NCH, true>(res, bias, part, n_part, part_stride, gamma, beta, hsum, mean_out, rstd_out, mc_out, peers,
                             n_out, out_row0, eps, row, lane);
  if (signal_out) {
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) sp_signal(sync, gridDim.x);
  }
}

template <int NCH, bool L2_COHERENT>
DTC_DEVICE void sp_ln_bwd_body(const bf16* __restrict__ part, int n_part, long long part_stride, const bf16* __restrict__ x,
                               const float* __restrict__ mean, const float* __restrict__ rstd,
                               const bf16* __restrict__ gamma, const bf16* __restrict__ dres, float* __restrict__ dgamma,
                               float* __restrict__ dbeta, bf16* mc_out, const SpPeers& peers, int n_out, long long out_row0,
                               int Ms) {
  constexpr int D = NCH * 256;
  // D <= 512: one [D] row of column partials per warp, summed by column after the row loop (plain 16-byte stores; the
  // shared-memory atomics of the general path hit the same banks 8 ways with the lane * 8 + j layout)
  constexpr bool WARP_ROWS = NCH <= 2;
  constexpr int S_ROWS = WARP_ROWS ? SP_WARPS : 1;
  __shared__ __align__(16) float s_dg[S_ROWS * D];
  __shared__ __align__(16) float s_db[S_ROWS * D];
  if constexpr (!WARP_ROWS) {
    for (int i = threadIdx.x; i < D; i += blockDim.x) {
      s_dg[i] = 0.f;
      s_db[i] = 0.f;
    }
    __syncthreads();
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float acc_dg[NCH * 8], acc_db[NCH * 8], gam[NCH * 8];
#pragma unroll
  for (int k = 0; k < NCH; ++k) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      acc_dg[k * 8 + j] = 0.f;
      acc_db[k * 8 + j] = 0.f;
      gam[k * 8 + j] = 0.f;
    }
    add_bf16x8(gam + k * 8, *reinterpret_cast<const uint4*>(gamma + k * 256 + lane * 8));
  }
  for (long long row = (long long)blockIdx.x * SP_WARPS + warp; row < Ms; row += (long long)gridDim.x * SP_WARPS) {
    const float mu = mean[row], rs = rstd[row];
    float xhat[NCH * 8], wdy[NCH * 8];
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      const long long off = row * D + k * 256 + lane * 8;
      float dy[8], xv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        dy[j] = 0.f;
        xv[j] = 0.f;
      }
      add_slots_bf16x8<L2_COHERENT>(dy, part, n_part, part_stride, off);
      add_bf16x8(xv, *reinterpret_cast<const uint4*>(x + off));
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int i = k * 8 + j;
        xhat[i] = (xv[j] - mu) * rs;
        acc_dg[i] += dy[j] * xhat[i];
        acc_db[i] += dy[j];
        wdy[i] = dy[j] * gam[i];
        c1 += wdy[i];
        c2 += wdy[i] * xhat[i];
      }
    }
    c1 = warp_sum(c1) * (1.f / D);
    c2 = warp_sum(c2) * (1.f / D);
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      const long long off = row * D + k * 256 + lane * 8;
      float r[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) r[j] = rs * (wdy[k * 8 + j] - c1 - xhat[k * 8 + j] * c2);
      if (dres != nullptr) add_bf16x8(r, *reinterpret_cast<const uint4*>(dres + off));
      sp_push16(mc_out, peers, n_out, (out_row0 + row) * D + k * 256 + lane * 8,
                make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7])));
    }
  }
  if constexpr (WARP_ROWS) {
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      float4* pg = reinterpret_cast<float4*>(s_dg + warp * D + k * 256 + lane * 8);
      float4* pb = reinterpret_cast<float4*>(s_db + warp * D + k * 256 + lane * 8);
      pg[0] = make_float4(acc_dg[k * 8], acc_dg[k * 8 + 1], acc_dg[k * 8 + 2], acc_dg[k * 8 + 3]);
      pg[1] = make_float4(acc_dg[k * 8 + 4], acc_dg[k * 8 + 5], acc_dg[k * 8 + 6], acc_dg[k * 8 + 7]);
      pb[0] = make_float4(acc_db[k * 8], acc_db[k * 8 + 1], acc_db[k * 8 + 2], acc_db[k * 8 + 3]);
      pb[1] = make_float4(acc_db[k * 8 + 4], acc_db[k * 8 + 5], acc_db[k * 8 + 6], acc_db[k * 8 + 7]);
    }
    __syncthreads();
    // one thread per 4 columns of dgamma (first D / 4 threads) or dbeta (next D / 4): sum over the warps, one 16-byte reduce-add
    for (int i = threadIdx.x; i < D / 2; i += blockDim.x) {
      const bool is_db = i >= D / 4;
      const int c = (is_db ? i - D / 4 : i) * 4;
      const float* src = (is_db ? s_db : s_dg) + c;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int w = 0; w < SP_WARPS; ++w) {
        const float4 v = *reinterpret_cast<const float4*>(src + w * D);
        a.x += v.x;
        a.y += v.y;
        a.z += v.z;
        a.w += v.w;
      }
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"((is_db ? dbeta : dgamma) + c), "f"(a.x), "f"(a.y),
                   "f"(a.z), "f"(a.w)
                   : "memory");
    }
  } else {
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        atomicAdd(&s_dg[k * 256 + lane * 8 + j], acc_dg[k * 8 + j]);
        atomicAdd(&s_db[k * 256 + lane * 8 + j], acc_db[k * 8 + j]);
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < D; i += blockDim.x) {
      atomicAdd(&dgamma[i], s_dg[i]);
      atomicAdd(&dbeta[i], s_db[i]);
    }
  }
}

// Barrier-kernel variant (default): the slots were completed before this kernel started, plain read-only loads are safe.
template <int NCH>
__global__ void __launch_bounds__(SP_WARPS * 32)
sp_ln_bwd_kernel(const bf16* __restrict__ part, int n_part, long long part_stride, const bf16* __restrict__ x,
                 const float* __restrict__ mean, const float* __restrict__ rstd, const bf16* __restrict__ gamma,
                 const bf16* __restrict__ dres, float* __restrict__ dgamma, float* __restrict__ dbeta, bf16* mc_out,
                 SpPeers peers, int n_out, long long out_row0, int Ms) {
  pdl_launch_dependents();
  pdl_wait();
  sp_ln_bwd_body<NCH, false>(part, n_part, part_stride, x, mean, rstd, gamma, dres, dgamma, dbeta, mc_out, peers, n_out,
                             out_row0, Ms);
}

// Folded-barrier variant (see sp_ln_fwd_sync_kernel).
template <int NCH>
__global__ void __launch_bounds__(SP_WARPS * 32)
sp_ln_bwd_sync_kernel(const bf16* __restrict__ part, int n_part, long long part_stride, const bf16* __restrict__ x,
                      const float* __restrict__ mean, const float* __restrict__ rstd, const bf16* __restrict__ gamma,
                      const bf16* __restrict__ dres, float* __restrict__ dgamma, float* __restrict__ dbeta, bf16* mc_out,
                      SpPeers peers, int n_out, long long out_row0, int Ms, SpSync sync, int wait_in, int signal_out) {
  pdl_launch_dependents();
  pdl_wait();
  if (wait_in) {
    sp_wait_signals(sync, threadIdx.x);
    __syncthreads();
    if (threadIdx.x == 0) sp_wait_done(sync, gridDim.x);
  }
  sp_ln_bwd_body<NCH, true>(part, n_part, part_stride, x, mean, rstd, gamma, dres, dgamma, dbeta, mc_out, peers, n_out,
                            out_row0, Ms);
  if (signal_out) {
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) sp_signal(sync, gridDim.x);
  }
}

// tiny consumer-side wait for kernels that cannot wait themselves (the stock GEMMs that read a gathered operand)
__global__ void sp_wait_kernel(SpSync sync) {
  pdl_launch_dependents();
  pdl_wait();
  sp_wait_signals(sync, threadIdx.x);
  __syncthreads();
  if (threadIdx.x == 0) sp_wait_done(sync, 1);
}

static int sp_fill_sync(SpSync* s, const unsigned long long* sig_ptrs, int* counters, int rank, int world, const char* who) {
  if (sig_ptrs == nullptr || counters == nullptr || world < 1 || world > SP_MAX_RANKS || rank < 0 || rank >= world) {
    dtc_set_error("%s: bad synchronisation block (world %d, rank %d)", who, world, rank);
    return -1;
  }
  for (int i = 0; i < SP_MAX_RANKS; ++i) s->sig[i] = i < world ? reinterpret_cast<int*>(sig_ptrs[i]) : nullptr;
  s->counters = counters;
  s->rank = rank;
  s->world = world;
  return 0;
}

static int sp_fill_peers(SpPeers* peers, unsigned long long mc_out, const unsigned long long* peer_out, int n_out,
                         const char* who) {
  for (int i = 0; i < SP_MAX_RANKS; ++i) peers->p[i] = nullptr;
  if (mc_out != 0) {
    if (mc_out & 15) {
      dtc_set_error("%s: multicast pointer must be 16-byte aligned", who);
      return -1;
    }
    return 0;
  }
  if (peer_out == nullptr || n_out < 1 || n_out > SP_MAX_RANKS) {
    dtc_set_error("%s: need a multicast pointer or 1..%d peer pointers (got %d)", who, SP_MAX_RANKS, n_out);
    return -1;
  }
  for (int i = 0; i < n_out; ++i) {
    if (peer_out[i] == 0 || (peer_out[i] & 15)) {
      dtc_set_error("%s: peer pointer %d null or unaligned", who, i);
      return -1;
    }
    peers->p[i] = reinterpret_cast<bf16*>(peer_out[i]);
  }
  return 0;
}

}  // namespace dtc

using namespace dtc;

// res [Ms, D] (own rows of the residual stream), bias [D] or null, part [n_part, Ms, D] or null (n_part = 0),
// hsum [Ms, D] or null (receives residual + bias + partial sums), mean / rstd [Ms];
// LayerNorm(hsum) is written to rows [out_row0, out_row0 + Ms) of the [*, D] destination reached through `mc_out`
// (multicast address: every rank's copy) or, when mc_out == 0, of each of the n_out `peer_out` base pointers.
extern "C" int dtc_sp_ln_fwd(const void* res, const void* bias, const void* part, int n_part, const void* gamma,
                             const void* beta, void* hsum, float* mean, float* rstd, unsigned long long mc_out,
                             const unsigned long long* peer_out, int n_out, long long out_row0, int Ms, int D, float eps,
                             cudaStream_t stream) {
  DTC_CHECK_ARG(Ms > 0 && D % 256 == 0 && D >= 256 && D <= 2048, "sp_ln_fwd: D=%d must be a multiple of 256, <= 2048", D);
  DTC_CHECK_ARG(n_part >= 0 && n_part <= SP_MAX_RANKS && (n_part == 0 || part != nullptr), "sp_ln_fwd: bad n_part %d", n_part);
  DTC_CHECK_ARG(res != nullptr && out_row0 >= 0, "sp_ln_fwd: residual rows missing");
  SpPeers peers;
  if (sp_fill_peers(&peers, mc_out, peer_out, n_out, "sp_ln_fwd")) return -1;
  const int blocks = (Ms + SPF_WARPS - 1) / SPF_WARPS;
  const long long stride = (long long)Ms * D;
#define DTC_SP_FWD(NCH)                                                                                              \
  case NCH:                                                                                                          \
    DTC_CHECK_CUDA(dtc::launch(sp_ln_fwd_kernel<NCH>, blocks, SPF_WARPS * 32, 0, stream,                              \
        reinterpret_cast<const bf16*>(res), reinterpret_cast<const bf16*>(bias), reinterpret_cast<const bf16*>(part), \
        n_part, stride, reinterpret_cast<const bf16*>(gamma), reinterpret_cast<const bf16*>(beta),                   \
        reinterpret_cast<bf16*>(hsum), mean, rstd, reinterpret_cast<bf16*>(mc_out), peers, n_out, out_row0, Ms, eps)); \
    break;
  switch (D / 256) {
    DTC_SP_FWD(1) DTC_SP_FWD(2) DTC_SP_FWD(3) DTC_SP_FWD(4) DTC_SP_FWD(5) DTC_SP_FWD(6) DTC_SP_FWD(7) DTC_SP_FWD(8)
    default: dtc_set_error("sp_ln_fwd: unsupported D=%d", D); return -1;
  }
#undef DTC_SP_FWD
  DTC_CHECK_LAUNCH();
  return 0;
}

// part [n_part, Ms, D] partial dL/dy slots, x [Ms, D] the rows that were normalised in the forward pass, dres [Ms, D]
// residual gradient or null; dgamma / dbeta (fp32 [D]) are ACCUMULATED into (sum over the owned rows only);
// dx goes to rows [out_row0, out_row0 + Ms) of the destination (see dtc_sp_ln_fwd).
extern "C" int dtc_sp_ln_bwd(const void* part, int n_part, const void* x, const float* mean, const float* rstd,
                             const void* gamma, const void* dres, float* dgamma, float* dbeta, unsigned long long mc_out,
                             const unsigned long long* peer_out, int n_out, long long out_row0, int Ms, int D,
                             cudaStream_t stream) {
  DTC_CHECK_ARG(Ms > 0 && D % 256 == 0 && D >= 256 && D <= 2048, "sp_ln_bwd: D=%d must be a multiple of 256, <= 2048", D);
  DTC_CHECK_ARG(n_part >= 1 && n_part <= SP_MAX_RANKS && part != nullptr, "sp_ln_bwd: bad n_part %d", n_part);
  DTC_CHECK_ARG((((uintptr_t)dgamma | (uintptr_t)dbeta) & 15) == 0, "sp_ln_bwd: dgamma / dbeta must be 16-byte aligned");
  SpPeers peers;
  if (sp_fill_peers(&peers, mc_out, peer_out, n_out, "sp_ln_bwd")) return -1;
  int blocks = (Ms + SP_WARPS - 1) / SP_WARPS;
  if (blocks > 148 * 2) blocks = 148 * 2;
  const long long stride = (long long)Ms * D;
#define DTC_SP_BWD(NCH)                                                                                             \
  case NCH:                                                                                                         \
    DTC_CHECK_CUDA(dtc::launch(sp_ln_bwd_kernel<NCH>, blocks, SP_WARPS * 32, 0, stream,                             \
        reinterpret_cast<const bf16*>(part), n_part, stride, reinterpret_cast<const bf16*>(x), mean, rstd,          \
        reinterpret_cast<const bf16*>(gamma), reinterpret_cast<const bf16*>(dres), dgamma, dbeta,                   \
        reinterpret_cast<bf16*>(mc_out), peers, n_out, out_row0, Ms));                                              \
    break;
  switch (D / 256) {
    DTC_SP_BWD(1) DTC_SP_BWD(2) DTC_SP_BWD(3) DTC_SP_BWD(4) DTC_SP_BWD(5) DTC_SP_BWD(6) DTC_SP_BWD(7) DTC_SP_BWD(8)
    default: dtc_set_error("sp_ln_bwd: unsupported D=%d", D); return -1;
  }
#undef DTC_SP_BWD
  DTC_CHECK_LAUNCH();
  return 0;
}

// ---- folded-barrier variants (experimental; `DTC_SP_FOLD=1`) --------------------------------------------------------------
// Same operations as dtc_sp_ln_fwd / dtc_sp_ln_bwd with the cross-rank synchronisation inside the kernel:
//   wait_in    the prologue waits until every rank has issued the signal that publishes its partial-sum slots
//   signal_out the last CTA publishes this kernel's pushes to every rank
// sig_ptrs: HOST array of `world` device addresses (every rank's int[world] flag array, symmetric memory, zero-initialised);
// counters: LOCAL int[4] device array, zero-initialised (signals issued, waits completed, two CTA ticket counters).
extern "C" int dtc_sp_ln_fwd_sync(const void* res, const void* bias, const void* part, int n_part, const void* gamma,
                                  const void* beta, void* hsum, float* mean, float* rstd, unsigned long long mc_out,
                                  const unsigned long long* peer_out, int n_out, long long out_row0, int Ms, int D,
                                  float eps, const unsigned long long* sig_ptrs, int* counters, int rank, int world,
                                  int wait_in, int signal_out, cudaStream_t stream) {
  DTC_CHECK_ARG(Ms > 0 && D % 256 == 0 && D >= 256 && D <= 2048, "sp_ln_fwd_sync: D=%d must be a multiple of 256, <= 2048", D);
  DTC_CHECK_ARG(n_part >= 0 && n_part <= SP_MAX_RANKS && (n_part == 0 || part != nullptr), "sp_ln_fwd_sync: bad n_part %d", n_part);
  DTC_CHECK_ARG(res != nullptr && out_row0 >= 0, "sp_ln_fwd_sync: residual rows missing");
  SpPeers peers;
  SpSync sync;
  if (sp_fill_peers(&peers, mc_out, peer_out, n_out, "sp_ln_fwd_sync")) return -1;
  if (sp_fill_sync(&sync, sig_ptrs, counters, rank, world, "sp_ln_fwd_sync")) return -1;
  const int blocks = (Ms + SPF_WARPS - 1) / SPF_WARPS;
  const long long stride = (long long)Ms * D;
#define DTC_SP_FWD(NCH)                                                                                              \
  case NCH:                                                                                                          \
    DTC_CHECK_CUDA(dtc::launch(sp_ln_fwd_sync_kernel<NCH>, blocks, SPF_WARPS * 32, 0, stream,                         \
        reinterpret_cast<const bf16*>(res), reinterpret_cast<const bf16*>(bias), reinterpret_cast<const bf16*>(part), \
        n_part, stride, reinterpret_cast<const bf16*>(gamma), reinterpret_cast<const bf16*>(beta),                   \
        reinterpret_cast<bf16*>(hsum), mean, rstd, reinterpret_cast<bf16*>(mc_out), peers, n_out, out_row0, Ms, eps,  \
        sync, wait_in, signal_out));                                                                                  \
    break;
  switch (D / 256) {
    DTC_SP_FWD(1) DTC_SP_FWD(2) DTC_SP_FWD(3) DTC_SP_FWD(4) DTC_SP_FWD(5) DTC_SP_FWD(6) DTC_SP_FWD(7) DTC_SP_FWD(8)
    default: dtc_set_error("sp_ln_fwd_sync: unsupported D=%d", D); return -1;
  }
#undef DTC_SP_FWD
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_sp_ln_bwd_sync(const void* part, int n_part, const void* x, const float* mean, const float* rstd,
                                  const void* gamma, const void* dres, float* dgamma, float* dbeta,
                                  unsigned long long mc_out, const unsigned long long* peer_out, int n_out,
                                  long long out_row0, int Ms, int D, const unsigned long long* sig_ptrs, int* counters,
                                  int rank, int world, int wait_in, int signal_out, cudaStream_t stream) {
  DTC_CHECK_ARG(Ms > 0 && D % 256 == 0 && D >= 256 && D <= 2048, "sp_ln_bwd_sync: D=%d must be a multiple of 256, <= 2048", D);
  DTC_CHECK_ARG(n_part >= 1 && n_part <= SP_MAX_RANKS && part != nullptr, "sp_ln_bwd_sync: bad n_part %d", n_part);
  DTC_CHECK_ARG((((uintptr_t)dgamma | (uintptr_t)dbeta) & 15) == 0, "sp_ln_bwd_sync: dgamma / dbeta must be 16-byte aligned");
  SpPeers peers;
  SpSync sync;
  if (sp_fill_peers(&peers, mc_out, peer_out, n_out, "sp_ln_bwd_sync")) return -1;
  if (sp_fill_sync(&sync, sig_ptrs, counters, rank, world, "sp_ln_bwd_sync")) return -1;
  int blocks = (Ms + SP_WARPS - 1) / SP_WARPS;
  if (blocks > 148 * 2) blocks = 148 * 2;
  const long long stride = (long long)Ms * D;
#define DTC_SP_BWD(NCH)                                                                                             \
  case NCH:                                                                                                         \
    DTC_CHECK_CUDA(dtc::launch(sp_ln_bwd_sync_kernel<NCH>, blocks, SP_WARPS * 32, 0, stream,                        \
        reinterpret_cast<const bf16*>(part), n_part, stride, reinterpret_cast<const bf16*>(x), mean, rstd,          \
        reinterpret_cast<const bf16*>(gamma), reinterpret_cast<const bf16*>(dres), dgamma, dbeta,                   \
        reinterpret_cast<bf16*>(mc_out), peers, n_out, out_row0, Ms, sync, wait_in, signal_out));                   \
    break;
  switch (D / 256) {
    DTC_SP_BWD(1) DTC_SP_BWD(2) DTC_SP_BWD(3) DTC_SP_BWD(4) DTC_SP_BWD(5) DTC_SP_BWD(6) DTC_SP_BWD(7) DTC_SP_BWD(8)
    default: dtc_set_error("sp_ln_bwd_sync: unsupported D=%d", D); return -1;
  }
#undef DTC_SP_BWD
  DTC_CHECK_LAUNCH();
  return 0;
}

// Waits (one tiny CTA, local polling only) until every rank has issued its next signal: in front of a stock GEMM that
// reads an operand gathered by dtc_sp_ln_*_sync(signal_out = 1).
extern "C" int dtc_sp_wait(const unsigned long long* sig_ptrs, int* counters, int rank, int world, cudaStream_t stream) {
  SpSync sync;
  if (sp_fill_sync(&sync, sig_ptrs, counters, rank, world, "sp_wait")) return -1;
  DTC_CHECK_CUDA(dtc::launch(sp_wait_kernel, 1, 32, 0, stream, sync));
  DTC_CHECK_LAUNCH();
  return 0;
}
