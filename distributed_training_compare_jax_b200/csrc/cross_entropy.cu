// Softmax cross-entropy with integer labels over (optionally vocab-sharded) bf16 logits
// (op site K12 + collective X7, reference: train/create_train_step.py:32-34,148; vocab sharding:
// parallel/sharding.py:33-39).  loss_i = logsumexp(logits_i) - logits_i[y_i]; the mean over all
// tokens and its gradient (softmax - onehot) / n_tokens are produced without an fp32 logits tensor:
// the bf16 logits buffer is overwritten in place by dlogits.
//
// Three entry points share the row scan:
//   dtc_ce_fwd_bwd  single shard: one CTA per row does stats + gradient back-to-back (row stays in L1/L2)
//   dtc_ce_stats    vocab-parallel phase 1: per-row (max, sum-exp, target-logit) of the local shard
//   dtc_ce_grad     vocab-parallel phase 2: given the global LSE per row, write dlogits of the local shard
#include "common.cuh"

namespace dtc {

constexpr int CE_THREADS = 256;

struct RowStat {
  float m, s;
};
DTC_DEVICE RowStat stat_merge(RowStat a, RowStat b) {
  const float m = fmaxf(a.m, b.m);
  RowStat r;
  r.m = m;
  r.s = (a.m == -INFINITY ? 0.f : a.s * __expf(a.m - m)) + (b.m == -INFINITY ? 0.f : b.s * __expf(b.m - m));
  return r;
}

// warp then CTA reduction of (m, s) pairs; every thread returns the merged statistics (s_red: 34 floats of shared memory)
DTC_DEVICE RowStat block_merge_stats(RowStat st, float* s_red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    RowStat other;
    other.m = __shfl_xor_sync(0xffffffffu, st.m, o);
    other.s = __shfl_xor_sync(0xffffffffu, st.s, o);
    st = stat_merge(st, other);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    s_red[warp * 2] = st.m;
    s_red[warp * 2 + 1] = st.s;
  }
  __syncthreads();
  if (warp == 0) {
    RowStat t;
    t.m = lane < CE_THREADS / 32 ? s_red[lane * 2] : -INFINITY;
    t.s = lane < CE_THREADS / 32 ? s_red[lane * 2 + 1] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      RowStat other;
      other.m = __shfl_xor_sync(0xffffffffu, t.m, o);
      other.s = __shfl_xor_sync(0xffffffffu, t.s, o);
      t = stat_merge(t, other);
    }
    if (lane == 0) {
      s_red[32] = t.m;
      s_red[33] = t.s;
    }
  }
  __syncthreads();
  st.m = s_red[32];
  st.s = s_red[33];
  return st;
}

DTC_DEVICE RowStat block_row_stats(const bf16* __restrict__ row, int V, float* s_red) {
  RowStat st;
  st.m = -INFINITY;
  st.s = 0.f;
  const int V8 = V & ~7;
  for (int c = threadIdx.x * 8; c < V8; c += CE_THREADS * 8) {
    const uint4 u = *reinterpret_cast<const uint4*>(row + c);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    float f[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 t = unpack_bf16x2(w[j]);
      f[2 * j] = t.x;
      f[2 * j + 1] = t.y;
    }
    float cm = f[0];
#pragma unroll
    for (int j = 1; j < 8; ++j) cm = fmaxf(cm, f[j]);
    if (cm > st.m) {
      st.s = (st.m == -INFINITY) ? 0.f : st.s * __expf(st.m - cm);
      st.m = cm;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) st.s += __expf(f[j] - st.m);
  }
  for (int c = V8 + threadIdx.x; c < V; c += CE_THREADS) {
    const float f = __bfloat162float(row[c]);
    if (f > st.m) {
      st.s = (st.m == -INFINITY) ? 0.f : st.s * __expf(st.m - f);
      st.m = f;
    }
    st.s += __expf(f - st.m);
  }
  return block_merge_stats(st, s_red);
}

// dlogits[c] = (exp(x - lse) - [c == y_local]) * gscale for c < V ; 0 for pad columns [V, ld)
DTC_DEVICE void block_row_grad(bf16* __restrict__ row, int V, int ld, float lse, int y_local, float gscale) {
  const int V8 = V & ~7;
  for (int c = threadIdx.x * 8; c < V8; c += CE_THREADS * 8) {
    const uint4 u = *reinterpret_cast<const uint4*>(row + c);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 t = unpack_bf16x2(w[j]);
      float p0 = __expf(t.x - lse), p1 = __expf(t.y - lse);
      if (c + 2 * j == y_local) p0 -= 1.f;
      if (c + 2 * j + 1 == y_local) p1 -= 1.f;
      o[j] = pack_bf16x2(p0 * gscale, p1 * gscale);
    }
    *reinterpret_cast<uint4*>(row + c) = make_uint4(o[0], o[1], o[2], o[3]);
  }
  for (int c = V8 + threadIdx.x; c < ld; c += CE_THREADS) {
    float p = 0.f;
    if (c < V) {
      p = __expf(__bfloat162float(row[c]) - lse);
      if (c == y_local) p -= 1.f;
      p *= gscale;
    }
    row[c] = __float2bfloat16(p);
  }
}

__global__ void __launch_bounds__(CE_THREADS)
ce_fwd_bwd_kernel(bf16* __restrict__ logits, long long ld, const int* __restrict__ labels, float* __restrict__ loss_rows,
                  int V, float gscale) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  __shared__ float s_red[34];
  bf16* row = logits + (long long)blockIdx.x * ld;
  int y = labels[blockIdx.x];
  y = y < 0 ? 0 : (y >= V ? V - 1 : y);
  const float target = __bfloat162float(row[y]);
  const RowStat st = block_row_stats(row, V, s_red);
  const float lse = st.m + __logf(st.s);
  if (threadIdx.x == 0) loss_rows[blockIdx.x] = lse - target;
  __syncthreads();  // all reads of row[y] / stats done before the in-place overwrite
  block_row_grad(row, V, (int)ld, lse, y, gscale);
}

// Loss + gradient from per-part statistics that the lm_head GEMM epilogue already produced (EPI_STORE_STATS):
// parts[row * n_parts + p] = {max, sum exp(x - max)} of part p of the row.  One pass over the logits instead of two.
__global__ void __launch_bounds__(CE_THREADS)
ce_bwd_from_parts_kernel(bf16* __restrict__ logits, long long ld, const int* __restrict__ labels,
                         const float2* __restrict__ parts, int n_parts, float* __restrict__ loss_rows, int V, float gscale) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s_red[34];
  bf16* row = logits + (long long)blockIdx.x * ld;
  int y = labels[blockIdx.x];
  y = y < 0 ? 0 : (y >= V ? V - 1 : y);
  const float target = __bfloat162float(row[y]);
  RowStat st;
  st.m = -INFINITY;
  st.s = 0.f;
  for (int p = threadIdx.x; p < n_parts; p += CE_THREADS) {
    const float2 v = parts[(long long)blockIdx.x * n_parts + p];
    RowStat o;
    o.m = v.x;
    o.s = v.y;
    st = stat_merge(st, o);
  }
  st = block_merge_stats(st, s_red);
  const float lse = st.m + __logf(st.s);
  if (threadIdx.x == 0) loss_rows[blockIdx.x] = lse - target;
  __syncthreads();  // the read of row[y] is done before the in-place overwrite
  block_row_grad(row, V, (int)ld, lse, y, gscale);
}

// stats[row] = {max, sumexp(relative to max), target logit or 0 if the label is not in this shard}
__global__ void __launch_bounds__(CE_THREADS)
ce_stats_kernel(const bf16* __restrict__ logits, long long ld, const int* __restrict__ labels, float* __restrict__ stats,
                int V_local, int vocab_start) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  __shared__ float s_red[34];
  const bf16* row = logits + (long long)blockIdx.x * ld;
  const RowStat st = block_row_stats(row, V_local, s_red);
  if (threadIdx.x == 0) {
    const int y = labels[blockIdx.x] - vocab_start;
    stats[blockIdx.x * 3 + 0] = st.m;
    stats[blockIdx.x * 3 + 1] = st.s;
    stats[blockIdx.x * 3 + 2] = (y >= 0 && y < V_local) ? __bfloat162float(row[y]) : 0.f;
  }
}

// Combine per-shard stats [n_shards, M, 3] into lse[M] and loss_rows[M].
__global__ void ce_combine_kernel(const float* __restrict__ stats, int n_shards, int M, float* __restrict__ lse,
                                  float* __restrict__ loss_rows) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M) return;
  float m = -INFINITY, s = 0.f, tgt = 0.f;
  for (int r = 0; r < n_shards; ++r) m = fmaxf(m, stats[((long long)r * M + i) * 3]);
  for (int r = 0; r < n_shards; ++r) {
    const float* p = stats + ((long long)r * M + i) * 3;
    s += p[1] * __expf(p[0] - m);
    tgt += p[2];
  }
  const float l = m + __logf(s);
  lse[i] = l;
  loss_rows[i] = l - tgt;
}

__global__ void __launch_bounds__(CE_THREADS)
ce_grad_kernel(bf16* __restrict__ logits, long long ld, const int* __restrict__ labels, const float* __restrict__ lse,
               int V_local, int vocab_start, float gscale) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  bf16* row = logits + (long long)blockIdx.x * ld;
  block_row_grad(row, V_local, (int)ld, lse[blockIdx.x], labels[blockIdx.x] - vocab_start, gscale);
}

// out[0] (+)= scale * sum(in[0..n))   — deterministic single-CTA tree reduction
__global__ void __launch_bounds__(1024) reduce_sum_kernel(const float* __restrict__ in, int n, float scale,
                                                          float* __restrict__ out, int accumulate) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  __shared__ float s[32];
  float v = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) v += in[i];
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    v = s[threadIdx.x];
    v = warp_sum(v);
    if (threadIdx.x == 0) {
      if (accumulate) atomicAdd(out, v * scale);  // several chains (streams) may accumulate into the same scalar
      else out[0] = v * scale;
    }
  }
}

}  // namespace dtc

using namespace dtc;

// logits [M, ld] bf16 (V valid columns, ld >= V, ld % 8 == 0) are overwritten by dlogits = dloss/dlogits
// for loss = gscale * sum_rows(loss_row).  loss_rows [M] fp32 receives the unscaled per-row loss.
extern "C" int dtc_ce_fwd_bwd(void* logits, long long ld, const int* labels, float* loss_rows, int M, int V, float gscale,
                              cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && V > 0 && ld >= V && ld % 8 == 0, "ce_fwd_bwd: bad shape M=%d V=%d ld=%lld", M, V, ld);
  DTC_CHECK_CUDA(dtc::launch(ce_fwd_bwd_kernel, M, CE_THREADS, 0, stream, reinterpret_cast<bf16*>(logits), ld, labels, loss_rows, V, gscale));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_ce_stats(const void* logits, long long ld, const int* labels, float* stats, int M, int V_local,
                            int vocab_start, cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && V_local > 0 && ld >= V_local && ld % 8 == 0, "ce_stats: bad shape");
  DTC_CHECK_CUDA(dtc::launch(ce_stats_kernel, M, CE_THREADS, 0, stream, reinterpret_cast<const bf16*>(logits), ld, labels, stats, V_local,
                                                vocab_start));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_ce_combine(const float* stats, int n_shards, int M, float* lse, float* loss_rows, cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && n_shards > 0, "ce_combine: bad shape");
  DTC_CHECK_CUDA(dtc::launch(ce_combine_kernel, (M + 255) / 256, 256, 0, stream, stats, n_shards, M, lse, loss_rows));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_ce_grad(void* logits, long long ld, const int* labels, const float* lse, int M, int V_local,
                           int vocab_start, float gscale, cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && V_local > 0 && ld >= V_local && ld % 8 == 0, "ce_grad: bad shape");
  DTC_CHECK_CUDA(dtc::launch(ce_grad_kernel, M, CE_THREADS, 0, stream, reinterpret_cast<bf16*>(logits), ld, labels, lse, V_local, vocab_start,
                                               gscale));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_reduce_sum(const float* in, int n, float scale, float* out, int accumulate, cudaStream_t stream) {
  DTC_CHECK_ARG(n > 0, "reduce_sum: empty input");
  DTC_CHECK_CUDA(dtc::launch(reduce_sum_kernel, 1, 1024, 0, stream, in, n, scale, out, accumulate));
  DTC_CHECK_LAUNCH();
  return 0;
}

// Loss and in-place gradient from the per-part statistics written by the lm_head GEMM (epilogue mode 8, see dtc_gemm_bf16_ex):
// parts = float2 [M, n_parts].  Same outputs as dtc_ce_fwd_bwd with one pass over the logits less.
extern "C" int dtc_ce_bwd_from_parts(void* logits, long long ld, const int* labels, const void* parts, int n_parts,
                                     float* loss_rows, int M, int V, float gscale, cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && V > 0 && ld >= V && ld % 8 == 0 && n_parts > 0 && parts != nullptr,
                "ce_bwd_from_parts: bad shape M=%d V=%d ld=%lld n_parts=%d", M, V, ld, n_parts);
  DTC_CHECK_CUDA(dtc::launch(ce_bwd_from_parts_kernel, M, CE_THREADS, 0, stream, reinterpret_cast<bf16*>(logits), ld, labels,
                             reinterpret_cast<const float2*>(parts), n_parts, loss_rows, V, gscale));
  DTC_CHECK_LAUNCH();
  return 0;
}
