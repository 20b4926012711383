// Symmetric-memory communication kernels over NVLink 5 / NVSwitch (components R10 / R11 / R13; collectives X1 / X2
// of SURVEY.md §2.5 — the gradient all-reduce GSPMD inserts for the reference's replicated parameters,
// parallel/sharding.py:25-27 + train/create_train_step.py:34-40 — and the cross-rank barriers the fused
// GEMM + all-reduce epilogues of gemm.cu rely on).
//
// Every rank maps every peer's buffers (torch symmetric memory: CUDA VMM handles exchanged at rendezvous) and,
// when the fabric supports it, a MULTICAST address bound to all ranks' copies.  The data-parallel optimizer
// step then is three launches instead of {NCCL all-reduce, norm, AdamW}:
//
//   dtc_dp_reduce_scatter   each rank owns 1/n of the flat fp32 gradient: `multimem.ld_reduce.add.v4.f32` makes
//                           the switch sum all ranks' copies (NVLS; peer loads when there is no multicast), the
//                           sum lands in the owner's buffer and its squares are accumulated per CTA
//   dtc_dp_norm_exchange    one CTA: local sum of squares -> written into every peer's slot (P2P stores) -> flag
//                           barrier -> global norm, clip factor, step += 1 (no host sync, graph-capturable)
//   dtc_dp_adamw_allgather  AdamW on the owned shard only (optimizer traffic / n), new bf16 parameters pushed to
//                           ALL ranks with `multimem.st` (all-gather through the switch), whole local gradient
//                           buffer zeroed for the next step
//   dtc_symm_barrier        flag barrier (one int per ordered rank pair, CAS 0->1 / wait 1 -> reset 0)
//
// Flag waits are bounded: a protocol error traps instead of hanging the node.
#include "common.cuh"

namespace dtc {

constexpr int MAX_RANKS = 16;

struct PeerPtrs {
  void* p[MAX_RANKS];
};

DTC_DEVICE int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
DTC_DEVICE void st_release_sys(int* p, int v) { asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

// thread t (< world) of ONE CTA: signal peer t, then wait for peer t's signal.  flags layout per rank:
// [n_slots][world] ints, flags[slot][src] is written by rank `src`.
DTC_DEVICE void flag_barrier(const PeerPtrs& flags, int rank, int world, int slot, int t) {
  if (t < world) {
    __threadfence_system();
    int* remote = reinterpret_cast<int*>(flags.p[t]) + slot * world + rank;
    uint32_t spins = 0;
    while (atomicCAS_system(remote, 0, 1) != 0) {  // previous use of this flag not consumed yet
      if (++spins > (1u << 28)) {
        printf("dtc: barrier signal timeout rank %d -> %d slot %d\n", rank, t, slot);
        __trap();
      }
    }
    int* mine = reinterpret_cast<int*>(flags.p[rank]) + slot * world + t;
    spins = 0;
    while (ld_acquire_sys(mine) != 1) {
      if (++spins > (1u << 28)) {
        printf("dtc: barrier wait timeout rank %d <- %d slot %d\n", rank, t, slot);
        __trap();
      }
    }
    st_release_sys(mine, 0);
    __threadfence_system();
  }
}

__global__ void symm_barrier_kernel(PeerPtrs flags, int rank, int world, int slot) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  flag_barrier(flags, rank, world, slot, threadIdx.x);
}

DTC_DEVICE float4 mc_ld_reduce_add_f32x4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
DTC_DEVICE void mc_st_b32x4(void* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}

// Reduce the owned shard [off, off+n) of the flat gradient over all ranks; n % 4 == 0, off % 4 == 0.
template <bool USE_MC>
__global__ void __launch_bounds__(256)
dp_reduce_scatter_kernel(float* __restrict__ local_grad, const float* __restrict__ mc_grad, PeerPtrs peers, int rank,
                         int world, long long off, long long n, float* __restrict__ partials) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  __shared__ float s[8];
  float acc = 0.f;
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const long long e = off + i * 4;
    float4 v;
    if (USE_MC) {
      v = mc_ld_reduce_add_f32x4(mc_grad + e);
    } else {
      v = *reinterpret_cast<const float4*>(local_grad + e);
      for (int r = 1; r < world; ++r) {  // start at the next rank to spread the load over the links
        const int pr = (rank + r) % world;
        const float4 u = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(peers.p[pr]) + e);
        v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w;
      }
    }
    *reinterpret_cast<float4*>(local_grad + e) = v;
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < 8 ? s[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) partials[blockIdx.x] = v;
  }
}

// norm_slots: symmetric fp32 [world] per rank.  One CTA of 1024 threads.
__global__ void __launch_bounds__(1024)
dp_norm_exchange_kernel(const float* __restrict__ partials, int n_partials, PeerPtrs norm_slots, PeerPtrs flags, int rank,
                        int world, int slot, float max_norm, float* __restrict__ scalars, int* __restrict__ step) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  __shared__ float s[32];
  __shared__ float total;
  float v = 0.f;
  for (int i = threadIdx.x; i < n_partials; i += blockDim.x) v += partials[i];
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    v = s[threadIdx.x];
    v = warp_sum(v);
    if (threadIdx.x == 0) total = v;
  }
  __syncthreads();
  if (threadIdx.x < world) reinterpret_cast<float*>(norm_slots.p[threadIdx.x])[rank] = total;  // push to every peer
  flag_barrier(flags, rank, world, slot, threadIdx.x);
  __syncthreads();
  if (threadIdx.x == 0) {
    const volatile float* mine = reinterpret_cast<const volatile float*>(norm_slots.p[rank]);
    float sq = 0.f;
    for (int r = 0; r < world; ++r) sq += mine[r];
    const float norm = sqrtf(sq);
    scalars[0] = norm;
    scalars[1] = (max_norm > 0.f && norm >= max_norm) ? max_norm / norm : 1.f;
    if (step) *step += 1;
  }
}

// AdamW on the owned shard + all-gather of the new bf16 parameters; zeroes the WHOLE local gradient buffer.
template <bool USE_MC>
__global__ void __launch_bounds__(256)
dp_adamw_allgather_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, float* __restrict__ g,
                          bf16* __restrict__ mc_compute, PeerPtrs compute_peers, int world, long long off, long long n,
                          long long total_n, const float* __restrict__ scalars, const int* __restrict__ step_ptr, float lr,
                          float b1, float b2, float eps, float wd) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  const float clip = scalars[1];
  const float t = (float)*step_ptr;
  const float bc1 = 1.f / (1.f - powf(b1, t));
  const float bc2 = 1.f / (1.f - powf(b2, t));
  const long long n8 = n >> 3;  // 8 elements per thread-iteration: one 16-byte bf16 store per destination
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const long long e = off + i * 8;
    float pa[8], ma[8], va[8], ga[8];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float4 pp = *reinterpret_cast<const float4*>(p + e + 4 * h);
      const float4 mm = *reinterpret_cast<const float4*>(m + e + 4 * h);
      const float4 vv = *reinterpret_cast<const float4*>(v + e + 4 * h);
      const float4 gg = *reinterpret_cast<const float4*>(g + e + 4 * h);
      pa[4 * h] = pp.x; pa[4 * h + 1] = pp.y; pa[4 * h + 2] = pp.z; pa[4 * h + 3] = pp.w;
      ma[4 * h] = mm.x; ma[4 * h + 1] = mm.y; ma[4 * h + 2] = mm.z; ma[4 * h + 3] = mm.w;
      va[4 * h] = vv.x; va[4 * h + 1] = vv.y; va[4 * h + 2] = vv.z; va[4 * h + 3] = vv.w;
      ga[4 * h] = gg.x; ga[4 * h + 1] = gg.y; ga[4 * h + 2] = gg.z; ga[4 * h + 3] = gg.w;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gj = ga[j] * clip;
      ma[j] = b1 * ma[j] + (1.f - b1) * gj;
      va[j] = b2 * va[j] + (1.f - b2) * gj * gj;
      pa[j] -= lr * ((ma[j] * bc1) / (sqrtf(va[j] * bc2) + eps) + wd * pa[j]);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      *reinterpret_cast<float4*>(p + e + 4 * h) = make_float4(pa[4 * h], pa[4 * h + 1], pa[4 * h + 2], pa[4 * h + 3]);
      *reinterpret_cast<float4*>(m + e + 4 * h) = make_float4(ma[4 * h], ma[4 * h + 1], ma[4 * h + 2], ma[4 * h + 3]);
      *reinterpret_cast<float4*>(v + e + 4 * h) = make_float4(va[4 * h], va[4 * h + 1], va[4 * h + 2], va[4 * h + 3]);
      *reinterpret_cast<float4*>(g + e + 4 * h) = make_float4(0.f, 0.f, 0.f, 0.f);  // owned shard: cleared once consumed
    }
    const uint4 packed = make_uint4(pack_bf16x2(pa[0], pa[1]), pack_bf16x2(pa[2], pa[3]), pack_bf16x2(pa[4], pa[5]),
                                    pack_bf16x2(pa[6], pa[7]));
    if (USE_MC) {
      mc_st_b32x4(mc_compute + e, packed);
    } else {
      for (int r = 0; r < world; ++r) *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(compute_peers.p[r]) + e) = packed;
    }
  }
  // every rank has finished reading this rank's gradient (norm-exchange barrier): clear the part owned by the
  // other ranks for the next step (the owned shard is cleared above, element by element, after it has been read)
  const long long t4 = total_n >> 2, o4 = off >> 2, e4 = (off + n) >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < t4; i += (long long)gridDim.x * blockDim.x)
    if (i < o4 || i >= e4) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}


// ---- bucketed variants: several disjoint ranges per launch, so that a layer's gradient can be reduce-scattered on a
// ---- communication stream as soon as that layer's backward has finished (overlap with the rest of the backward).
// Work is balanced over the CTAs: the ranges form one virtual index space (prefix sums), every thread maps its
// group index back to (range, offset).
constexpr int MAX_RANGES = 16;
struct RangeList {
  long long off[MAX_RANGES];     // first element of the sub-range THIS rank owns (multiple of 4)
  long long prefix[MAX_RANGES + 1];  // prefix sums of the range lengths in 4-element groups
  int count;
};

// grid.x CTAs; partials[blockIdx.x]
template <bool USE_MC>
__global__ void __launch_bounds__(256)
dp_reduce_scatter_ranges_kernel(float* __restrict__ local_grad, const float* __restrict__ mc_grad, PeerPtrs peers, int rank,
                                int world, RangeList ranges, float* __restrict__ partials) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s[8];
  const long long total4 = ranges.prefix[ranges.count];
  float acc = 0.f;
  // NVLink round trips are ~2 us: keep 4 independent 16-byte reductions in flight per thread
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < total4; i0 += 4 * stride) {
    float4 v[4];
    long long e[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + u * stride;
      e[u] = -1;
      if (i < total4) {
        int r = 0;
        while (r + 1 < ranges.count && i >= ranges.prefix[r + 1]) ++r;
        e[u] = ranges.off[r] + (i - ranges.prefix[r]) * 4;
        if (USE_MC) {
          v[u] = mc_ld_reduce_add_f32x4(mc_grad + e[u]);
        } else {
          v[u] = *reinterpret_cast<const float4*>(local_grad + e[u]);
          for (int pr0 = 1; pr0 < world; ++pr0) {
            const int pr = (rank + pr0) % world;
            const float4 w = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(peers.p[pr]) + e[u]);
            v[u].x += w.x; v[u].y += w.y; v[u].z += w.z; v[u].w += w.w;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (e[u] >= 0) {
        *reinterpret_cast<float4*>(local_grad + e[u]) = v[u];
        acc += v[u].x * v[u].x + v[u].y * v[u].y + v[u].z * v[u].z + v[u].w * v[u].w;
      }
    }
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < 8 ? s[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) partials[blockIdx.x] = v;
  }
}

// AdamW + all-gather over a device-resident table of the owned sub-ranges: offs[n_ranges] (multiples of 8) and
// prefix[n_ranges + 1] (prefix sums of the lengths in 8-element groups).  The gradient buffer is cleared by the caller.
template <bool USE_MC>
__global__ void __launch_bounds__(256)
dp_adamw_allgather_ranges_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, float* __restrict__ g,
                                 bf16* __restrict__ mc_compute, PeerPtrs compute_peers, int world,
                                 const long long* __restrict__ offs, const long long* __restrict__ prefix, int n_ranges,
                                 const float* __restrict__ scalars, const int* __restrict__ step_ptr, float lr, float b1,
                                 float b2, float eps, float wd) {
  pdl_launch_dependents();
  pdl_wait();
  const float clip = scalars[1];
  const float t = (float)*step_ptr;
  const float bc1 = 1.f / (1.f - powf(b1, t));
  const float bc2 = 1.f / (1.f - powf(b2, t));
  const long long total8 = prefix[n_ranges];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total8; i += (long long)gridDim.x * blockDim.x) {
    int lo = 0, hi = n_ranges - 1;  // last range whose prefix <= i
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (prefix[mid] <= i) lo = mid; else hi = mid - 1;
    }
    const long long e = offs[lo] + (i - prefix[lo]) * 8;
    float pa[8], ma[8], va[8], ga[8];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float4 pp = *reinterpret_cast<const float4*>(p + e + 4 * h);
      const float4 mm = *reinterpret_cast<const float4*>(m + e + 4 * h);
      const float4 vv = *reinterpret_cast<const float4*>(v + e + 4 * h);
      const float4 gg = *reinterpret_cast<const float4*>(g + e + 4 * h);
      pa[4 * h] = pp.x; pa[4 * h + 1] = pp.y; pa[4 * h + 2] = pp.z; pa[4 * h + 3] = pp.w;
      ma[4 * h] = mm.x; ma[4 * h + 1] = mm.y; ma[4 * h + 2] = mm.z; ma[4 * h + 3] = mm.w;
      va[4 * h] = vv.x; va[4 * h + 1] = vv.y; va[4 * h + 2] = vv.z; va[4 * h + 3] = vv.w;
      ga[4 * h] = gg.x; ga[4 * h + 1] = gg.y; ga[4 * h + 2] = gg.z; ga[4 * h + 3] = gg.w;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gj = ga[j] * clip;
      ma[j] = b1 * ma[j] + (1.f - b1) * gj;
      va[j] = b2 * va[j] + (1.f - b2) * gj * gj;
      pa[j] -= lr * ((ma[j] * bc1) / (sqrtf(va[j] * bc2) + eps) + wd * pa[j]);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      *reinterpret_cast<float4*>(p + e + 4 * h) = make_float4(pa[4 * h], pa[4 * h + 1], pa[4 * h + 2], pa[4 * h + 3]);
      *reinterpret_cast<float4*>(m + e + 4 * h) = make_float4(ma[4 * h], ma[4 * h + 1], ma[4 * h + 2], ma[4 * h + 3]);
      *reinterpret_cast<float4*>(v + e + 4 * h) = make_float4(va[4 * h], va[4 * h + 1], va[4 * h + 2], va[4 * h + 3]);
    }
    const uint4 packed = make_uint4(pack_bf16x2(pa[0], pa[1]), pack_bf16x2(pa[2], pa[3]), pack_bf16x2(pa[4], pa[5]),
                                    pack_bf16x2(pa[6], pa[7]));
    if (USE_MC) {
      mc_st_b32x4(mc_compute + e, packed);
    } else {
      for (int r = 0; r < world; ++r) *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(compute_peers.p[r]) + e) = packed;
    }
  }
}

}  // namespace dtc

using namespace dtc;

static int fill_peers(PeerPtrs* out, const unsigned long long* ptrs, int world) {
  DTC_CHECK_ARG(world >= 1 && world <= MAX_RANKS, "comm: world size %d not in [1, %d]", world, MAX_RANKS);
  for (int i = 0; i < MAX_RANKS; ++i) out->p[i] = i < world ? reinterpret_cast<void*>(ptrs[i]) : nullptr;
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Small all-gather / all-reduce through symmetric memory (tensor-parallel step: cross-entropy statistics X7, partial
// gradients of the replicated LayerNorm / row-bias leaves).  Messages of a few 10-100 KB: latency, not bandwidth.
//   push : dst[slot_off + i] = src[i] on EVERY rank (multimem.st through the switch, or one store per peer)
//   sum  : out[i] (+)= sum_r slots[r * stride + i]                      (local, after a flag barrier)
// all-gather = push into slot `rank` + barrier;  all-reduce = push + barrier + sum.
// ---------------------------------------------------------------------------------------------------------
namespace dtc {
__global__ void symm_push_f32_kernel(const float* __restrict__ src, float* mc_dst, PeerPtrs peers, int world,
                                     long long dst_off, long long n4) {
  pdl_launch_dependents();
  pdl_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(src)[i];
    if (mc_dst != nullptr) {
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_dst + dst_off + i * 4), "f"(v.x),
                   "f"(v.y), "f"(v.z), "f"(v.w)
                   : "memory");
    } else {
      for (int r = 0; r < world; ++r) reinterpret_cast<float4*>(reinterpret_cast<float*>(peers.p[r]) + dst_off)[i] = v;
    }
  }
}
__global__ void sum_slots_f32_kernel(const float* __restrict__ slots, long long stride, int world, float* __restrict__ out,
                                     long long n4, int accumulate) {
  pdl_launch_dependents();
  pdl_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 a = accumulate ? reinterpret_cast<const float4*>(out)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < world; ++r) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(slots + r * stride) + i);  // written by peers: L2-coherent path
      a.x += v.x;
      a.y += v.y;
      a.z += v.z;
      a.w += v.w;
    }
    reinterpret_cast<float4*>(out)[i] = a;
  }
}
}  // namespace dtc

// dst_ptrs: HOST array of `world` device addresses of the symmetric destination buffer; mc_dst: its multicast address or 0.
// n and dst_off in floats, multiples of 4; src / destinations 16-byte aligned.
extern "C" int dtc_symm_push_f32(const float* src, unsigned long long mc_dst, const unsigned long long* dst_ptrs, int world,
                                 long long dst_off, long long n, cudaStream_t stream) {
  DTC_CHECK_ARG(n > 0 && n % 4 == 0 && dst_off % 4 == 0, "symm_push: n=%lld dst_off=%lld must be multiples of 4", n, dst_off);
  dtc::PeerPtrs peers;
  int rc = fill_peers(&peers, dst_ptrs, world);
  if (rc) return rc;
  const long long n4 = n / 4;
  int blocks = (int)((n4 + 255) / 256);
  if (blocks > 296) blocks = 296;
  DTC_CHECK_CUDA(dtc::launch(dtc::symm_push_f32_kernel, blocks, 256, 0, stream, src, reinterpret_cast<float*>(mc_dst), peers, world,
                             dst_off, n4));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_sum_slots_f32(const float* slots, long long stride, int world, float* out, long long n, int accumulate,
                                 cudaStream_t stream) {
  DTC_CHECK_ARG(n > 0 && n % 4 == 0 && stride % 4 == 0 && world >= 1, "sum_slots: n=%lld stride=%lld world=%d", n, stride, world);
  const long long n4 = n / 4;
  int blocks = (int)((n4 + 255) / 256);
  if (blocks > 296) blocks = 296;
  DTC_CHECK_CUDA(dtc::launch(dtc::sum_slots_f32_kernel, blocks, 256, 0, stream, slots, stride, world, out, n4, accumulate));
  DTC_CHECK_LAUNCH();
  return 0;
}

// flag_ptrs: HOST array of `world` device addresses (each rank's symmetric flag buffer, n_slots*world ints, zeroed).
extern "C" int dtc_symm_barrier(const unsigned long long* flag_ptrs, int rank, int world, int slot, cudaStream_t stream) {
  PeerPtrs f;
  int rc = fill_peers(&f, flag_ptrs, world);
  if (rc) return rc;
  DTC_CHECK_CUDA(dtc::launch(symm_barrier_kernel, 1, 32, 0, stream, f, rank, world, slot));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_dp_num_partials(long long n) {
  long long blocks = (n / 4 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 4) blocks = 148 * 4;
  return (int)blocks;
}

// mc_grad: multicast address of the flat gradient (0 = no multicast: peer loads through grad_ptrs).
extern "C" int dtc_dp_reduce_scatter(float* local_grad, unsigned long long mc_grad, const unsigned long long* grad_ptrs,
                                     int rank, int world, long long off, long long n, float* partials,
                                     cudaStream_t stream) {
  DTC_CHECK_ARG(n > 0 && n % 4 == 0 && off % 4 == 0, "dp_reduce_scatter: shard offset/size must be multiples of 4");
  PeerPtrs peers;
  int rc = fill_peers(&peers, grad_ptrs, world);
  if (rc) return rc;
  const int blocks = dtc_dp_num_partials(n);
  if (mc_grad)
    DTC_CHECK_CUDA(dtc::launch(dp_reduce_scatter_kernel<true>, blocks, 256, 0, stream, local_grad, reinterpret_cast<const float*>(mc_grad), peers,
                                                               rank, world, off, n, partials));
  else
    DTC_CHECK_CUDA(dtc::launch(dp_reduce_scatter_kernel<false>, blocks, 256, 0, stream, local_grad, nullptr, peers, rank, world, off, n, partials));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_dp_norm_exchange(const float* partials, int n_partials, const unsigned long long* norm_slot_ptrs,
                                    const unsigned long long* flag_ptrs, int rank, int world, int slot, float max_norm,
                                    float* scalars, int* step, cudaStream_t stream) {
  PeerPtrs ns, fl;
  int rc = fill_peers(&ns, norm_slot_ptrs, world);
  if (rc) return rc;
  if ((rc = fill_peers(&fl, flag_ptrs, world))) return rc;
  DTC_CHECK_CUDA(dtc::launch(dp_norm_exchange_kernel, 1, 1024, 0, stream, partials, n_partials, ns, fl, rank, world, slot, max_norm, scalars, step));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_dp_adamw_allgather(float* p, float* m, float* v, float* g, unsigned long long mc_compute,
                                      const unsigned long long* compute_ptrs, int world, long long off, long long n,
                                      long long total_n, const float* scalars, const int* step_ptr, float lr, float b1,
                                      float b2, float eps, float wd, cudaStream_t stream) {
  DTC_CHECK_ARG(n > 0 && n % 8 == 0 && off % 8 == 0 && total_n % 4 == 0, "dp_adamw_allgather: shard alignment");
  PeerPtrs peers;
  int rc = fill_peers(&peers, compute_ptrs, world);
  if (rc) return rc;
  long long blocks = (n / 8 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 148) blocks = 148;  // the gradient clear covers the whole buffer
  if (mc_compute)
    DTC_CHECK_CUDA(dtc::launch(dp_adamw_allgather_kernel<true>, (int)blocks, 256, 0, stream, p, m, v, g, reinterpret_cast<bf16*>(mc_compute), peers,
                                                                      world, off, n, total_n, scalars, step_ptr, lr, b1, b2,
                                                                      eps, wd));
  else
    DTC_CHECK_CUDA(dtc::launch(dp_adamw_allgather_kernel<false>, (int)blocks, 256, 0, stream, p, m, v, g, nullptr, peers, world, off, n, total_n,
                                                                       scalars, step_ptr, lr, b1, b2, eps, wd));
  DTC_CHECK_LAUNCH();
  return 0;
}

// ---- bucketed API ------------------------------------------------------------------------------------------------
// offs / lens: HOST arrays (count <= 16) of the sub-ranges this rank owns inside one bucket (lengths multiples of 4);
// `blocks` CTAs share the work evenly; partials must hold `blocks` floats.
extern "C" int dtc_dp_reduce_scatter_ranges(float* local_grad, unsigned long long mc_grad, const unsigned long long* grad_ptrs,
                                            int rank, int world, const long long* offs, const long long* lens, int count,
                                            int blocks, float* partials, cudaStream_t stream) {
  DTC_CHECK_ARG(count >= 1 && count <= MAX_RANGES, "dp_reduce_scatter_ranges: %d ranges (max %d)", count, MAX_RANGES);
  DTC_CHECK_ARG(blocks >= 1, "dp_reduce_scatter_ranges: blocks");
  PeerPtrs peers;
  int rc = fill_peers(&peers, grad_ptrs, world);
  if (rc) return rc;
  RangeList rl;
  rl.count = count;
  rl.prefix[0] = 0;
  for (int i = 0; i < MAX_RANGES; ++i) {
    rl.off[i] = i < count ? offs[i] : 0;
    const long long n = i < count ? lens[i] : 0;
    DTC_CHECK_ARG(rl.off[i] % 4 == 0 && n % 4 == 0, "dp_reduce_scatter_ranges: range %d not 4-aligned", i);
    rl.prefix[i + 1] = rl.prefix[i] + n / 4;
  }
  if (mc_grad)
    DTC_CHECK_CUDA(dtc::launch(dp_reduce_scatter_ranges_kernel<true>, blocks, 256, 0, stream, local_grad,
                               reinterpret_cast<const float*>(mc_grad), peers, rank, world, rl, partials));
  else
    DTC_CHECK_CUDA(dtc::launch(dp_reduce_scatter_ranges_kernel<false>, blocks, 256, 0, stream, local_grad, nullptr, peers, rank,
                               world, rl, partials));
  return 0;
}

// offs [n_ranges] / prefix [n_ranges + 1]: DEVICE arrays (long long); prefix in 8-element groups.
extern "C" int dtc_dp_adamw_allgather_ranges(float* p, float* m, float* v, float* g, unsigned long long mc_compute,
                                             const unsigned long long* compute_ptrs, int world, const long long* offs,
                                             const long long* prefix, int n_ranges, int blocks, const float* scalars,
                                             const int* step_ptr, float lr, float b1, float b2, float eps, float wd,
                                             cudaStream_t stream) {
  DTC_CHECK_ARG(n_ranges >= 1 && blocks >= 1, "dp_adamw_allgather_ranges: empty launch");
  PeerPtrs peers;
  int rc = fill_peers(&peers, compute_ptrs, world);
  if (rc) return rc;
  if (mc_compute)
    DTC_CHECK_CUDA(dtc::launch(dp_adamw_allgather_ranges_kernel<true>, blocks, 256, 0, stream, p, m, v, g,
                               reinterpret_cast<bf16*>(mc_compute), peers, world, offs, prefix, n_ranges, scalars, step_ptr, lr,
                               b1, b2, eps, wd));
  else
    DTC_CHECK_CUDA(dtc::launch(dp_adamw_allgather_ranges_kernel<false>, blocks, 256, 0, stream, p, m, v, g, nullptr, peers,
                               world, offs, prefix, n_ranges, scalars, step_ptr, lr, b1, b2, eps, wd));
  return 0;
}
