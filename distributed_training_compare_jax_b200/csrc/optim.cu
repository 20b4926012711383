// Optimizer-side kernels over FLAT parameter / gradient buffers (op sites K13/K14, reference:
// train/create_optimizer.py:9-12 = optax.chain(clip_by_global_norm, adamw)):
//   dtc_sumsq_partials   per-CTA partial sums of g^2 over a flat fp32 range
//   dtc_clip_finalize    total norm, clip factor min(1, c/||g||) (optax: where(norm < c, 1, c/norm)), step += 1
//   dtc_adamw            multi-tensor AdamW in ONE launch over the flat buffer: m, v, bias correction,
//                        decoupled weight decay on every element, clip factor folded in, bf16 compute
//                        copy written, gradient optionally zeroed for the next step
//   dtc_colsum           bias gradient: column sums of a bf16 [M,N] matrix accumulated into fp32 [N]
// Scalars that change every step (step count, clip factor) are read from device memory so that the
// whole training step can be replayed as one CUDA graph.
#include "common.cuh"

namespace dtc {

__global__ void __launch_bounds__(256) sumsq_partials_kernel(const float* __restrict__ g, long long n,
                                                             float* __restrict__ partials) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  __shared__ float s[8];
  float acc = 0.f;
  const long long n4 = n >> 2;
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = g4[i];
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long i = n4 << 2; i < n; ++i) acc += g[i] * g[i];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < 8 ? s[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) partials[blockIdx.x] = v;
  }
}

// scalars[0] = ||g|| ; scalars[1] = clip factor ; *step += 1 (if step != null)
__global__ void __launch_bounds__(1024) clip_finalize_kernel(const float* __restrict__ partials, int n, float grad_scale,
                                                             float max_norm, float* __restrict__ scalars,
                                                             int* __restrict__ step) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  __shared__ float s[32];
  float v = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) v += partials[i];
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    v = s[threadIdx.x];
    v = warp_sum(v);
    if (threadIdx.x == 0) {
      const float norm = sqrtf(v) * grad_scale;
      scalars[0] = norm;
      scalars[1] = (max_norm > 0.f && norm >= max_norm) ? max_norm / norm : 1.f;
      if (step) *step += 1;
    }
  }
}

__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, float* __restrict__ g,
             bf16* __restrict__ p_bf16, long long n, const float* __restrict__ scalars, const int* __restrict__ step_ptr,
             float lr, float b1, float b2, float eps, float wd, float grad_scale, int zero_grad) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  const float clip = scalars ? scalars[1] : 1.f;
  const float gs = grad_scale * clip;
  const float t = (float)(step_ptr ? *step_ptr : 1);
  const float bc1 = 1.f / (1.f - powf(b1, t));
  const float bc2 = 1.f / (1.f - powf(b2, t));
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    const float4 gg = reinterpret_cast<float4*>(g)[i];
    float* pa = reinterpret_cast<float*>(&pp);
    float* ma = reinterpret_cast<float*>(&mm);
    float* va = reinterpret_cast<float*>(&vv);
    const float* ga = reinterpret_cast<const float*>(&gg);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gj = ga[j] * gs;
      ma[j] = b1 * ma[j] + (1.f - b1) * gj;
      va[j] = b2 * va[j] + (1.f - b2) * gj * gj;
      const float upd = (ma[j] * bc1) / (sqrtf(va[j] * bc2) + eps) + wd * pa[j];
      pa[j] -= lr * upd;
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p_bf16) reinterpret_cast<uint2*>(p_bf16)[i] = make_uint2(pack_bf16x2(pa[0], pa[1]), pack_bf16x2(pa[2], pa[3]));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    for (long long i = n4 << 2; i < n; ++i) {
      const float gj = g[i] * gs;
      m[i] = b1 * m[i] + (1.f - b1) * gj;
      v[i] = b2 * v[i] + (1.f - b2) * gj * gj;
      p[i] -= lr * ((m[i] * bc1) / (sqrtf(v[i] * bc2) + eps) + wd * p[i]);
      if (zero_grad) g[i] = 0.f;
      if (p_bf16) p_bf16[i] = __float2bfloat16(p[i]);
    }
  }
}

// grid.x over 256-column slabs, grid.y row splits; block (32 column-groups of 8) x 8 row lanes.
__global__ void __launch_bounds__(256) colsum_kernel(const bf16* __restrict__ x, long long ld, int M, int N,
                                                     float* __restrict__ out) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  __shared__ float s[8][256 + 8];
  const int cg = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int c = blockIdx.x * 256 + cg * 8;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (c < N) {
    for (int r = blockIdx.y * 8 + rl; r < M; r += gridDim.y * 8) {
      const uint4 u = *reinterpret_cast<const uint4*>(x + (long long)r * ld + c);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(w[j]);
        acc[2 * j] += f.x;
        acc[2 * j + 1] += f.y;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) s[rl][cg * 8 + j] = acc[j];
  __syncthreads();
  const int col = threadIdx.x;
  if (blockIdx.x * 256 + col < N) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) t += s[r][col];
    atomicAdd(out + blockIdx.x * 256 + col, t);
  }
}

// The bias gradients of one layer in one launch: blockIdx.z selects the problem (same body as colsum_kernel).
constexpr int CG_MAX = 4;
struct ColsumGroup {
  const bf16* x[CG_MAX];
  long long ld[CG_MAX];
  int M[CG_MAX];
  int N[CG_MAX];
  float* out[CG_MAX];
};

__global__ void __launch_bounds__(256) colsum_grouped_kernel(const __grid_constant__ ColsumGroup grp) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s[8][256 + 8];
  const int p = blockIdx.z;
  const bf16* __restrict__ x = grp.x[p];
  const long long ld = grp.ld[p];
  const int M = grp.M[p], N = grp.N[p];
  if ((int)blockIdx.x * 256 >= N) return;   // uniform per CTA: narrower problems use fewer column blocks
  const int cg = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int c = blockIdx.x * 256 + cg * 8;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (c < N) {
    for (int r = blockIdx.y * 8 + rl; r < M; r += gridDim.y * 8) {
      const uint4 u = *reinterpret_cast<const uint4*>(x + (long long)r * ld + c);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(w[j]);
        acc[2 * j] += f.x;
        acc[2 * j + 1] += f.y;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) s[rl][cg * 8 + j] = acc[j];
  __syncthreads();
  const int col = threadIdx.x;
  if (blockIdx.x * 256 + col < N) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) t += s[r][col];
    atomicAdd(grp.out[p] + blockIdx.x * 256 + col, t);
  }
}

}  // namespace dtc

using namespace dtc;

extern "C" int dtc_sumsq_num_partials(long long n) {
  long long blocks = (n / 4 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  return (int)blocks;
}

// partials must hold dtc_sumsq_num_partials(n) floats.
extern "C" int dtc_sumsq_partials(const float* g, long long n, float* partials, cudaStream_t stream) {
  DTC_CHECK_ARG(n > 0 && ((uintptr_t)g & 15) == 0, "sumsq: n must be > 0 and g 16-byte aligned");
  DTC_CHECK_CUDA(dtc::launch(sumsq_partials_kernel, dtc_sumsq_num_partials(n), 256, 0, stream, g, n, partials));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_clip_finalize(const float* partials, int n_partials, float grad_scale, float max_norm, float* scalars,
                                 int* step, cudaStream_t stream) {
  DTC_CHECK_ARG(n_partials > 0, "clip_finalize: no partials");
  DTC_CHECK_CUDA(dtc::launch(clip_finalize_kernel, 1, 1024, 0, stream, partials, n_partials, grad_scale, max_norm, scalars, step));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_adamw(float* p, float* m, float* v, float* g, void* p_bf16, long long n, const float* scalars,
                         const int* step_ptr, float lr, float b1, float b2, float eps, float wd, float grad_scale,
                         int zero_grad, cudaStream_t stream) {
  DTC_CHECK_ARG(n > 0, "adamw: empty range");
  DTC_CHECK_ARG((((uintptr_t)p | (uintptr_t)m | (uintptr_t)v | (uintptr_t)g) & 15) == 0 && ((uintptr_t)p_bf16 & 7) == 0,
                "adamw: buffers must be 16-byte aligned");
  long long blocks = (n / 4 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  DTC_CHECK_CUDA(dtc::launch(adamw_kernel, (int)blocks, 256, 0, stream, p, m, v, g, reinterpret_cast<bf16*>(p_bf16), n, scalars, step_ptr, lr, b1,
                                                b2, eps, wd, grad_scale, zero_grad));
  DTC_CHECK_LAUNCH();
  return 0;
}

// out[n] += sum_m x[m, n]   (x bf16 [M, N] with leading dimension ld, N % 8 == 0)
extern "C" int dtc_colsum(const void* x, long long ld, int M, int N, float* out, cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && N > 0 && N % 8 == 0 && ld % 8 == 0, "colsum: bad shape M=%d N=%d ld=%lld", M, N, ld);
  int ysplit = (M + 31) / 32;  // >= 4 rows per thread; enough CTAs to fill 148 SMs even for N = 512
  if (ysplit > 128) ysplit = 128;
  dim3 grid((N + 255) / 256, ysplit);
  DTC_CHECK_CUDA(dtc::launch(colsum_kernel, grid, 256, 0, stream, reinterpret_cast<const bf16*>(x), ld, M, N, out));
  DTC_CHECK_LAUNCH();
  return 0;
}

// out[p][n] += sum_m x[p][m, n] for p < count (<= 4): the bias gradients of one layer in one launch.
// All arrays are HOST arrays; x[p] bf16 [M[p], N[p]] with leading dimension ld[p], N[p] % 8 == 0.
extern "C" int dtc_colsum_grouped(int count, const unsigned long long* x, const long long* ld, const int* M, const int* N,
                                  const unsigned long long* out, cudaStream_t stream) {
  DTC_CHECK_ARG(count >= 1 && count <= CG_MAX, "colsum_grouped: count=%d (1..%d)", count, CG_MAX);
  ColsumGroup grp;
  int max_n = 0, max_m = 0;
  for (int p = 0; p < CG_MAX; ++p) {
    const int q = p < count ? p : 0;
    DTC_CHECK_ARG(M[q] > 0 && N[q] > 0 && N[q] % 8 == 0 && ld[q] % 8 == 0 && x[q] && out[q] && (x[q] & 15) == 0,
                  "colsum_grouped: problem %d: bad shape M=%d N=%d ld=%lld or pointer", q, M[q], N[q], ld[q]);
    grp.x[p] = reinterpret_cast<const bf16*>(x[q]);
    grp.ld[p] = ld[q];
    grp.M[p] = M[q];
    grp.N[p] = N[q];
    grp.out[p] = reinterpret_cast<float*>(out[q]);
    if (p < count && N[p] > max_n) max_n = N[p];
    if (p < count && M[p] > max_m) max_m = M[p];
  }
  int ysplit = (max_m + 31) / 32;
  if (ysplit > 64) ysplit = 64;
  dim3 grid((max_n + 255) / 256, ysplit, count);
  DTC_CHECK_CUDA(dtc::launch(colsum_grouped_kernel, grid, 256, 0, stream, grp));
  DTC_CHECK_LAUNCH();
  return 0;
}
