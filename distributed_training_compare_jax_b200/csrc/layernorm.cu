// LayerNorm forward / backward (op site K4, reference: model/TransformerBlock.py:16,22 and model/GPTModel.py:71).
// Semantics follow flax.linen.LayerNorm defaults: eps 1e-6, learnable scale + bias, "fast variance"
// var = E[x^2] - E[x]^2, statistics in fp32.  One warp per row, 16-byte loads, row re-read served by L1.
// The backward kernel also adds the incoming residual gradient (dx = dres + LN'(dy)) and reduces
// dgamma / dbeta per CTA before one fp32 reduce-add per column.
#include "common.cuh"

namespace dtc {

constexpr int LN_WARPS = 8;

__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ gamma, const bf16* __restrict__ beta,
                     bf16* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out, int M, int D,
                     float eps) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * LN_WARPS + warp;
  if (row >= M) return;
  const bf16* xr = x + row * D;
  float s = 0.f, ss = 0.f;
  for (int c = lane * 8; c < D; c += 256) {
    const uint4 u = *reinterpret_cast<const uint4*>(xr + c);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(w[j]);
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
  bf16* yr = y + row * D;
  for (int c = lane * 8; c < D; c += 256) {
    const uint4 u = *reinterpret_cast<const uint4*>(xr + c);
    const uint4 g = *reinterpret_cast<const uint4*>(gamma + c);
    const uint4 b = *reinterpret_cast<const uint4*>(beta + c);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w}, gw[4] = {g.x, g.y, g.z, g.w}, bw[4] = {b.x, b.y, b.z, b.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(w[j]), gg = unpack_bf16x2(gw[j]), bb = unpack_bf16x2(bw[j]);
      o[j] = pack_bf16x2((f.x - mean) * rstd * gg.x + bb.x, (f.y - mean) * rstd * gg.y + bb.y);
    }
    *reinterpret_cast<uint4*>(yr + c) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// NCH = D / 256 chunks of 8 columns per lane.
template <int NCH>
__global__ void __launch_bounds__(LN_WARPS * 32)
layernorm_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, const float* __restrict__ mean,
                     const float* __restrict__ rstd, const bf16* __restrict__ gamma, const bf16* __restrict__ dres,
                     bf16* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta, int M) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  constexpr int D = NCH * 256;
  __shared__ float s_dg[D];
  __shared__ float s_db[D];
  for (int i = threadIdx.x; i < D; i += blockDim.x) {
    s_dg[i] = 0.f;
    s_db[i] = 0.f;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float acc_dg[NCH * 8], acc_db[NCH * 8], gam[NCH * 8];
#pragma unroll
  for (int k = 0; k < NCH; ++k) {
    const uint4 g = *reinterpret_cast<const uint4*>(gamma + k * 256 + lane * 8);
    const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(gw[j]);
      gam[k * 8 + 2 * j] = f.x;
      gam[k * 8 + 2 * j + 1] = f.y;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      acc_dg[k * 8 + j] = 0.f;
      acc_db[k * 8 + j] = 0.f;
    }
  }
  for (long long row = (long long)blockIdx.x * LN_WARPS + warp; row < M; row += (long long)gridDim.x * LN_WARPS) {
    const float mu = mean[row], rs = rstd[row];
    float xhat[NCH * 8], wdy[NCH * 8];
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      const long long off = row * D + k * 256 + lane * 8;
      const uint4 ux = *reinterpret_cast<const uint4*>(x + off);
      const uint4 ud = *reinterpret_cast<const uint4*>(dy + off);
      const uint32_t xw[4] = {ux.x, ux.y, ux.z, ux.w}, dw[4] = {ud.x, ud.y, ud.z, ud.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 xf = unpack_bf16x2(xw[j]), df = unpack_bf16x2(dw[j]);
        const int i0 = k * 8 + 2 * j, i1 = i0 + 1;
        xhat[i0] = (xf.x - mu) * rs;
        xhat[i1] = (xf.y - mu) * rs;
        acc_dg[i0] += df.x * xhat[i0];
        acc_dg[i1] += df.y * xhat[i1];
        acc_db[i0] += df.x;
        acc_db[i1] += df.y;
        wdy[i0] = df.x * gam[i0];
        wdy[i1] = df.y * gam[i1];
        c1 += wdy[i0] + wdy[i1];
        c2 += wdy[i0] * xhat[i0] + wdy[i1] * xhat[i1];
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
      if (dres != nullptr) {
        const uint4 ur = *reinterpret_cast<const uint4*>(dres + off);
        const uint32_t rw[4] = {ur.x, ur.y, ur.z, ur.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 f = unpack_bf16x2(rw[j]);
          r[2 * j] += f.x;
          r[2 * j + 1] += f.y;
        }
      }
      *reinterpret_cast<uint4*>(dx + off) =
          make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7]));
    }
  }
  // CTA-level reduction of the per-lane column partials, then one global reduce-add per column.
#pragma unroll
  for (int k = 0; k < NCH; ++k) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(&s_dg[k * 256 + lane * 8 + j], acc_dg[k * 8 + j]);
      atomicAdd(&s_db[k * 256 + lane * 8 + j], acc_db[k * 8 + j]);
    }
  }
  __syncthreads();
  // 16-byte vector reductions: a quarter of the L2 atomic operations of a scalar atomicAdd per column (296 CTAs add into
  // the same D columns, so this tail is serialised per address in L2)
  for (int i = threadIdx.x * 4; i < D; i += blockDim.x * 4) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dgamma + i), "f"(s_dg[i]), "f"(s_dg[i + 1]),
                 "f"(s_dg[i + 2]), "f"(s_dg[i + 3])
                 : "memory");
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dbeta + i), "f"(s_db[i]), "f"(s_db[i + 1]),
                 "f"(s_db[i + 2]), "f"(s_db[i + 3])
                 : "memory");
  }
}

// Backward for D <= 512 (D = 768 would spill at 128 registers per thread), one 16-warp CTA per SM.  Against the kernel above: the next row's x / dy (and this row's residual
// gradient) are in flight while the current row is reduced; the per-lane column partials go to shared memory as plain
// 16-byte stores, one [D] row per warp, and are summed over the warps by column (no shared-memory atomics, which hit the
// same banks 8 ways with the lane * 8 + j layout); and half as many CTAs add into the same dgamma / dbeta addresses.
constexpr int LNB_WARPS = 16;

template <int NCH>
__global__ void __launch_bounds__(LNB_WARPS * 32, 1)
layernorm_bwd_wide_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x, const float* __restrict__ mean,
                          const float* __restrict__ rstd, const bf16* __restrict__ gamma, const bf16* __restrict__ dres,
                          bf16* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta, int M) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int D = NCH * 256;
  extern __shared__ float s_part[];  // [2][LNB_WARPS][D]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long stride = (long long)gridDim.x * LNB_WARPS;
  long long row = (long long)blockIdx.x * LNB_WARPS + warp;
  uint4 ux[NCH], ud[NCH];
  float mu = 0.f, rs = 0.f;
  if (row < M) {
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      const long long off = row * D + k * 256 + lane * 8;
      ux[k] = *reinterpret_cast<const uint4*>(x + off);
      ud[k] = *reinterpret_cast<const uint4*>(dy + off);
    }
    mu = mean[row];
    rs = rstd[row];
  }
  float acc_dg[NCH * 8], acc_db[NCH * 8], gam[NCH * 8];
#pragma unroll
  for (int k = 0; k < NCH; ++k) {
    const uint4 g = *reinterpret_cast<const uint4*>(gamma + k * 256 + lane * 8);
    const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(gw[j]);
      gam[k * 8 + 2 * j] = f.x;
      gam[k * 8 + 2 * j + 1] = f.y;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      acc_dg[k * 8 + j] = 0.f;
      acc_db[k * 8 + j] = 0.f;
    }
  }
  while (row < M) {
    const long long nrow = row + stride;
    uint4 nx[NCH], nd[NCH], ur[NCH];
    float nmu = 0.f, nrs = 0.f;
    if (nrow < M) {
#pragma unroll
      for (int k = 0; k < NCH; ++k) {
        const long long off = nrow * D + k * 256 + lane * 8;
        nx[k] = *reinterpret_cast<const uint4*>(x + off);
        nd[k] = *reinterpret_cast<const uint4*>(dy + off);
      }
      nmu = mean[nrow];
      nrs = rstd[nrow];
    }
    if (dres != nullptr) {
#pragma unroll
      for (int k = 0; k < NCH; ++k) ur[k] = *reinterpret_cast<const uint4*>(dres + row * D + k * 256 + lane * 8);
    }
    float xhat[NCH * 8], wdy[NCH * 8];
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      const uint32_t xw[4] = {ux[k].x, ux[k].y, ux[k].z, ux[k].w}, dw[4] = {ud[k].x, ud[k].y, ud[k].z, ud[k].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 xf = unpack_bf16x2(xw[j]), df = unpack_bf16x2(dw[j]);
        const int i0 = k * 8 + 2 * j, i1 = i0 + 1;
        xhat[i0] = (xf.x - mu) * rs;
        xhat[i1] = (xf.y - mu) * rs;
        acc_dg[i0] += df.x * xhat[i0];
        acc_dg[i1] += df.y * xhat[i1];
        acc_db[i0] += df.x;
        acc_db[i1] += df.y;
        wdy[i0] = df.x * gam[i0];
        wdy[i1] = df.y * gam[i1];
        c1 += wdy[i0] + wdy[i1];
        c2 += wdy[i0] * xhat[i0] + wdy[i1] * xhat[i1];
      }
    }
    c1 = warp_sum(c1) * (1.f / D);
    c2 = warp_sum(c2) * (1.f / D);
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      float r[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) r[j] = rs * (wdy[k * 8 + j] - c1 - xhat[k * 8 + j] * c2);
      if (dres != nullptr) {
        const uint32_t rw[4] = {ur[k].x, ur[k].y, ur[k].z, ur[k].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 f = unpack_bf16x2(rw[j]);
          r[2 * j] += f.x;
          r[2 * j + 1] += f.y;
        }
      }
      *reinterpret_cast<uint4*>(dx + row * D + k * 256 + lane * 8) =
          make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7]));
    }
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
      ux[k] = nx[k];
      ud[k] = nd[k];
    }
    mu = nmu;
    rs = nrs;
    row = nrow;
  }
  float* s_dg = s_part;
  float* s_db = s_part + LNB_WARPS * D;
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
    for (int w = 0; w < LNB_WARPS; ++w) {
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
}

}  // namespace dtc

using namespace dtc;

extern "C" int dtc_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                                 int M, int D, float eps, cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && D > 0 && D % 8 == 0, "layernorm_fwd: need D %% 8 == 0 (M=%d D=%d)", M, D);
  const int blocks = (M + LN_WARPS - 1) / LN_WARPS;
  DTC_CHECK_CUDA(dtc::launch(layernorm_fwd_kernel, blocks, LN_WARPS * 32, 0, stream, 
      reinterpret_cast<const bf16*>(x), reinterpret_cast<const bf16*>(gamma), reinterpret_cast<const bf16*>(beta),
      reinterpret_cast<bf16*>(y), mean, rstd, M, D, eps));
  DTC_CHECK_LAUNCH();
  return 0;
}

// dx = (dres ? dres : 0) + LayerNorm'(dy); dgamma/dbeta (fp32, [D]) are ACCUMULATED into.
extern "C" int dtc_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd, const void* gamma,
                                 const void* dres, void* dx, float* dgamma, float* dbeta, int M, int D,
                                 cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && D % 256 == 0 && D >= 256 && D <= 2048, "layernorm_bwd: D=%d must be a multiple of 256, <= 2048", D);
  DTC_CHECK_ARG((((uintptr_t)dgamma | (uintptr_t)dbeta) & 15) == 0, "layernorm_bwd: dgamma / dbeta must be 16-byte aligned");
  static const bool v1 = getenv("DTC_LN_BWD_V1") != nullptr;  // A/B: the 8-warp kernel for every D
  if (D <= 512 && !v1) {
    int sms = 148, dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int wblocks = (M + LNB_WARPS - 1) / LNB_WARPS;
    if (wblocks > sms) wblocks = sms;
    const int smem = 2 * LNB_WARPS * D * (int)sizeof(float);
#define DTC_LN_BWD_WIDE(NCH)                                                                                          \
  case NCH: {                                                                                                         \
    static dtc::SmemOptIn configured;                                                                                 \
    DTC_CHECK_CUDA(dtc::ensure_dyn_smem(layernorm_bwd_wide_kernel<NCH>, smem, configured));                           \
    DTC_CHECK_CUDA(dtc::launch(layernorm_bwd_wide_kernel<NCH>, wblocks, LNB_WARPS * 32, smem, stream,                 \
        reinterpret_cast<const bf16*>(dy), reinterpret_cast<const bf16*>(x), mean, rstd,                              \
        reinterpret_cast<const bf16*>(gamma), reinterpret_cast<const bf16*>(dres), reinterpret_cast<bf16*>(dx),       \
        dgamma, dbeta, M));                                                                                            \
    break;                                                                                                            \
  }
    switch (D / 256) { DTC_LN_BWD_WIDE(1) DTC_LN_BWD_WIDE(2) }
#undef DTC_LN_BWD_WIDE
    DTC_CHECK_LAUNCH();
    return 0;
  }
  int blocks = (M + LN_WARPS - 1) / LN_WARPS;
  if (blocks > 148 * 2) blocks = 148 * 2;
#define DTC_LN_BWD(NCH)                                                                                             \
  case NCH:                                                                                                         \
    DTC_CHECK_CUDA(dtc::launch(layernorm_bwd_kernel<NCH>, blocks, LN_WARPS * 32, 0, stream,                                                 \
        reinterpret_cast<const bf16*>(dy), reinterpret_cast<const bf16*>(x), mean, rstd,                            \
        reinterpret_cast<const bf16*>(gamma), reinterpret_cast<const bf16*>(dres), reinterpret_cast<bf16*>(dx),     \
        dgamma, dbeta, M));                                                                                          \
    break;
  switch (D / 256) {
    DTC_LN_BWD(1) DTC_LN_BWD(2) DTC_LN_BWD(3) DTC_LN_BWD(4) DTC_LN_BWD(5) DTC_LN_BWD(6) DTC_LN_BWD(7) DTC_LN_BWD(8)
    default: dtc_set_error("layernorm_bwd: unsupported D=%d", D); return -1;
  }
#undef DTC_LN_BWD
  DTC_CHECK_LAUNCH();
  return 0;
}
