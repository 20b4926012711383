// Fused embedding: token gather + positional add + (always-on) dropout, and its backward
// (op sites K1/K2, reference: model/GPTModel.py:30-38).  The dropout mask is never stored: it is
// regenerated in the backward pass from a Philox-4x32-10 counter keyed by (seed, step, stream, element).
// `step_ptr` lives in device memory so a captured CUDA graph sees a fresh mask every replay.
#include "common.cuh"

namespace dtc {

DTC_DEVICE uint4 philox4x32_10(uint4 ctr, uint2 key) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

// keep-mask for 8 consecutive elements starting at flat element index e0 (e0 % 8 == 0)
DTC_DEVICE void dropout_keep8(unsigned long long seed, uint32_t step, uint32_t stream, unsigned long long e0,
                              uint32_t threshold, bool keep[8]) {
  const uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
  const unsigned long long c = e0 >> 2;
  const uint4 r0 = philox4x32_10(make_uint4((uint32_t)c, (uint32_t)(c >> 32), step, stream), key);
  const uint4 r1 = philox4x32_10(make_uint4((uint32_t)(c + 1), (uint32_t)((c + 1) >> 32), step, stream), key);
  keep[0] = r0.x >= threshold; keep[1] = r0.y >= threshold; keep[2] = r0.z >= threshold; keep[3] = r0.w >= threshold;
  keep[4] = r1.x >= threshold; keep[5] = r1.y >= threshold; keep[6] = r1.z >= threshold; keep[7] = r1.w >= threshold;
}

// one thread per 8 columns of one token
__global__ void embed_fwd_kernel(const int* __restrict__ ids, const bf16* __restrict__ wte, const bf16* __restrict__ wpe,
                                 bf16* __restrict__ out, int M, int T, int D, int V, unsigned long long seed,
                                 const int* __restrict__ step_ptr, uint32_t stream, uint32_t threshold, float keep_scale,
                                 long long row_offset) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  const int cpr = D / 8;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)M * cpr) return;
  const int m = (int)(idx / cpr), c = (int)(idx % cpr) * 8;
  int tok = ids[m];
  tok = tok < 0 ? 0 : (tok >= V ? V - 1 : tok);
  const int t = m % T;
  const uint4 a = *reinterpret_cast<const uint4*>(wte + (long long)tok * D + c);
  const uint4 b = *reinterpret_cast<const uint4*>(wpe + (long long)t * D + c);
  const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
  bool keep[8];
  const uint32_t step = step_ptr ? (uint32_t)*step_ptr : 0u;
  dropout_keep8(seed, step, stream, (unsigned long long)(row_offset + m) * D + c, threshold, keep);
  uint32_t o[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 fa = unpack_bf16x2(aw[j]), fb = unpack_bf16x2(bw[j]);
    const float v0 = keep[2 * j] ? (fa.x + fb.x) * keep_scale : 0.f;
    const float v1 = keep[2 * j + 1] ? (fa.y + fb.y) * keep_scale : 0.f;
    o[j] = pack_bf16x2(v0, v1);
  }
  *reinterpret_cast<uint4*>(out + (long long)m * D + c) = make_uint4(o[0], o[1], o[2], o[3]);
}

// one thread per (position t, 8 columns): loops over the batch, scatter-adds token rows, keeps the
// positional sum in registers.
__global__ void embed_bwd_kernel(const int* __restrict__ ids, const bf16* __restrict__ dh, float* __restrict__ dwte,
                                 float* __restrict__ dwpe, int Bsz, int T, int D, int V, unsigned long long seed,
                                 const int* __restrict__ step_ptr, uint32_t stream, uint32_t threshold, float keep_scale,
                                 long long row_offset) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  const int cpr = D / 8;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)T * cpr) return;
  const int t = (int)(idx / cpr), c = (int)(idx % cpr) * 8;
  const uint32_t step = step_ptr ? (uint32_t)*step_ptr : 0u;
  float pos[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) pos[j] = 0.f;
  for (int b = 0; b < Bsz; ++b) {
    const int m = b * T + t;
    int tok = ids[m];
    tok = tok < 0 ? 0 : (tok >= V ? V - 1 : tok);
    const uint4 u = *reinterpret_cast<const uint4*>(dh + (long long)m * D + c);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    bool keep[8];
    dropout_keep8(seed, step, stream, (unsigned long long)(row_offset + m) * D + c, threshold, keep);
    float g[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(w[j]);
      g[2 * j] = keep[2 * j] ? f.x * keep_scale : 0.f;
      g[2 * j + 1] = keep[2 * j + 1] ? f.y * keep_scale : 0.f;
    }
    float* dst = dwte + (long long)tok * D + c;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(dst + j, g[j]);
      pos[j] += g[j];
    }
  }
  float* dp = dwpe + (long long)t * D + c;
#pragma unroll
  for (int j = 0; j < 8; ++j) atomicAdd(dp + j, pos[j]);
}

}  // namespace dtc

using namespace dtc;

static inline uint32_t dropout_threshold(float p) {
  double th = (double)p * 4294967296.0;
  if (th < 0) th = 0;
  if (th > 4294967295.0) th = 4294967295.0;
  return (uint32_t)th;
}

// ids [M] int32 (M = B*T, row-major (b,t)), wte [V,D], wpe [>=T, D] bf16, out [M,D] bf16.
// row_offset: global row index of local row 0 (keeps masks distinct across data-parallel ranks /
// microbatches while a tensor-parallel group, which passes the same offset, sees identical masks).
extern "C" int dtc_embed_fwd(const int* ids, const void* wte, const void* wpe, void* out, int M, int T, int D, int V,
                             float dropout_p, unsigned long long seed, const int* step_ptr, unsigned int stream_id,
                             long long row_offset, cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && T > 0 && D % 8 == 0 && M % T == 0, "embed_fwd: bad shape M=%d T=%d D=%d", M, T, D);
  DTC_CHECK_ARG(dropout_p >= 0.f && dropout_p < 1.f, "embed_fwd: dropout must be in [0,1)");
  const long long total = (long long)M * (D / 8);
  const int blocks = (int)((total + 255) / 256);
  DTC_CHECK_CUDA(dtc::launch(embed_fwd_kernel, blocks, 256, 0, stream, ids, reinterpret_cast<const bf16*>(wte), reinterpret_cast<const bf16*>(wpe),
                                               reinterpret_cast<bf16*>(out), M, T, D, V, seed, step_ptr, stream_id,
                                               dropout_threshold(dropout_p), 1.f / (1.f - dropout_p), row_offset));
  DTC_CHECK_LAUNCH();
  return 0;
}

// dwte [V,D], dwpe [>=T,D] fp32 are ACCUMULATED into.
extern "C" int dtc_embed_bwd(const int* ids, const void* dh, float* dwte, float* dwpe, int M, int T, int D, int V,
                             float dropout_p, unsigned long long seed, const int* step_ptr, unsigned int stream_id,
                             long long row_offset, cudaStream_t stream) {
  DTC_CHECK_ARG(M > 0 && T > 0 && D % 8 == 0 && M % T == 0, "embed_bwd: bad shape M=%d T=%d D=%d", M, T, D);
  const long long total = (long long)T * (D / 8);
  const int blocks = (int)((total + 127) / 128);
  DTC_CHECK_CUDA(dtc::launch(embed_bwd_kernel, blocks, 128, 0, stream, ids, reinterpret_cast<const bf16*>(dh), dwte, dwpe, M / T, T, D, V, seed,
                                               step_ptr, stream_id, dropout_threshold(dropout_p), 1.f / (1.f - dropout_p),
                                               row_offset));
  DTC_CHECK_LAUNCH();
  return 0;
}
