// Library-wide runtime helpers: error string, version, device probe, small utility kernels.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[1024] = "";

void dtc_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// Off by default: measured on B200 inside the captured step graph it changes nothing (6.56 ms with, 6.46 ms without —
// graph kernel nodes already start back to back); DTC_PDL=1 / dtc_set_pdl(1) enables it for eager launch loops.
int g_dtc_pdl_enabled = []() {
  const char* e = getenv("DTC_PDL");
  return (e && e[0] == '1') ? 1 : 0;
}();

// 1 = launch kernels with programmatic dependent launch, 0 = plain stream order (default).
extern "C" int dtc_set_pdl(int enabled) {
  g_dtc_pdl_enabled = enabled ? 1 : 0;
  return 0;
}

extern "C" const char* dtc_last_error() { return g_err; }
extern "C" int dtc_version() { return 100; }

// Returns the compute capability major*10+minor of the current device (e.g. 100 for B200), or <0.
extern "C" int dtc_device_arch() {
  int dev = 0, major = 0, minor = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev) != cudaSuccess) return -1;
  return major * 10 + minor;
}

namespace dtc {

__global__ void cast_f32_to_bf16_kernel(const float* __restrict__ in, bf16* __restrict__ out, long long n) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  const long long stride = (long long)gridDim.x * blockDim.x * 8;
  for (; i + 8 <= n; i += stride) {
    const float4 a = *reinterpret_cast<const float4*>(in + i);
    const float4 b = *reinterpret_cast<const float4*>(in + i + 4);
    *reinterpret_cast<uint4*>(out + i) =
        make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
  }
  // tail (n % 8) handled by the thread whose chunk is partial
  if (i < n && i + 8 > n) {
    for (long long j = i; j < n; ++j) out[j] = __float2bfloat16(in[j]);
  }
}

// dst_bf16[m, dst_col0 + c] = bf16(src_f32[m, c])   (attention-backward dQ accumulator -> dqkv slice)
// clear_src != 0: the source is zeroed behind the read (the attention backward's dQ accumulator is then clean for the
// next layer without a separate fill kernel).
__global__ void cast_f32_to_bf16_2d_kernel(float* __restrict__ in, long long ld_in, bf16* __restrict__ out,
                                           long long ld_out, int rows, int cols, int clear_src) {
  pdl_launch_dependents();  // PDL: let the next kernel's prologue overlap this kernel
  pdl_wait();               // ... and do not touch global memory before the previous kernel has completed
  const int cpr = cols / 8;
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)rows * cpr;
  for (; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(idx / cpr), c = (int)(idx % cpr) * 8;
    const float4 a = *reinterpret_cast<const float4*>(in + r * ld_in + c);
    const float4 b = *reinterpret_cast<const float4*>(in + r * ld_in + c + 4);
    *reinterpret_cast<uint4*>(out + r * ld_out + c) =
        make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
    if (clear_src) {
      *reinterpret_cast<float4*>(in + r * ld_in + c) = make_float4(0.f, 0.f, 0.f, 0.f);
      *reinterpret_cast<float4*>(in + r * ld_in + c + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// Test aid for the cross-rank flag protocols: stall the stream for a pseudo-random time that differs per
// rank, per call site and per training step, so that ranks reach every barrier / signal in a different order on every
// replay of the captured step graph.  `step` is the device-side optimizer step counter (changes between replays).
__global__ void jitter_kernel(unsigned seed, unsigned max_ns, const int* __restrict__ step) {
  unsigned x = seed * 2654435761u + (unsigned)(step ? *step : 0) * 40503u + 0x9e3779b9u;
  x ^= x >> 16;
  x *= 0x85ebca6bu;
  x ^= x >> 13;
  x *= 0xc2b2ae35u;
  x ^= x >> 16;
  const unsigned long long want = max_ns ? (x % max_ns) : 0;
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do {
    __nanosleep(200);
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  } while (t - t0 < want);
}

}  // namespace dtc

// Delay the stream by hash(seed, *step) % max_ns nanoseconds (one thread; see jitter_kernel).  max_ns <= 1 ms.
extern "C" int dtc_jitter(unsigned seed, unsigned max_ns, const int* step, cudaStream_t stream) {
  DTC_CHECK_ARG(max_ns <= 1000000u, "jitter: max_ns %u too large (<= 1000000)", max_ns);
  DTC_CHECK_CUDA(dtc::launch(dtc::jitter_kernel, 1, 1, 0, stream, seed, max_ns, step));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_cast_f32_to_bf16(const float* in, void* out, long long n, cudaStream_t stream) {
  DTC_CHECK_ARG(n >= 0, "cast: negative size");
  if (n == 0) return 0;
  DTC_CHECK_ARG(((uintptr_t)in & 15) == 0 && ((uintptr_t)out & 15) == 0, "cast: pointers must be 16-byte aligned");
  long long chunks = (n + 7) / 8;
  int blocks = (int)((chunks + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  DTC_CHECK_CUDA(dtc::launch(dtc::cast_f32_to_bf16_kernel, blocks, 256, 0, stream, in, reinterpret_cast<bf16*>(out), n));
  DTC_CHECK_LAUNCH();
  return 0;
}

static int cast2d(float* in, long long ld_in, void* out, long long ld_out, int rows, int cols, int clear_src,
                  cudaStream_t stream) {
  DTC_CHECK_ARG(cols % 8 == 0 && ld_in % 4 == 0 && ld_out % 8 == 0, "cast2d: cols/ld alignment");
  long long total = (long long)rows * (cols / 8);
  if (total == 0) return 0;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  DTC_CHECK_CUDA(dtc::launch(dtc::cast_f32_to_bf16_2d_kernel, blocks, 256, 0, stream, in, ld_in, reinterpret_cast<bf16*>(out), ld_out, rows, cols,
                             clear_src));
  DTC_CHECK_LAUNCH();
  return 0;
}

extern "C" int dtc_cast_f32_to_bf16_2d(const float* in, long long ld_in, void* out, long long ld_out, int rows, int cols,
                                       cudaStream_t stream) {
  return cast2d(const_cast<float*>(in), ld_in, out, ld_out, rows, cols, 0, stream);
}

// Same, and the fp32 source is set to zero behind the read.
extern "C" int dtc_cast_f32_to_bf16_2d_clear(float* in, long long ld_in, void* out, long long ld_out, int rows, int cols,
                                             cudaStream_t stream) {
  return cast2d(in, ld_in, out, ld_out, rows, cols, 1, stream);
}
