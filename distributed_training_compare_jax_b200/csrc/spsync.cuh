// In-kernel cross-rank synchronisation for the sequence-parallel tensor-parallel path (experimental "folded barrier"
// variants: dtc_sp_ln_*_sync, dtc_gemm_bf16_scatter_signal, dtc_sp_wait).  Shared by seqpar.cu and gemm.cu.
#pragma once
#include "common.cuh"

namespace dtc {

constexpr int SP_MAX_RANKS = 8;

// Every rank owns sig[world] ints in symmetric memory; sig[src] holds the sequence number of the last signal rank `src`
// has issued.  All ranks run the same sequence of kernels, so the k-th signal of every rank pairs with the k-th wait of
// every rank: counters[0] = signals issued, counters[1] = waits completed (device memory, so CUDA-graph replays keep
// counting), counters[2] / counters[3] = CTA tickets of the kernel that is currently signalling / waiting.
struct SpSync {
  int* sig[SP_MAX_RANKS];
  int* counters;
  int rank, world;
};

DTC_DEVICE int sp_ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
DTC_DEVICE void sp_st_release_sys(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Threads t < world of a CTA: wait until rank t has issued the signal this kernel consumes.  Call before touching the
// data, follow with a CTA barrier.  Bounded: a protocol error traps instead of hanging the node.
DTC_DEVICE void sp_wait_signals(const SpSync& s, int t) {
  if (t < s.world) {
    const int expected = *reinterpret_cast<volatile int*>(s.counters + 1) + 1;
    const int* flag = s.sig[s.rank] + t;
    uint32_t spins = 0;
    while (sp_ld_acquire_sys(flag) < expected) {
      if (++spins > (1u << 28)) {
        printf("dtc: sequence-parallel wait timeout: rank %d <- %d, expected %d\n", s.rank, t, expected);
        __trap();
      }
    }
  }
}

// ONE thread per CTA, after every thread of the CTA has waited (CTA barrier in between): the last CTA of the grid
// advances the wait counter.
DTC_DEVICE void sp_wait_done(const SpSync& s, unsigned n_ctas) {
  if (atomicAdd(reinterpret_cast<unsigned*>(s.counters + 3), 1u) == n_ctas - 1) {
    s.counters[3] = 0;
    s.counters[1] += 1;
  }
}

// ONE thread per CTA, after every thread of the CTA that stored remote data has executed __threadfence_system() and the
// CTA has synchronised: the last CTA of the grid publishes the next sequence number to every rank.
DTC_DEVICE void sp_signal(const SpSync& s, unsigned n_ctas) {
  if (atomicAdd(reinterpret_cast<unsigned*>(s.counters + 2), 1u) == n_ctas - 1) {
    __threadfence_system();  // the tickets of the other CTAs (and the stores they fenced) are ordered before the flags
    s.counters[2] = 0;
    const int seq = s.counters[0] + 1;
    s.counters[0] = seq;
#pragma unroll
    for (int p = 0; p < SP_MAX_RANKS; ++p)
      if (p < s.world) sp_st_release_sys(s.sig[p] + s.rank, seq);
  }
}


}  // namespace dtc
