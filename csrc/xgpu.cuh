// Cross-GPU context used inside kernels: peer-mapped data windows + signal pads of every rank (symmetric
// memory set up by csrc/symm.cu) and a device-side barrier built on system-scope release/acquire flags with
// monotonically increasing epochs (never reset, so CUDA-graph replays need no host bookkeeping).
#pragma once
#include "b200ddp.h"
#include "ptx.cuh"

namespace b2 {

struct XgpuCtx {
  int rank, world;
  float* data[B2_MAX_WORLD];
  uint32_t* flags[B2_MAX_WORLD];
  uint32_t* epoch;
};

static inline XgpuCtx xgpu_from_c(const b2_xgpu_ctx* c) {
  XgpuCtx x{};
  if (!c) {
    x.rank = 0;
    x.world = 1;
    return x;
  }
  x.rank = c->rank;
  x.world = c->world;
  for (int i = 0; i < B2_MAX_WORLD; ++i) {
    x.data[i] = reinterpret_cast<float*>(c->data[i]);
    x.flags[i] = reinterpret_cast<uint32_t*>(c->flags[i]);
  }
  x.epoch = reinterpret_cast<uint32_t*>(c->epoch);
  return x;
}

B2_DEVICE float ld_relaxed_sys_f32(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

// Debug aid for the signal protocols (SURVEY 5.2 "poisoned buffer / skew" mode): with a non-zero skew a
// pseudo-random subset of the CTAs (different per rank and per phase) sleeps for that long at the marked points of
// the collective kernels, so CTAs of one rank are far apart in progress. A kernel that consumes data which a
// DIFFERENTLY numbered CTA of a peer produced - the barrier below does not order that - then reads stale values
// and the numerical tests fail instead of passing by lock-step luck (tests/test_multigpu.py).
static __device__ unsigned long long g_debug_skew_ns = 0;
#define B2_DEFINE_DEBUG_SKEW_SETTER(name)                                                          \
  extern "C" int name(unsigned long long ns) {                                                     \
    return (int)cudaMemcpyToSymbol(b2::g_debug_skew_ns, &ns, sizeof(ns), 0, cudaMemcpyHostToDevice); \
  }
B2_DEVICE void debug_skew(int rank, int phase) {
  const unsigned long long ns = g_debug_skew_ns;
  if (ns == 0) return;
  if (((blockIdx.x * 7u + (unsigned)rank * 3u + (unsigned)phase * 5u) & 3u) != 1u) return;
  const uint64_t t0 = global_timer_ns();
  while (global_timer_ns() - t0 < ns) __nanosleep(20000);
}

// Barrier across the same-numbered CTA of every rank. Must be reached by all threads of the CTA.
// Pad layout on rank r: flags[r][channel * B2_MAX_WORLD + src] = last epoch signalled by rank `src`.
// Everything this CTA wrote before the call is visible to every peer CTA after it returns.
B2_DEVICE void xgpu_barrier(const XgpuCtx& xg, int channel) {
  __shared__ uint32_t s_epoch;
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t e = xg.epoch[channel] + 1;
    xg.epoch[channel] = e;
    s_epoch = e;
  }
  __syncthreads();
  const uint32_t e = s_epoch;
  if (threadIdx.x < xg.world) {
    const int peer = threadIdx.x;
    fence_acq_rel_sys();
    st_release_sys(xg.flags[peer] + channel * B2_MAX_WORLD + xg.rank, e);
    const uint32_t* mine = xg.flags[xg.rank] + channel * B2_MAX_WORLD + peer;
    SpinGuard guard;
    while ((int32_t)(ld_acquire_sys(mine) - e) < 0) guard.tick();
  }
  __syncthreads();
}

}  // namespace b2
