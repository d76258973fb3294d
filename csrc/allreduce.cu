// Gradient all-reduce and broadcast over peer-mapped (symmetric) memory, fused with the 1/world scale and the
// SGD-momentum update.  Kernels read and write other GPUs' bucket buffers directly over NVLink 5 / NVSwitch;
// NCCL is never called.
//
//   one-shot : every rank reads the whole bucket from every peer and reduces locally (latency-bound sizes)
//   two-shot : reduce-scatter by peer reads of the rank's own 1/W slice (written back in place), then an
//              all-gather by peer reads of every owner's reduced slice, with the SGD update applied on the fly
//   multimem : NVLS path (multimem.ld_reduce / multimem.st through a multicast mapping), when one is bound
//   broadcast: peers pull a region of the source rank's window (parameter / BN-buffer sync)
//
// Reductions run in fixed rank order, so results are bit-identical on every rank and run to run.
// Replaces ProcessGroupNCCL allreduce/broadcast used by DDP's Reducer and the foreach SGD step of the reference
// (reference resnet/main.py:74,80,123,124).
#include "b200ddp.h"
#include "ptx.cuh"
#include "xgpu.cuh"

namespace b2 {

constexpr int kArThreads = 512;
constexpr int kOneShotMaxVecPerThread = 8;

struct SgdArgs {
  float* p32;                 // nullptr -> reduce only
  float* mom;
  __nv_bfloat16* shadow;
  float lr, mu, wd;
  int first;
  const float* hyper;         // optional device block {lr, momentum, weight_decay}: overrides the by-value copies
};
// Hyper-parameters come from device memory when a block is given, so a learning-rate schedule changes three
// floats on the device instead of the arguments baked into a captured CUDA graph.
B2_DEVICE SgdArgs load_hyper(SgdArgs a) {
  if (a.p32 && a.hyper) {
    a.lr = a.hyper[0]; a.mu = a.hyper[1]; a.wd = a.hyper[2];
  }
  return a;
}

B2_DEVICE float4 ld_peer_f4(const float* p) {
  uint4 u = ld_volatile_v4(p);
  return make_float4(__uint_as_float(u.x), __uint_as_float(u.y), __uint_as_float(u.z), __uint_as_float(u.w));
}
B2_DEVICE void st_f4(float* p, float4 v) {
  st_v4(p, make_uint4(__float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w)));
}

// g' = g + wd*p ; buf = first ? g' : mu*buf + g' ; p -= lr*buf ; shadow = bf16(p)      (4 elements)
B2_DEVICE void sgd_apply4(const SgdArgs& a, long long vi, float4 g) {
  float4 pp = reinterpret_cast<float4*>(a.p32)[vi];
  float4 bb = a.first ? make_float4(0.f, 0.f, 0.f, 0.f) : reinterpret_cast<float4*>(a.mom)[vi];
  const float gx = fmaf(a.wd, pp.x, g.x), gy = fmaf(a.wd, pp.y, g.y), gz = fmaf(a.wd, pp.z, g.z),
              gw = fmaf(a.wd, pp.w, g.w);
  bb.x = a.first ? gx : fmaf(a.mu, bb.x, gx); bb.y = a.first ? gy : fmaf(a.mu, bb.y, gy);
  bb.z = a.first ? gz : fmaf(a.mu, bb.z, gz); bb.w = a.first ? gw : fmaf(a.mu, bb.w, gw);
  pp.x -= a.lr * bb.x; pp.y -= a.lr * bb.y; pp.z -= a.lr * bb.z; pp.w -= a.lr * bb.w;
  reinterpret_cast<float4*>(a.p32)[vi] = pp;
  reinterpret_cast<float4*>(a.mom)[vi] = bb;
  if (a.shadow) {
    uint2 s;
    s.x = pack_bf16x2(pp.x, pp.y); s.y = pack_bf16x2(pp.z, pp.w);
    reinterpret_cast<uint2*>(a.shadow)[vi] = s;
  }
}

constexpr int kUnroll = 4;  // independent 16-byte peer loads in flight per thread and peer (NVLink latency ~2 us)

// Reduce U vectors (vi, vi+stride, ...) over W ranks: issue every load first, then add in rank order.
// W is a compile-time bound on the world size (2, 4, 8) so the staging registers are sized exactly.
template <int W, int U>
B2_DEVICE void reduce_peers_u(const XgpuCtx& xg, long long off_f, long long vi, long long stride, long long vend,
                              float scale, float4 (&out)[U]) {
  float4 v[W][U];
#pragma unroll
  for (int r = 0; r < W; ++r) {
    if (r < xg.world) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = vi + u * stride;
        v[r][u] = (i < vend) ? ld_peer_f4(xg.data[r] + off_f + i * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int r = 0; r < W; ++r) {
      if (r < xg.world) {
        acc.x += v[r][u].x; acc.y += v[r][u].y; acc.z += v[r][u].z; acc.w += v[r][u].w;
      }
    }
    acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
    out[u] = acc;
  }
}

B2_DEVICE float4 reduce_peers(const XgpuCtx& xg, long long off_f, long long vi, float scale) {
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int r = 0; r < B2_MAX_WORLD; ++r) {
    if (r < xg.world) {
      const float4 v = ld_peer_f4(xg.data[r] + off_f + vi * 4);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
  return acc;
}

// ------------------------------------------------------------------------------------------------ one-shot
__global__ void __launch_bounds__(kArThreads)
    allreduce_oneshot_kernel(XgpuCtx xg, long long off_f, long long nvec, float scale, SgdArgs sgd_in, int write_grad) {
  const SgdArgs sgd = load_hyper(sgd_in);
  debug_skew(xg.rank, 0);
  xgpu_barrier(xg, blockIdx.x);
  debug_skew(xg.rank, 1);
  float4 keep[kOneShotMaxVecPerThread];
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long v0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
#pragma unroll
  for (int k = 0; k < kOneShotMaxVecPerThread; ++k) {
    const long long vi = v0 + k * stride;
    if (vi < nvec) {
      keep[k] = reduce_peers(xg, off_f, vi, scale);
      if (sgd.p32) sgd_apply4(sgd, vi, keep[k]);
    }
  }
  xgpu_barrier(xg, blockIdx.x);  // every peer has finished reading this rank's raw gradients
  if (write_grad) {
    float* mine = xg.data[xg.rank] + off_f;
#pragma unroll
    for (int k = 0; k < kOneShotMaxVecPerThread; ++k) {
      const long long vi = v0 + k * stride;
      if (vi < nvec) st_f4(mine + vi * 4, keep[k]);
    }
  }
}

// ------------------------------------------------------------------------------------------------ two-shot
template <int W>
__global__ void __launch_bounds__(kArThreads)
    allreduce_twoshot_kernel(XgpuCtx xg, long long off_f, long long nvec, float scale, SgdArgs sgd_in, int write_grad) {
  const SgdArgs sgd = load_hyper(sgd_in);
  constexpr int U = (W <= 4) ? 4 : 2;  // peer loads in flight per thread: W*U vectors
  const long long slice = (nvec + xg.world - 1) / xg.world;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long t0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  float* mine = xg.data[xg.rank] + off_f;

  debug_skew(xg.rank, 0);
  xgpu_barrier(xg, blockIdx.x);
  debug_skew(xg.rank, 1);
  {  // reduce-scatter: this rank owns [s0, s1)
    const long long s0 = slice * xg.rank, s1 = min(nvec, s0 + slice);
    for (long long vi = s0 + t0; vi < s1; vi += stride * U) {
      float4 r[U];
      reduce_peers_u<W, U>(xg, off_f, vi, stride, s1, scale, r);
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (vi + u * stride < s1) st_f4(mine + (vi + u * stride) * 4, r[u]);
    }
  }
  xgpu_barrier(xg, blockIdx.x);
  debug_skew(xg.rank, 2);
  // all-gather (+ update): pull every owner's reduced slice; own slice is already local
  for (int q = 0; q < xg.world; ++q) {
    const int owner = (xg.rank + q) % xg.world;  // start with the local slice, then stagger peers
    const long long s0 = slice * owner, s1 = min(nvec, s0 + slice);
    const float* src = xg.data[owner] + off_f;
    for (long long vi = s0 + t0; vi < s1; vi += stride * kUnroll) {
      float4 g[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u)
        if (vi + u * stride < s1) g[u] = ld_peer_f4(src + (vi + u * stride) * 4);
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = vi + u * stride;
        if (i < s1) {
          if (sgd.p32) sgd_apply4(sgd, i, g[u]);
          if (write_grad && owner != xg.rank) st_f4(mine + i * 4, g[u]);
        }
      }
    }
  }
  xgpu_barrier(xg, blockIdx.x);  // peers are done reading this rank's slice before it is overwritten again
}

// ------------------------------------------------------------------------------------------------ NVLS multimem
// mc: multicast mapping of the same window (all ranks' copies at once). ld_reduce pulls a switch-reduced value,
// st pushes to every rank.
B2_DEVICE float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
B2_DEVICE void multimem_st_f4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
// Every cross-GPU barrier pairs the SAME-NUMBERED CTA of each rank and nothing else, so after a barrier CTA b may
// only consume what CTA b (of any rank) produced before it. All loops below therefore walk a slice as
// s0(owner) + (b * blockDim + t) + k * stride: the vector set CTA b of the owner stored in phase 1 is exactly the
// set CTA b of every rank loads in phase 2 (csrc/xgpu.cuh; B2_DEBUG_SKEW delays some CTAs to prove it, tests).
__global__ void __launch_bounds__(kArThreads)
    allreduce_multimem_kernel(XgpuCtx xg, float* mc_base, long long off_f, long long nvec, float scale, SgdArgs sgd_in) {
  const SgdArgs sgd = load_hyper(sgd_in);
  const long long slice = (nvec + xg.world - 1) / xg.world;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long t0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  float* mc = mc_base + off_f;
  float* mine = xg.data[xg.rank] + off_f;
  debug_skew(xg.rank, 0);
  xgpu_barrier(xg, blockIdx.x);
  debug_skew(xg.rank, 1);
  {  // own slice: switch-reduced load, scale, multicast store to every rank; the SGD update of this slice is
     // applied straight from the registers (no second read)
    // four switch-reduced loads in flight per thread. Eight were tried against torch's multimem_all_reduce_ (which is
    // 15-25 % faster between 8 and 32 MB): no gain (8.2 MB 43.0 vs 41.6 us, 31.5 MB 118.8 vs 105.0, 100 MB 273.7 vs
    // 258.9 on two different 8-GPU nodes) - the kernel is not bound by loads in flight at 64 CTAs.
    constexpr int kMcUnroll = kUnroll;
    const long long s0 = slice * xg.rank, s1 = min(nvec, s0 + slice);
    for (long long vi = s0 + t0; vi < s1; vi += stride * kMcUnroll) {
      float4 g[kMcUnroll];
#pragma unroll
      for (int u = 0; u < kMcUnroll; ++u)
        if (vi + u * stride < s1) g[u] = multimem_ld_reduce_f4(mc + (vi + u * stride) * 4);
#pragma unroll
      for (int u = 0; u < kMcUnroll; ++u) {
        if (vi + u * stride < s1) {
          g[u].x *= scale; g[u].y *= scale; g[u].z *= scale; g[u].w *= scale;
          multimem_st_f4(mc + (vi + u * stride) * 4, g[u]);  // lands in every rank's buffer
          if (sgd.p32) sgd_apply4(sgd, vi + u * stride, g[u]);
        }
      }
    }
  }
  xgpu_barrier(xg, blockIdx.x);
  debug_skew(xg.rank, 2);
  if (sgd.p32) {
    // the other owners' slices of the local copy, filled by their multicast stores; read through L2 (never cached
    // in this SM's L1), owner by owner in the order their same-numbered CTAs wrote them
    for (int q = 1; q < xg.world; ++q) {
      const int owner = (xg.rank + q) % xg.world;
      const long long s0 = slice * owner, s1 = min(nvec, s0 + slice);
      for (long long vi = s0 + t0; vi < s1; vi += stride * kUnroll) {
        uint4 u[kUnroll];
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
          if (vi + k * stride < s1) u[k] = ld_cg_v4(mine + (vi + k * stride) * 4);
#pragma unroll
        for (int k = 0; k < kUnroll; ++k)
          if (vi + k * stride < s1)
            sgd_apply4(sgd, vi + k * stride,
                       make_float4(__uint_as_float(u[k].x), __uint_as_float(u[k].y), __uint_as_float(u[k].z),
                                   __uint_as_float(u[k].w)));
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ broadcast
__global__ void __launch_bounds__(kArThreads)
    broadcast_kernel(XgpuCtx xg, long long off_bytes, long long nvec16, int src, uint4* dst_override) {
  xgpu_barrier(xg, blockIdx.x);
  if (xg.rank != src || dst_override) {
    const uint4* s = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(xg.data[src]) + off_bytes);
    uint4* d = dst_override ? dst_override
                            : reinterpret_cast<uint4*>(reinterpret_cast<char*>(xg.data[xg.rank]) + off_bytes);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec16; i += stride * kUnroll) {
      uint4 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u)
        if (i + u * stride < nvec16) v[u] = ld_volatile_v4(s + i + u * stride);
#pragma unroll
      for (int u = 0; u < kUnroll; ++u)
        if (i + u * stride < nvec16) st_v4(d + i + u * stride, v[u]);
    }
  }
  xgpu_barrier(xg, blockIdx.x);
}

// Multicast broadcast: the source rank pushes the region through the NVSwitch multicast address, one store
// reaches every rank's window (broadcast_mc).
__global__ void __launch_bounds__(kArThreads)
    broadcast_mc_kernel(XgpuCtx xg, char* mc_base, long long off_bytes, long long nvec16, int src) {
  xgpu_barrier(xg, blockIdx.x);
  if (xg.rank == src) {
    const uint4* s = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(xg.data[src]) + off_bytes);
    float* d = reinterpret_cast<float*>(mc_base + off_bytes);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec16; i += stride) {
      const uint4 v = ld_cg_v4(s + i);
      multimem_st_f4(d + i * 4, make_float4(__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z),
                                            __uint_as_float(v.w)));
    }
  }
  xgpu_barrier(xg, blockIdx.x);
}

// ------------------------------------------------------------------------------------------------ bf16 wire
// Same algorithms with the gradients travelling as bf16 (half the NVLink bytes; the counterpart of torch's
// bf16_compress_hook): each rank first casts its fp32 bucket into a bf16 staging area of the symmetric window,
// peers' values are accumulated in fp32, the average is rounded to bf16 once, and the SGD epilogue / the fp32
// gradient written back use that value on every rank (so replicas stay bit-identical).
struct Bf16x8 {
  float v[8];
};
B2_DEVICE Bf16x8 unpack_bf16x8(uint4 u) {
  Bf16x8 r;
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = b.x; r.v[3] = b.y; r.v[4] = c.x; r.v[5] = c.y; r.v[6] = d.x; r.v[7] = d.y;
  return r;
}
B2_DEVICE uint4 pack_bf16x8(const Bf16x8& r) {
  return make_uint4(pack_bf16x2(r.v[0], r.v[1]), pack_bf16x2(r.v[2], r.v[3]), pack_bf16x2(r.v[4], r.v[5]),
                    pack_bf16x2(r.v[6], r.v[7]));
}
// vector index v8 covers elements [8*v8, 8*v8+8)
B2_DEVICE void consume_avg8(const SgdArgs& sgd, float* g32_out, long long v8, const Bf16x8& g) {
  if (sgd.p32) {
    sgd_apply4(sgd, 2 * v8, make_float4(g.v[0], g.v[1], g.v[2], g.v[3]));
    sgd_apply4(sgd, 2 * v8 + 1, make_float4(g.v[4], g.v[5], g.v[6], g.v[7]));
  }
  if (g32_out) {
    st_f4(g32_out + v8 * 8, make_float4(g.v[0], g.v[1], g.v[2], g.v[3]));
    st_f4(g32_out + v8 * 8 + 4, make_float4(g.v[4], g.v[5], g.v[6], g.v[7]));
  }
}
B2_DEVICE void cast_local_to_wire(const float* __restrict__ g32, __nv_bfloat16* wire, long long nvec8,
                                  long long t0, long long stride) {
  for (long long v = t0; v < nvec8; v += stride) {
    const float4 a = *reinterpret_cast<const float4*>(g32 + v * 8);
    const float4 b = *reinterpret_cast<const float4*>(g32 + v * 8 + 4);
    st_v4(wire + v * 8, make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y),
                                   pack_bf16x2(b.z, b.w)));
  }
}
// Same cast, walked slice by slice (s0(owner) + t0 + k * stride): what CTA b of an owner reduces after the first
// barrier is then exactly what CTA b of every rank cast before it.
B2_DEVICE void cast_local_to_wire_sliced(const float* __restrict__ g32, __nv_bfloat16* wire, long long nvec8,
                                         long long slice, int world, long long t0, long long stride) {
  for (int q = 0; q < world; ++q) {
    const long long s0 = slice * q, s1 = min(nvec8, s0 + slice);
    for (long long v = s0 + t0; v < s1; v += stride) {
      const float4 a = *reinterpret_cast<const float4*>(g32 + v * 8);
      const float4 b = *reinterpret_cast<const float4*>(g32 + v * 8 + 4);
      st_v4(wire + v * 8, make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y),
                                     pack_bf16x2(b.z, b.w)));
    }
  }
}
template <int W>
B2_DEVICE Bf16x8 reduce_wire(const XgpuCtx& xg, long long off_h, long long v8, float scale) {
  uint4 u[W];
#pragma unroll
  for (int r = 0; r < W; ++r)
    if (r < xg.world) u[r] = ld_volatile_v4(reinterpret_cast<const __nv_bfloat16*>(xg.data[r]) + off_h + v8 * 8);
  Bf16x8 acc;
#pragma unroll
  for (int j = 0; j < 8; ++j) acc.v[j] = 0.f;
#pragma unroll
  for (int r = 0; r < W; ++r) {
    if (r < xg.world) {
      const Bf16x8 x = unpack_bf16x8(u[r]);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc.v[j] += x.v[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) acc.v[j] *= scale;
  return unpack_bf16x8(pack_bf16x8(acc));  // round once: every consumer sees the same bf16 value
}

template <int W>
__global__ void __launch_bounds__(kArThreads)
    allreduce_oneshot_bf16_kernel(XgpuCtx xg, long long off_h, long long nvec8, float scale, const float* g32_in,
                                  float* g32_out, SgdArgs sgd_in) {
  const SgdArgs sgd = load_hyper(sgd_in);
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long t0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  debug_skew(xg.rank, 0);
  cast_local_to_wire(g32_in, reinterpret_cast<__nv_bfloat16*>(xg.data[xg.rank]) + off_h, nvec8, t0, stride);
  xgpu_barrier(xg, blockIdx.x);
  debug_skew(xg.rank, 1);
  for (long long v = t0; v < nvec8; v += stride) consume_avg8(sgd, g32_out, v, reduce_wire<W>(xg, off_h, v, scale));
  xgpu_barrier(xg, blockIdx.x);  // peers are done with this rank's staging area
}

template <int W>
__global__ void __launch_bounds__(kArThreads)
    allreduce_twoshot_bf16_kernel(XgpuCtx xg, long long off_h, long long nvec8, float scale, const float* g32_in,
                                  float* g32_out, SgdArgs sgd_in) {
  const SgdArgs sgd = load_hyper(sgd_in);
  const long long slice = (nvec8 + xg.world - 1) / xg.world;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long t0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  __nv_bfloat16* mine = reinterpret_cast<__nv_bfloat16*>(xg.data[xg.rank]) + off_h;
  debug_skew(xg.rank, 0);
  cast_local_to_wire_sliced(g32_in, mine, nvec8, slice, xg.world, t0, stride);
  xgpu_barrier(xg, blockIdx.x);
  debug_skew(xg.rank, 1);
  {
    const long long s0 = slice * xg.rank, s1 = min(nvec8, s0 + slice);
    for (long long v = s0 + t0; v < s1; v += stride) st_v4(mine + v * 8, pack_bf16x8(reduce_wire<W>(xg, off_h, v, scale)));
  }
  xgpu_barrier(xg, blockIdx.x);
  debug_skew(xg.rank, 2);
  for (int q = 0; q < xg.world; ++q) {
    const int owner = (xg.rank + q) % xg.world;
    const long long s0 = slice * owner, s1 = min(nvec8, s0 + slice);
    const __nv_bfloat16* src = reinterpret_cast<const __nv_bfloat16*>(xg.data[owner]) + off_h;
    for (long long v = s0 + t0; v < s1; v += stride * 2) {
      uint4 u0 = ld_volatile_v4(src + v * 8), u1 = make_uint4(0, 0, 0, 0);
      const bool two = (v + stride) < s1;
      if (two) u1 = ld_volatile_v4(src + (v + stride) * 8);
      consume_avg8(sgd, g32_out, v, unpack_bf16x8(u0));
      if (two) consume_avg8(sgd, g32_out, v + stride, unpack_bf16x8(u1));
    }
  }
  xgpu_barrier(xg, blockIdx.x);
}

B2_DEVICE uint4 multimem_ld_reduce_bf16x8(const __nv_bfloat16* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
B2_DEVICE void multimem_st_bf16x8(__nv_bfloat16* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}
__global__ void __launch_bounds__(kArThreads)
    allreduce_multimem_bf16_kernel(XgpuCtx xg, __nv_bfloat16* mc_base, long long off_h, long long nvec8, float scale,
                                   const float* g32_in, float* g32_out, SgdArgs sgd_in) {
  const SgdArgs sgd = load_hyper(sgd_in);
  const long long slice = (nvec8 + xg.world - 1) / xg.world;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long t0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  __nv_bfloat16* mine = reinterpret_cast<__nv_bfloat16*>(xg.data[xg.rank]) + off_h;
  __nv_bfloat16* mc = mc_base + off_h;
  debug_skew(xg.rank, 0);
  cast_local_to_wire_sliced(g32_in, mine, nvec8, slice, xg.world, t0, stride);
  xgpu_barrier(xg, blockIdx.x);
  debug_skew(xg.rank, 1);
  {
    const long long s0 = slice * xg.rank, s1 = min(nvec8, s0 + slice);
    for (long long v = s0 + t0; v < s1; v += stride) {
      Bf16x8 g = unpack_bf16x8(multimem_ld_reduce_bf16x8(mc + v * 8));
#pragma unroll
      for (int j = 0; j < 8; ++j) g.v[j] *= scale;
      const uint4 packed = pack_bf16x8(g);
      multimem_st_bf16x8(mc + v * 8, packed);
      consume_avg8(sgd, g32_out, v, unpack_bf16x8(packed));   // own slice: from the registers, rounded once
    }
  }
  xgpu_barrier(xg, blockIdx.x);
  debug_skew(xg.rank, 2);
  for (int q = 1; q < xg.world; ++q) {   // owner-relative: see allreduce_multimem_kernel
    const int owner = (xg.rank + q) % xg.world;
    const long long s0 = slice * owner, s1 = min(nvec8, s0 + slice);
    for (long long v = s0 + t0; v < s1; v += stride) consume_avg8(sgd, g32_out, v, unpack_bf16x8(ld_cg_v4(mine + v * 8)));
  }
}

// Plain device-side barrier (one CTA per rank): orders everything before it on all ranks' streams.
__global__ void barrier_kernel(XgpuCtx xg, int channel) { xgpu_barrier(xg, channel); }

}  // namespace b2

using namespace b2;

B2_DEFINE_SPIN_TIMEOUT_SETTER(b2_set_spin_timeout_ns_allreduce)
B2_DEFINE_DEBUG_SKEW_SETTER(b2_set_debug_skew_ns_allreduce)

static int clamp_blocks(int nblocks, long long nvec) {
  long long need = (nvec + kArThreads - 1) / kArThreads;
  if (nblocks <= 0) nblocks = 64;
  if (nblocks > B2_XGPU_CHANNELS) nblocks = B2_XGPU_CHANNELS;
  if (need < nblocks) nblocks = (int)(need < 1 ? 1 : need);
  return nblocks;
}

extern "C" long long b2_allreduce_oneshot_max_elems(int nblocks) {
  if (nblocks <= 0 || nblocks > B2_XGPU_CHANNELS) nblocks = B2_XGPU_CHANNELS;
  return (long long)nblocks * kArThreads * kOneShotMaxVecPerThread * 4;
}

extern "C" int b2_allreduce(const b2_xgpu_ctx* ctx, void* mc_base, long long offset_bytes, long long n, int algo,
                            float scale, float* p32, float* mom, void* shadow_bf16, float lr, float momentum,
                            float weight_decay, int first_step, const float* hyper_dev, int write_grad, int nblocks,
                            void* stream) {
  if (!ctx || n <= 0 || (n & 3) || (offset_bytes & 15)) return cudaErrorInvalidValue;
  XgpuCtx xg = xgpu_from_c(ctx);
  const long long nvec = n >> 2;
  const long long off_f = offset_bytes >> 2;
  SgdArgs sgd{p32, mom, (__nv_bfloat16*)shadow_bf16, lr, momentum, weight_decay, first_step, hyper_dev};
  cudaStream_t st = (cudaStream_t)stream;
  if (algo == B2_AR_AUTO) {
    // One-shot moves (W-1)*S inbound per GPU but has two barriers and no second pass: wins only when small.
    algo = (n * 4 <= (512 << 10)) ? B2_AR_ONESHOT : B2_AR_TWOSHOT;
    // In-switch reduction pays off once several peers would otherwise be read one by one; with two ranks the
    // plain peer read moves the same bytes without the extra local pass (measured, tools/comm_bench.py).
    if (mc_base && xg.world > 2 && n * 4 >= (32 << 10)) algo = B2_AR_MULTIMEM;
  }
  if (algo == B2_AR_ONESHOT) {
    int nb = nblocks > 0 ? nblocks : B2_XGPU_CHANNELS;
    if (nb > B2_XGPU_CHANNELS) nb = B2_XGPU_CHANNELS;
    if (n > b2_allreduce_oneshot_max_elems(nb)) algo = B2_AR_TWOSHOT;
    else {
      // keep per-thread work balanced: as many blocks as needed for <= 8 vec4 per thread, at least 1
      long long need = (nvec + (long long)kArThreads * kOneShotMaxVecPerThread - 1) /
                       ((long long)kArThreads * kOneShotMaxVecPerThread);
      long long spread = (nvec + kArThreads - 1) / kArThreads;
      int b = (int)(spread < nb ? spread : nb);
      if (b < need) b = (int)need;
      if (b < 1) b = 1;
      allreduce_oneshot_kernel<<<b, kArThreads, 0, st>>>(xg, off_f, nvec, scale, sgd, write_grad);
      return cudaGetLastError();
    }
  }
  if (algo == B2_AR_MULTIMEM) {
    if (!mc_base) return cudaErrorInvalidValue;
    allreduce_multimem_kernel<<<clamp_blocks(nblocks, nvec / xg.world + 1), kArThreads, 0, st>>>(
        xg, (float*)mc_base, off_f, nvec, scale, sgd);
    return cudaGetLastError();
  }
  const int nb = clamp_blocks(nblocks, nvec / xg.world + 1);
  if (xg.world <= 2) allreduce_twoshot_kernel<2><<<nb, kArThreads, 0, st>>>(xg, off_f, nvec, scale, sgd, write_grad);
  else if (xg.world <= 4) allreduce_twoshot_kernel<4><<<nb, kArThreads, 0, st>>>(xg, off_f, nvec, scale, sgd, write_grad);
  else allreduce_twoshot_kernel<8><<<nb, kArThreads, 0, st>>>(xg, off_f, nvec, scale, sgd, write_grad);
  return cudaGetLastError();
}

extern "C" int b2_broadcast(const b2_xgpu_ctx* ctx, long long offset_bytes, long long nbytes, int src,
                            void* dst_override, int nblocks, void* stream) {
  if (!ctx || nbytes <= 0 || (nbytes & 15) || (offset_bytes & 15)) return cudaErrorInvalidValue;
  XgpuCtx xg = xgpu_from_c(ctx);
  const long long nvec = nbytes >> 4;
  broadcast_kernel<<<clamp_blocks(nblocks, nvec), kArThreads, 0, (cudaStream_t)stream>>>(xg, offset_bytes, nvec, src,
                                                                                        (uint4*)dst_override);
  return cudaGetLastError();
}

extern "C" int b2_allreduce_bf16(const b2_xgpu_ctx* ctx, void* mc_base, long long wire_offset_bytes, long long n,
                                 int algo, float scale, const float* g32_in, float* g32_out, float* p32, float* mom,
                                 void* shadow_bf16, float lr, float momentum, float weight_decay, int first_step,
                                 const float* hyper_dev, int nblocks, void* stream) {
  if (!ctx || n <= 0 || (n & 7) || (wire_offset_bytes & 15) || !g32_in) return cudaErrorInvalidValue;
  XgpuCtx xg = xgpu_from_c(ctx);
  const long long nvec8 = n >> 3;
  const long long off_h = wire_offset_bytes >> 1;
  SgdArgs sgd{p32, mom, (__nv_bfloat16*)shadow_bf16, lr, momentum, weight_decay, first_step, hyper_dev};
  cudaStream_t st = (cudaStream_t)stream;
  if (algo == B2_AR_AUTO) {
    algo = (n * 2 <= (256 << 10)) ? B2_AR_ONESHOT : B2_AR_TWOSHOT;
    if (mc_base && xg.world > 2 && n * 2 >= (32 << 10)) algo = B2_AR_MULTIMEM;
  }
  const int nb1 = clamp_blocks(nblocks, nvec8);
#define B2_BF16_DISPATCH(KERN, NB)                                                                          \
  do {                                                                                                      \
    if (xg.world <= 2) KERN<2><<<NB, kArThreads, 0, st>>>(xg, off_h, nvec8, scale, g32_in, g32_out, sgd);   \
    else if (xg.world <= 4) KERN<4><<<NB, kArThreads, 0, st>>>(xg, off_h, nvec8, scale, g32_in, g32_out, sgd); \
    else KERN<8><<<NB, kArThreads, 0, st>>>(xg, off_h, nvec8, scale, g32_in, g32_out, sgd);                 \
  } while (0)
  if (algo == B2_AR_ONESHOT) {
    B2_BF16_DISPATCH(allreduce_oneshot_bf16_kernel, nb1);
  } else if (algo == B2_AR_MULTIMEM) {
    if (!mc_base) return cudaErrorInvalidValue;
    allreduce_multimem_bf16_kernel<<<nb1, kArThreads, 0, st>>>(xg, (__nv_bfloat16*)mc_base, off_h, nvec8, scale,
                                                               g32_in, g32_out, sgd);
  } else {
    B2_BF16_DISPATCH(allreduce_twoshot_bf16_kernel, nb1);
  }
#undef B2_BF16_DISPATCH
  return cudaGetLastError();
}

extern "C" int b2_broadcast_mc(const b2_xgpu_ctx* ctx, void* mc_base, long long offset_bytes, long long nbytes,
                               int src, int nblocks, void* stream) {
  if (!ctx || !mc_base || nbytes <= 0 || (nbytes & 15) || (offset_bytes & 15)) return cudaErrorInvalidValue;
  XgpuCtx xg = xgpu_from_c(ctx);
  const long long nvec = nbytes >> 4;
  broadcast_mc_kernel<<<clamp_blocks(nblocks, nvec), kArThreads, 0, (cudaStream_t)stream>>>(
      xg, (char*)mc_base, offset_bytes, nvec, src);
  return cudaGetLastError();
}

extern "C" int b2_xgpu_barrier(const b2_xgpu_ctx* ctx, int channel, void* stream) {
  if (!ctx) return cudaErrorInvalidValue;
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(xgpu_from_c(ctx), channel);
  return cudaGetLastError();
}
