// BatchNorm (training + eval) fused with ReLU and the residual add, forward and backward, on NHWC bf16.
// Memory-bound streaming kernels: 16-byte vectors (8 channels) per thread, fp32 statistics, two-pass
// deterministic reductions (per-CTA partials -> finalize).  The finalize kernels optionally reduce the
// statistics across GPUs through peer-mapped symmetric memory (SyncBatchNorm) without a separate collective.
//
// Replaces aten::native_batch_norm(+_backward), relu_, add_, threshold_backward reached by the reference through
// torchvision's BasicBlock/Bottleneck (reference resnet/main.py:76,121,123) and, for the cross-GPU variant,
// torch.nn.SyncBatchNorm's all_gather / all_reduce of statistics.
#include "b200ddp.h"
#include "ptx.cuh"
#include "xgpu.cuh"

#include <cstdio>
#include <mutex>

namespace b2 {

constexpr int kBnThreads = 256;
#ifndef B2_BN_BLOCKS_PER_SM
#define B2_BN_BLOCKS_PER_SM 4
#endif
// Capacity of the partial-sum scratch in rows (sized for a 148-SM B200; the launch grid is derived from the
// device's actual SM count and clamped to this).
constexpr int kMaxPartialBlocks = 148 * B2_BN_BLOCKS_PER_SM;

struct Vec8 {
  float v[8];
};
B2_DEVICE Vec8 load_bf16x8(const __nv_bfloat16* p) {
  uint4 u = *reinterpret_cast<const uint4*>(p);
  Vec8 r;
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = b.x; r.v[3] = b.y; r.v[4] = c.x; r.v[5] = c.y; r.v[6] = d.x; r.v[7] = d.y;
  return r;
}
B2_DEVICE void store_bf16x8(__nv_bfloat16* p, const Vec8& r) {
  uint4 u;
  u.x = pack_bf16x2(r.v[0], r.v[1]); u.y = pack_bf16x2(r.v[2], r.v[3]);
  u.z = pack_bf16x2(r.v[4], r.v[5]); u.w = pack_bf16x2(r.v[6], r.v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}
B2_DEVICE Vec8 load_f32x8(const float* p) {
  Vec8 r;
  float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}

// Reduce per-thread 8-channel accumulators over the rows handled by one CTA and write partial[blk][which][C].
template <int NACC>
B2_DEVICE void block_reduce_store(float (&acc)[NACC][8], float* smem, float* partial, int C, int C8, int RP, int cv,
                                  int r, int nblk_stride /* = C */) {
  // smem layout: [NACC][RP][C]
  for (int a = 0; a < NACC; ++a)
#pragma unroll
    for (int j = 0; j < 8; ++j) smem[(a * RP + r) * C + cv * 8 + j] = acc[a][j];
  __syncthreads();
  for (int idx = threadIdx.x; idx < NACC * C; idx += blockDim.x) {
    const int a = idx / C, c = idx - a * C;
    float s = 0.f;
    for (int rr = 0; rr < RP; ++rr) s += smem[(a * RP + rr) * C + c];
    partial[(size_t)blockIdx.x * NACC * C + idx] = s;
  }
}

// ------------------------------------------------------------------------------------------------ forward stats
__global__ void __launch_bounds__(kBnThreads, B2_BN_BLOCKS_PER_SM) bn_stats_kernel(const __nv_bfloat16* __restrict__ y, float* partial,
                                                              long long M, int C, long long rows_per_blk) {
  extern __shared__ float smem[];
  const int C8 = C >> 3, RP = kBnThreads / C8;
  const int cv = threadIdx.x % C8, r = threadIdx.x / C8;
  float acc[2][8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[0][j] = acc[1][j] = 0.f;
  // Blocks interleave at row-group granularity (block b takes groups b, b+grid, ...): at any moment the whole
  // grid streams one contiguous window of the tensor, which keeps HBM pages open (a per-block contiguous range
  // costs ~30% of the bandwidth). The assignment is static, so the result stays deterministic.
  (void)rows_per_blk;
  if (r < RP) {
#pragma unroll 4
    for (long long row = (long long)blockIdx.x * RP + r; row < M; row += (long long)gridDim.x * RP) {
      Vec8 v = load_bf16x8(y + row * C + cv * 8);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        acc[0][j] += v.v[j];
        acc[1][j] += v.v[j] * v.v[j];
      }
    }
    block_reduce_store<2>(acc, smem, partial, C, C8, RP, cv, r, C);
  } else {
    __syncthreads();
  }
}

// Finalize: partial sums -> mean / invstd / running stats / fused scale+shift.  With world > 1 the per-GPU
// totals (sum, sumsq, count) are exchanged through symmetric memory inside this kernel (SyncBatchNorm).
// One CTA per 8 channels: 32 row-groups x 8 channels read the partials sector-coalesced, then a fixed-order
// shared-memory reduction (deterministic).
constexpr int kFinCh = 8;
constexpr int kFinGroups = 32;
constexpr int kFinThreads = kFinCh * kFinGroups;

B2_DEVICE void fin_reduce2(const float* __restrict__ partial, int nblk, int C, int c, int rg, int cg, float* sm,
                           float& out0, float& out1) {
  float a0 = 0.f, a1 = 0.f;
#pragma unroll 4
  for (int b = rg; b < nblk; b += kFinGroups) {
    a0 += partial[(size_t)b * 2 * C + c];
    a1 += partial[(size_t)b * 2 * C + C + c];
  }
  sm[rg * kFinCh + cg] = a0;
  sm[kFinThreads + rg * kFinCh + cg] = a1;
  __syncthreads();
  if (rg == 0) {
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int g = 0; g < kFinGroups; ++g) {
      s0 += sm[g * kFinCh + cg];
      s1 += sm[kFinThreads + g * kFinCh + cg];
    }
    out0 = s0;
    out1 = s1;
  }
}

__global__ void __launch_bounds__(kFinThreads)
    bn_finalize_kernel(const float* __restrict__ partial, int nblk, int C, float count,
                       const float* __restrict__ gamma, const float* __restrict__ beta, float* running_mean,
                       float* running_var, long long* num_batches_tracked, float momentum, float eps,
                       float* mean_out, float* invstd_out, float* scale_out, float* shift_out, XgpuCtx xg,
                       int slot_off) {
  __shared__ float sm[2 * kFinThreads];
  const int cg = threadIdx.x % kFinCh, rg = threadIdx.x / kFinCh;
  const int c = blockIdx.x * kFinCh + cg;
  float s = 0.f, ss = 0.f;
  fin_reduce2(partial, nblk, C, c, rg, cg, sm, s, ss);
  float total = count;
  if (xg.world > 1) {
    // Publish this GPU's totals, then read everybody's in rank order (identical result on every rank).
    float* mine = xg.data[xg.rank] + slot_off;
    if (rg == 0) {
      mine[c] = s;
      mine[C + c] = ss;
    }
    if (threadIdx.x == 0) mine[2 * C + blockIdx.x] = count;
    xgpu_barrier(xg, /*channel=*/blockIdx.x);
    s = 0.f; ss = 0.f; total = 0.f;
    if (rg == 0) {
      // all 3 * W peer loads are issued before the first one is consumed (one NVLink round trip, not W of them),
      // then added in rank order
      float v0[B2_MAX_WORLD], v1[B2_MAX_WORLD], v2[B2_MAX_WORLD];
#pragma unroll
      for (int rk = 0; rk < B2_MAX_WORLD; ++rk) {
        if (rk < xg.world) {
          const float* peer = xg.data[rk] + slot_off;
          v0[rk] = ld_relaxed_sys_f32(peer + c);
          v1[rk] = ld_relaxed_sys_f32(peer + C + c);
          v2[rk] = ld_relaxed_sys_f32(peer + 2 * C + blockIdx.x);
        }
      }
#pragma unroll
      for (int rk = 0; rk < B2_MAX_WORLD; ++rk) {
        if (rk < xg.world) {
          s += v0[rk];
          ss += v1[rk];
          total += v2[rk];
        }
      }
    }
    // No trailing barrier: the caller alternates slot_off between two exchange areas from call to call, and a rank
    // can only be two calls ahead of a peer after that peer has arrived at the barrier in between, i.e. after the
    // peer's previous kernel (whose reads of this area are the last ones) has completed. One device barrier per
    // BatchNorm instead of two (SyncBatchNorm of torch: a blocking all_gather + a host sync per layer).
  }
  if (rg != 0) return;
  const float mean = s / total;
  float var = ss / total - mean * mean;
  var = fmaxf(var, 0.f);
  const float invstd = rsqrtf(var + eps);
  mean_out[c] = mean;
  invstd_out[c] = invstd;
  const float sc = gamma[c] * invstd;
  scale_out[c] = sc;
  shift_out[c] = beta[c] - mean * sc;
  if (running_mean) {
    const float unbiased = total > 1.f ? var * total / (total - 1.f) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
    if (c == 0 && num_batches_tracked) *num_batches_tracked += 1;
  }
}

// Eval mode: scale/shift from running statistics.
__global__ void bn_eval_coeffs_kernel(int C, const float* gamma, const float* beta, const float* rm, const float* rv,
                                      float eps, float* scale, float* shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float sc = gamma[c] * rsqrtf(rv[c] + eps);
  scale[c] = sc;
  shift[c] = beta[c] - rm[c] * sc;
}

// ------------------------------------------------------------------------------------------------ forward apply
template <bool RELU, bool RES>
__global__ void __launch_bounds__(256) bn_apply_kernel(const __nv_bfloat16* __restrict__ y,
                                                       const __nv_bfloat16* __restrict__ res,
                                                       const float* __restrict__ scale,
                                                       const float* __restrict__ shift, __nv_bfloat16* __restrict__ z,
                                                       unsigned char* __restrict__ mask, long long nvec, int C8) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    const int cv = (int)(i % C8);
    Vec8 v = load_bf16x8(y + i * 8);
    Vec8 sc = load_f32x8(scale + cv * 8), sh = load_f32x8(shift + cv * 8);
    Vec8 o;
    if (RES) {
      Vec8 rr = load_bf16x8(res + i * 8);
#pragma unroll
      for (int j = 0; j < 8; ++j) o.v[j] = fmaf(v.v[j], sc.v[j], sh.v[j]) + rr.v[j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) o.v[j] = fmaf(v.v[j], sc.v[j], sh.v[j]);
    }
    if (RELU) {
      if (mask) {  // one bit per element: all the backward pass needs to know about z
        unsigned bits = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) bits |= (o.v[j] > 0.f ? 1u : 0u) << j;
        mask[i] = (unsigned char)bits;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) o.v[j] = fmaxf(o.v[j], 0.f);
    }
    store_bf16x8(z + i * 8, o);
  }
}

// Same pass, every thread owning one 8-channel vector (scale / shift in registers) and issuing the loads of U row
// groups back to back; the grid-stride kernel above re-reads scale / shift from L1 for every 16 bytes and reaches only
// ~58 % of the copy bandwidth on the narrow (64..512-channel) tensors.
template <bool RELU, bool RES, int U, int BPS>
__global__ void __launch_bounds__(kBnThreads, BPS)
    bn_apply_batched_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ res,
                            const float* __restrict__ scale, const float* __restrict__ shift,
                            __nv_bfloat16* __restrict__ z, unsigned char* __restrict__ mask, long long M, int C) {
  const int C8 = C >> 3, RP = kBnThreads / C8;
  const int cv = threadIdx.x % C8, r = threadIdx.x / C8;
  if (r >= RP) return;
  const Vec8 sc = load_f32x8(scale + cv * 8), sh = load_f32x8(shift + cv * 8);
  const long long step = (long long)gridDim.x * RP;
  for (long long row0 = (long long)blockIdx.x * RP + r; row0 < M; row0 += step * U) {
    uint4 uy[U], ur[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long row = row0 + u * step;
      if (row < M) {
        const long long off = row * C + cv * 8;
        uy[u] = ld_nc_v4(y + off);
        if (RES) ur[u] = ld_nc_v4(res + off);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long row = row0 + u * step;
      if (row < M) {
        const long long off = row * C + cv * 8;
        const uint32_t yw[4] = {uy[u].x, uy[u].y, uy[u].z, uy[u].w};
        uint32_t ow[4];
        unsigned bits = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float o0 = fmaf(__uint_as_float(yw[q] << 16), sc.v[2 * q], sh.v[2 * q]);
          float o1 = fmaf(__uint_as_float(yw[q] & 0xffff0000u), sc.v[2 * q + 1], sh.v[2 * q + 1]);
          if (RES) {
            const uint32_t rw = q == 0 ? ur[u].x : (q == 1 ? ur[u].y : (q == 2 ? ur[u].z : ur[u].w));
            o0 += __uint_as_float(rw << 16);
            o1 += __uint_as_float(rw & 0xffff0000u);
          }
          if (RELU) {
            bits |= (o0 > 0.f ? 1u : 0u) << (2 * q);
            bits |= (o1 > 0.f ? 1u : 0u) << (2 * q + 1);
            o0 = fmaxf(o0, 0.f);
            o1 = fmaxf(o1, 0.f);
          }
          ow[q] = pack_bf16x2(o0, o1);
        }
        if (RELU && mask) mask[off >> 3] = (unsigned char)bits;
        *reinterpret_cast<uint4*>(z + off) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ backward
// dz' = (dz_a [+ dz_b]) * (z > 0 if RELU).  Accumulates sum(dz') and sum(dz' * (y - mean)) per channel.
// GOUT: dz' (rounded to bf16, which is what the sums then use) is also written out - it is the gradient of the
// residual branch, and the apply pass then reads that one tensor instead of dz_a, dz_b and the mask again.
template <bool RELU, bool TWO, bool GOUT>
__global__ void __launch_bounds__(kBnThreads, B2_BN_BLOCKS_PER_SM)
    bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dza, const __nv_bfloat16* __restrict__ dzb,
                         const __nv_bfloat16* __restrict__ z, const unsigned char* __restrict__ mask,
                         const __nv_bfloat16* __restrict__ y, const float* __restrict__ mean, float* partial,
                         __nv_bfloat16* __restrict__ gout, long long M, int C, long long rows_per_blk) {
  extern __shared__ float smem[];
  const int C8 = C >> 3, RP = kBnThreads / C8;
  const int cv = threadIdx.x % C8, r = threadIdx.x / C8;
  float acc[2][8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[0][j] = acc[1][j] = 0.f;
  (void)rows_per_blk;
  if (r < RP) {
    Vec8 mu = load_f32x8(mean + cv * 8);
#pragma unroll 4
    for (long long row = (long long)blockIdx.x * RP + r; row < M; row += (long long)gridDim.x * RP) {
      const long long off = row * C + cv * 8;
      Vec8 g = load_bf16x8(dza + off);
      if (TWO) {
        Vec8 g2 = load_bf16x8(dzb + off);
#pragma unroll
        for (int j = 0; j < 8; ++j) g.v[j] += g2.v[j];
      }
      if (RELU) {
        if (mask) {
          const unsigned bits = mask[off >> 3];
#pragma unroll
          for (int j = 0; j < 8; ++j) g.v[j] = ((bits >> j) & 1u) ? g.v[j] : 0.f;
        } else {
          Vec8 zz = load_bf16x8(z + off);
#pragma unroll
          for (int j = 0; j < 8; ++j) g.v[j] = zz.v[j] > 0.f ? g.v[j] : 0.f;
        }
      }
      if (GOUT) {
#pragma unroll
        for (int j = 0; j < 8; ++j) g.v[j] = __bfloat162float(__float2bfloat16(g.v[j]));
        store_bf16x8(gout + off, g);
      }
      Vec8 yy = load_bf16x8(y + off);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        acc[0][j] += g.v[j];
        acc[1][j] += g.v[j] * (yy.v[j] - mu.v[j]);
      }
    }
    block_reduce_store<2>(acc, smem, partial, C, C8, RP, cv, r, C);
  } else {
    __syncthreads();
  }
}

// Same pass with the loads of U row groups issued back to back before any of them is consumed (explicit software
// batching: U * (2..3) independent 16-byte loads + U mask bytes in flight per thread). The plain loop above leaves it
// to the compiler to hoist loads across the unrolled iterations, which the 64-register cap of 4 CTAs/SM prevents:
// ncu showed the kernel waiting on long_scoreboard at 49 % occupancy and 75 % of the measured HBM peak.
// BPS = CTAs per SM the register budget is sized for (fewer CTAs, more registers, more loads in flight each).
// TWO = 2: the second incoming gradient exists at HALF resolution only - it is the input gradient of a 1x1 / stride 2
// downsample convolution, non-zero at the even pixels of the [N, qH, qW] grid - and is read from its compact
// [N, qH/2, qW/2, C] form instead of from a zero-filled full-resolution tensor.
template <bool RELU, int TWO, bool GOUT, int U, int BPS>
__global__ void __launch_bounds__(kBnThreads, BPS)
    bn_bwd_reduce_batched_kernel(const __nv_bfloat16* __restrict__ dza, const __nv_bfloat16* __restrict__ dzb,
                                 const unsigned char* __restrict__ mask, const __nv_bfloat16* __restrict__ y,
                                 const float* __restrict__ mean, float* partial, __nv_bfloat16* __restrict__ gout,
                                 long long M, int C, int qH, int qW) {
  extern __shared__ float smem[];
  const int C8 = C >> 3, RP = kBnThreads / C8;
  const int cv = threadIdx.x % C8, r = threadIdx.x / C8;
  float acc[2][8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[0][j] = acc[1][j] = 0.f;
  if (r < RP) {
    const Vec8 mu = load_f32x8(mean + cv * 8);
    const long long step = (long long)gridDim.x * RP;
    for (long long row0 = (long long)blockIdx.x * RP + r; row0 < M; row0 += step * U) {
      uint4 ua[U], ub[U], uy[U];
      unsigned bits[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long row = row0 + u * step;
        if (row < M) {
          const long long off = row * C + cv * 8;
          ua[u] = ld_nc_v4(dza + off);
          if (TWO == 1) ub[u] = ld_nc_v4(dzb + off);
          if (TWO == 2) {
            const int w = (int)(row % qW);
            const long long t = row / qW;
            const int h = (int)(t % qH);
            const long long n = t / qH;
            ub[u] = make_uint4(0u, 0u, 0u, 0u);
            if (((h | w) & 1) == 0) {
              const long long qrow = (n * ((qH + 1) >> 1) + (h >> 1)) * ((qW + 1) >> 1) + (w >> 1);
              ub[u] = ld_nc_v4(dzb + qrow * C + cv * 8);
            }
          }
          uy[u] = ld_nc_v4(y + off);
          if (RELU) bits[u] = mask[off >> 3];
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long row = row0 + u * step;
        if (row < M) {
          const long long off = row * C + cv * 8;
          float g[8], yy[8];
          {
            const float2 a0 = unpack_bf16x2(ua[u].x), a1 = unpack_bf16x2(ua[u].y), a2 = unpack_bf16x2(ua[u].z),
                         a3 = unpack_bf16x2(ua[u].w);
            g[0] = a0.x; g[1] = a0.y; g[2] = a1.x; g[3] = a1.y; g[4] = a2.x; g[5] = a2.y; g[6] = a3.x; g[7] = a3.y;
          }
          if (TWO) {
            const float2 a0 = unpack_bf16x2(ub[u].x), a1 = unpack_bf16x2(ub[u].y), a2 = unpack_bf16x2(ub[u].z),
                         a3 = unpack_bf16x2(ub[u].w);
            g[0] += a0.x; g[1] += a0.y; g[2] += a1.x; g[3] += a1.y; g[4] += a2.x; g[5] += a2.y; g[6] += a3.x;
            g[7] += a3.y;
          }
          if (RELU) {
#pragma unroll
            for (int j = 0; j < 8; ++j) g[j] = ((bits[u] >> j) & 1u) ? g[j] : 0.f;
          }
          if (GOUT) {
            uint4 o;
            o.x = pack_bf16x2(g[0], g[1]); o.y = pack_bf16x2(g[2], g[3]);
            o.z = pack_bf16x2(g[4], g[5]); o.w = pack_bf16x2(g[6], g[7]);
            *reinterpret_cast<uint4*>(gout + off) = o;
            const float2 a0 = unpack_bf16x2(o.x), a1 = unpack_bf16x2(o.y), a2 = unpack_bf16x2(o.z),
                         a3 = unpack_bf16x2(o.w);   // the sums use the rounded values, as the apply pass will
            g[0] = a0.x; g[1] = a0.y; g[2] = a1.x; g[3] = a1.y; g[4] = a2.x; g[5] = a2.y; g[6] = a3.x; g[7] = a3.y;
          }
          {
            const float2 a0 = unpack_bf16x2(uy[u].x), a1 = unpack_bf16x2(uy[u].y), a2 = unpack_bf16x2(uy[u].z),
                         a3 = unpack_bf16x2(uy[u].w);
            yy[0] = a0.x; yy[1] = a0.y; yy[2] = a1.x; yy[3] = a1.y; yy[4] = a2.x; yy[5] = a2.y; yy[6] = a3.x;
            yy[7] = a3.y;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            acc[0][j] += g[j];
            acc[1][j] += g[j] * (yy[j] - mu.v[j]);
          }
        }
      }
    }
    block_reduce_store<2>(acc, smem, partial, C, C8, RP, cv, r, C);
  } else {
    __syncthreads();
  }
}

// partials -> dgamma/dbeta (fp32, optionally accumulated into existing grads) and the three per-channel
// coefficients of  dy = k1*dz' + k2*y + k3.  With world > 1 the two sums are reduced across GPUs in-kernel.
__global__ void __launch_bounds__(kFinThreads)
    bn_bwd_finalize_kernel(const float* __restrict__ partial, int nblk, int C, float count,
                           const float* __restrict__ gamma, const float* __restrict__ mean,
                           const float* __restrict__ invstd, float* dgamma, float* dbeta, int accumulate,
                           float* coef, XgpuCtx xg, int slot_off) {
  __shared__ float sm[2 * kFinThreads];
  const int cg = threadIdx.x % kFinCh, rg = threadIdx.x / kFinCh;
  const int c = blockIdx.x * kFinCh + cg;
  float sdy = 0.f, sdx = 0.f;
  fin_reduce2(partial, nblk, C, c, rg, cg, sm, sdy, sdx);
  // Local sums give this rank's dgamma/dbeta (DDP averages them later, as torch's SyncBatchNorm does).
  const float sdy_local = sdy, sdx_local = sdx;
  float total = count;
  if (xg.world > 1) {
    float* mine = xg.data[xg.rank] + slot_off;
    if (rg == 0) {
      mine[c] = sdy;
      mine[C + c] = sdx;
    }
    if (threadIdx.x == 0) mine[2 * C + blockIdx.x] = count;
    xgpu_barrier(xg, blockIdx.x);
    sdy = 0.f; sdx = 0.f; total = 0.f;
    if (rg == 0) {
      float v0[B2_MAX_WORLD], v1[B2_MAX_WORLD], v2[B2_MAX_WORLD];   // loads first, adds after (see bn_finalize_kernel)
#pragma unroll
      for (int rk = 0; rk < B2_MAX_WORLD; ++rk) {
        if (rk < xg.world) {
          const float* peer = xg.data[rk] + slot_off;
          v0[rk] = ld_relaxed_sys_f32(peer + c);
          v1[rk] = ld_relaxed_sys_f32(peer + C + c);
          v2[rk] = ld_relaxed_sys_f32(peer + 2 * C + blockIdx.x);
        }
      }
#pragma unroll
      for (int rk = 0; rk < B2_MAX_WORLD; ++rk) {
        if (rk < xg.world) {
          sdy += v0[rk];
          sdx += v1[rk];
          total += v2[rk];
        }
      }
    }
    // (no trailing barrier: alternating exchange areas, see bn_finalize_kernel)
  }
  if (rg != 0) return;
  const float is = invstd[c], mu = mean[c], g = gamma[c];
  const float dg = sdx_local * is, db = sdy_local;
  if (dgamma) dgamma[c] = accumulate ? dgamma[c] + dg : dg;
  if (dbeta) dbeta[c] = accumulate ? dbeta[c] + db : db;
  const float A = g * is;
  const float k2 = -A * is * is * sdx / total;
  coef[c] = A;
  coef[C + c] = k2;
  coef[2 * C + c] = -A * sdy / total - k2 * mu;
}

// dy = k1*dz' + k2*y + k3 ; optionally also emits dz' (the gradient that flows into the residual branch).
template <bool RELU, bool TWO, bool DRES>
__global__ void __launch_bounds__(256)
    bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dza, const __nv_bfloat16* __restrict__ dzb,
                        const __nv_bfloat16* __restrict__ z, const unsigned char* __restrict__ mask,
                        const __nv_bfloat16* __restrict__ y, const float* __restrict__ coef,
                        __nv_bfloat16* __restrict__ dy,
                        __nv_bfloat16* __restrict__ dres, long long nvec, int C8) {
  const int C = C8 * 8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec;
       i += (long long)gridDim.x * blockDim.x) {
    const int cv = (int)(i % C8);
    Vec8 g = load_bf16x8(dza + i * 8);
    if (TWO) {
      Vec8 g2 = load_bf16x8(dzb + i * 8);
#pragma unroll
      for (int j = 0; j < 8; ++j) g.v[j] += g2.v[j];
    }
    if (RELU) {
      if (mask) {
        const unsigned bits = mask[i];
#pragma unroll
        for (int j = 0; j < 8; ++j) g.v[j] = ((bits >> j) & 1u) ? g.v[j] : 0.f;
      } else {
        Vec8 zz = load_bf16x8(z + i * 8);
#pragma unroll
        for (int j = 0; j < 8; ++j) g.v[j] = zz.v[j] > 0.f ? g.v[j] : 0.f;
      }
    }
    if (DRES) store_bf16x8(dres + i * 8, g);
    Vec8 yy = load_bf16x8(y + i * 8);
    Vec8 k1 = load_f32x8(coef + cv * 8), k2 = load_f32x8(coef + C + cv * 8), k3 = load_f32x8(coef + 2 * C + cv * 8);
    Vec8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.v[j] = fmaf(k1.v[j], g.v[j], fmaf(k2.v[j], yy.v[j], k3.v[j]));
    store_bf16x8(dy + i * 8, o);
  }
}

// 0 = plain loop, 2 = batched loads: 4 row groups in flight, 2 CTAs/SM (b2_bn_set_reduce_variant). Default 2: measured
// on B200 (tools/bn_bench.py, batch 256, reduce + apply of a residual block output), C=256 @56x56: 0.546 ms plain loop,
// 0.475 ms batched; the two other batchings that were tried (2 groups @ 4 CTAs/SM: 0.535, 3 groups @ 3 CTAs/SM: 0.497)
// lost on every ResNet-50 shape and are no longer built.
static int g_bn_reduce_variant = 2;
static int g_bn_apply_batched = 1;   // b2_bn_set_apply_batched(0): grid-stride apply kernels again

// dy = k1*g + k2*y + k3 with g = dz (optionally masked by the 1-bit ReLU mask): the apply pass in the shape of the
// batched reduce pass - every thread owns one 8-channel vector for the whole kernel (the three coefficient vectors
// live in registers instead of being re-read from L1 for every 16 bytes of data) and issues the loads of U row groups
// back to back. The grid-stride version above reaches 80 % of the copy bandwidth on the wide tensors.
template <bool RELU, int U, int BPS>
__global__ void __launch_bounds__(kBnThreads, BPS)
    bn_bwd_apply_batched_kernel(const __nv_bfloat16* __restrict__ dz, const unsigned char* __restrict__ mask,
                                const __nv_bfloat16* __restrict__ y, const float* __restrict__ coef,
                                __nv_bfloat16* __restrict__ dy, long long M, int C) {
  const int C8 = C >> 3, RP = kBnThreads / C8;
  const int cv = threadIdx.x % C8, r = threadIdx.x / C8;
  if (r >= RP) return;
  const Vec8 k1 = load_f32x8(coef + cv * 8), k2 = load_f32x8(coef + C + cv * 8), k3 = load_f32x8(coef + 2 * C + cv * 8);
  const long long step = (long long)gridDim.x * RP;
  for (long long row0 = (long long)blockIdx.x * RP + r; row0 < M; row0 += step * U) {
    uint4 ug[U], uy[U];
    unsigned bits[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long row = row0 + u * step;
      if (row < M) {
        const long long off = row * C + cv * 8;
        ug[u] = ld_nc_v4(dz + off);
        uy[u] = ld_nc_v4(y + off);
        if (RELU) bits[u] = mask[off >> 3];
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long row = row0 + u * step;
      if (row < M) {
        const uint32_t gw[4] = {ug[u].x, ug[u].y, ug[u].z, ug[u].w};
        const uint32_t yw[4] = {uy[u].x, uy[u].y, uy[u].z, uy[u].w};
        uint32_t ow[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float g0 = __uint_as_float(gw[q] << 16), g1 = __uint_as_float(gw[q] & 0xffff0000u);
          if (RELU) {
            g0 = ((bits[u] >> (2 * q)) & 1u) ? g0 : 0.f;
            g1 = ((bits[u] >> (2 * q + 1)) & 1u) ? g1 : 0.f;
          }
          const float y0 = __uint_as_float(yw[q] << 16), y1 = __uint_as_float(yw[q] & 0xffff0000u);
          ow[q] = pack_bf16x2(fmaf(k1.v[2 * q], g0, fmaf(k2.v[2 * q], y0, k3.v[2 * q])),
                              fmaf(k1.v[2 * q + 1], g1, fmaf(k2.v[2 * q + 1], y1, k3.v[2 * q + 1])));
        }
        *reinterpret_cast<uint4*>(dy + row * C + cv * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
      }
    }
  }
}

static int plan_blocks(long long M, int C, long long* rows_per_blk) {
  const int C8 = C / 8, RP = kBnThreads / C8;
  long long nblk = (M + (long long)RP * 8 - 1) / ((long long)RP * 8);
  long long cap = (long long)host_num_sms() * B2_BN_BLOCKS_PER_SM;
  if (cap > kMaxPartialBlocks) cap = kMaxPartialBlocks;
  if (nblk > cap) nblk = cap;
  if (nblk < 1) nblk = 1;
  long long rpb = (M + nblk - 1) / nblk;
  rpb = (rpb + RP - 1) / RP * RP;
  nblk = (M + rpb - 1) / rpb;
  *rows_per_blk = rpb;
  return (int)nblk;
}

static int stream_grid(long long nvec) {
  long long b = (nvec + 255) / 256;
  const long long cap = (long long)host_num_sms() * 16;
  return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
}

}  // namespace b2

using namespace b2;

B2_DEFINE_SPIN_TIMEOUT_SETTER(b2_set_spin_timeout_ns_bn)

extern "C" void b2_bn_set_apply_batched(int on) { g_bn_apply_batched = on; }
extern "C" void b2_bn_set_reduce_variant(int v) { g_bn_reduce_variant = (v == 0) ? 0 : 2; }
extern "C" long long b2_bn_partial_floats(int C) { return (long long)kMaxPartialBlocks * 2 * C; }

static bool bn_shape_ok(int C) { return C % 8 == 0 && C >= 8 && C <= 2048 && (kBnThreads % (C / 8) == 0); }

// given_blocks > 0: `partial` already holds that many [2][C] rows of (sum, sum of squares) — produced by the
// convolution's epilogue — and the statistics pass over y is skipped.
static int bn_fwd_train_impl(const void* y, const void* res, void* z, void* relu_mask, long long M, int C,
                             const float* gamma, const float* beta, float* running_mean, float* running_var,
                             long long* num_batches_tracked, float momentum, float eps, int relu, float* mean,
                             float* invstd, float* scale_shift /*2C*/, float* partial, int given_blocks,
                             const b2_xgpu_ctx* xg, int xg_slot, void* stream) {
  if (!bn_shape_ok(C)) return cudaErrorInvalidValue;
  cudaStream_t st = (cudaStream_t)stream;
  int nblk = given_blocks;
  if (given_blocks <= 0) {
    long long rpb;
    nblk = plan_blocks(M, C, &rpb);
    const int RP = kBnThreads / (C / 8);
    bn_stats_kernel<<<nblk, kBnThreads, 2 * RP * C * sizeof(float), st>>>((const __nv_bfloat16*)y, partial, M, C,
                                                                         rpb);
  }
  XgpuCtx ctx = xgpu_from_c(xg);
  bn_finalize_kernel<<<C / kFinCh, kFinThreads, 0, st>>>(
      partial, nblk, C, (float)M, gamma, beta, running_mean, running_var, num_batches_tracked, momentum, eps, mean,
      invstd, scale_shift, scale_shift + C, ctx, xg_slot);
  if (!z) return cudaGetLastError();   // coefficients only: the consumer applies them itself (b2_bn_relu_maxpool_fwd)
  const long long nvec = M * (C / 8);
  const int grid = stream_grid(nvec);
  const __nv_bfloat16* yy = (const __nv_bfloat16*)y;
  const __nv_bfloat16* rr = (const __nv_bfloat16*)res;
  __nv_bfloat16* zz = (__nv_bfloat16*)z;
  unsigned char* mk = (unsigned char*)relu_mask;
  if (g_bn_apply_batched && bn_shape_ok(C)) {
    const int RPa = kBnThreads / (C / 8);
    long long gb = (long long)host_num_sms() * 2, need = (M + RPa - 1) / RPa;
    if (gb > need) gb = need < 1 ? 1 : need;
    const float *scp = scale_shift, *shp = scale_shift + C;
    if (relu && res) bn_apply_batched_kernel<true, true, 4, 2><<<(int)gb, kBnThreads, 0, st>>>(yy, rr, scp, shp, zz, mk, M, C);
    else if (relu) bn_apply_batched_kernel<true, false, 4, 2><<<(int)gb, kBnThreads, 0, st>>>(yy, rr, scp, shp, zz, mk, M, C);
    else if (res) bn_apply_batched_kernel<false, true, 4, 2><<<(int)gb, kBnThreads, 0, st>>>(yy, rr, scp, shp, zz, mk, M, C);
    else bn_apply_batched_kernel<false, false, 4, 2><<<(int)gb, kBnThreads, 0, st>>>(yy, rr, scp, shp, zz, mk, M, C);
    return cudaGetLastError();
  }
  if (relu && res) bn_apply_kernel<true, true><<<grid, 256, 0, st>>>(yy, rr, scale_shift, scale_shift + C, zz, mk, nvec, C / 8);
  else if (relu) bn_apply_kernel<true, false><<<grid, 256, 0, st>>>(yy, rr, scale_shift, scale_shift + C, zz, mk, nvec, C / 8);
  else if (res) bn_apply_kernel<false, true><<<grid, 256, 0, st>>>(yy, rr, scale_shift, scale_shift + C, zz, mk, nvec, C / 8);
  else bn_apply_kernel<false, false><<<grid, 256, 0, st>>>(yy, rr, scale_shift, scale_shift + C, zz, mk, nvec, C / 8);
  return cudaGetLastError();
}

extern "C" int b2_bn_fwd_train(const void* y, const void* res, void* z, void* relu_mask, long long M, int C,
                               const float* gamma, const float* beta, float* running_mean, float* running_var,
                               long long* num_batches_tracked, float momentum, float eps, int relu, float* mean,
                               float* invstd, float* scale_shift, float* partial, const b2_xgpu_ctx* xg, int xg_slot,
                               void* stream) {
  return bn_fwd_train_impl(y, res, z, relu_mask, M, C, gamma, beta, running_mean, running_var, num_batches_tracked,
                           momentum, eps, relu, mean, invstd, scale_shift, partial, 0, xg, xg_slot, stream);
}
extern "C" int b2_bn_fwd_train_partial(const void* y, const void* res, void* z, void* relu_mask, long long M, int C,
                                       const float* gamma, const float* beta, float* running_mean,
                                       float* running_var, long long* num_batches_tracked, float momentum, float eps,
                                       int relu, float* mean, float* invstd, float* scale_shift, float* partial,
                                       int partial_blocks, const b2_xgpu_ctx* xg, int xg_slot, void* stream) {
  if (partial_blocks <= 0) return cudaErrorInvalidValue;
  return bn_fwd_train_impl(y, res, z, relu_mask, M, C, gamma, beta, running_mean, running_var, num_batches_tracked,
                           momentum, eps, relu, mean, invstd, scale_shift, partial, partial_blocks, xg, xg_slot,
                           stream);
}

extern "C" int b2_bn_fwd_eval(const void* y, const void* res, void* z, long long M, int C, const float* gamma,
                              const float* beta, const float* running_mean, const float* running_var, float eps,
                              int relu, float* scale_shift, void* stream) {
  if (C % 8) return cudaErrorInvalidValue;
  cudaStream_t st = (cudaStream_t)stream;
  bn_eval_coeffs_kernel<<<(C + 127) / 128, 128, 0, st>>>(C, gamma, beta, running_mean, running_var, eps, scale_shift,
                                                       scale_shift + C);
  if (!z) return cudaGetLastError();   // coefficients only
  const long long nvec = M * (C / 8);
  const int grid = stream_grid(nvec);
  const __nv_bfloat16* yy = (const __nv_bfloat16*)y;
  const __nv_bfloat16* rr = (const __nv_bfloat16*)res;
  __nv_bfloat16* zz = (__nv_bfloat16*)z;
  unsigned char* mk = nullptr;
  if (g_bn_apply_batched && bn_shape_ok(C)) {
    const int RPa = kBnThreads / (C / 8);
    long long gb = (long long)host_num_sms() * 2, need = (M + RPa - 1) / RPa;
    if (gb > need) gb = need < 1 ? 1 : need;
    const float *scp = scale_shift, *shp = scale_shift + C;
    if (relu && res) bn_apply_batched_kernel<true, true, 4, 2><<<(int)gb, kBnThreads, 0, st>>>(yy, rr, scp, shp, zz, mk, M, C);
    else if (relu) bn_apply_batched_kernel<true, false, 4, 2><<<(int)gb, kBnThreads, 0, st>>>(yy, rr, scp, shp, zz, mk, M, C);
    else if (res) bn_apply_batched_kernel<false, true, 4, 2><<<(int)gb, kBnThreads, 0, st>>>(yy, rr, scp, shp, zz, mk, M, C);
    else bn_apply_batched_kernel<false, false, 4, 2><<<(int)gb, kBnThreads, 0, st>>>(yy, rr, scp, shp, zz, mk, M, C);
    return cudaGetLastError();
  }
  if (relu && res) bn_apply_kernel<true, true><<<grid, 256, 0, st>>>(yy, rr, scale_shift, scale_shift + C, zz, mk, nvec, C / 8);
  else if (relu) bn_apply_kernel<true, false><<<grid, 256, 0, st>>>(yy, rr, scale_shift, scale_shift + C, zz, mk, nvec, C / 8);
  else if (res) bn_apply_kernel<false, true><<<grid, 256, 0, st>>>(yy, rr, scale_shift, scale_shift + C, zz, mk, nvec, C / 8);
  else bn_apply_kernel<false, false><<<grid, 256, 0, st>>>(yy, rr, scale_shift, scale_shift + C, zz, mk, nvec, C / 8);
  return cudaGetLastError();
}

#define B2_BWD_DISPATCH(KERNEL, ...)                                            \
  do {                                                                          \
    if (relu) {                                                                 \
      if (dzb) KERNEL(true, true, __VA_ARGS__); else KERNEL(true, false, __VA_ARGS__);   \
    } else {                                                                    \
      if (dzb) KERNEL(false, true, __VA_ARGS__); else KERNEL(false, false, __VA_ARGS__); \
    }                                                                           \
  } while (0)

static int bn_bwd_impl(const void* dz_a, const void* dz_b, const void* z, const void* relu_mask, const void* y,
                       void* dy, void* dres, long long M, int C, const float* gamma, const float* mean,
                       const float* invstd, float* dgamma, float* dbeta, int accumulate, int relu, float* coef,
                       float* partial, int given_blocks, const b2_xgpu_ctx* xg, int xg_slot, void* stream, int qH = 0,
                       int qW = 0);

extern "C" int b2_bn_bwd(const void* dz_a, const void* dz_b, const void* z, const void* relu_mask, const void* y,
                         void* dy, void* dres, long long M, int C, const float* gamma, const float* mean, const float* invstd,
                         float* dgamma, float* dbeta, int accumulate, int relu, float* coef /*3C*/, float* partial,
                         const b2_xgpu_ctx* xg, int xg_slot, void* stream) {
  return bn_bwd_impl(dz_a, dz_b, z, relu_mask, y, dy, dres, M, C, gamma, mean, invstd, dgamma, dbeta, accumulate, relu,
                     coef, partial, 0, xg, xg_slot, stream);
}

// Same with the second incoming gradient at half resolution: dz_b_q is [N, ceil(H/2), ceil(W/2), C] and stands for
// a full-resolution [N, H, W, C] tensor that is dz_b_q at the even pixels and zero elsewhere (the input gradient of a
// 1x1 / stride 2 convolution). Needs the 1-bit ReLU mask form and a dres output.
extern "C" int b2_bn_bwd_q(const void* dz_a, const void* dz_b_q, int H, int W, const void* relu_mask, const void* y,
                           void* dy, void* dres, long long M, int C, const float* gamma, const float* mean,
                           const float* invstd, float* dgamma, float* dbeta, int accumulate, int relu, float* coef,
                           float* partial, const b2_xgpu_ctx* xg, int xg_slot, void* stream) {
  if (!dz_b_q || H <= 0 || W <= 0 || M % ((long long)H * W)) return cudaErrorInvalidValue;
  return bn_bwd_impl(dz_a, dz_b_q, nullptr, relu_mask, y, dy, dres, M, C, gamma, mean, invstd, dgamma, dbeta,
                     accumulate, relu, coef, partial, 0, xg, xg_slot, stream, H, W);
}

// Same, but `partial` already holds partial_blocks rows of [2][C] (sum g, sum g*(y-mean)) - produced by the dgrad
// that computed dz (b2_conv_dgrad_bnstats) - so the reduce pass over dz / y is skipped. Single incoming gradient,
// no residual-branch output.
extern "C" int b2_bn_bwd_partial(const void* dz, const void* z, const void* relu_mask, const void* y, void* dy,
                                 long long M, int C, const float* gamma, const float* mean, const float* invstd,
                                 float* dgamma, float* dbeta, int accumulate, int relu, float* coef, float* partial,
                                 int partial_blocks, const b2_xgpu_ctx* xg, int xg_slot, void* stream) {
  if (partial_blocks <= 0 || partial_blocks > kMaxPartialBlocks) return cudaErrorInvalidValue;
  return bn_bwd_impl(dz, nullptr, z, relu_mask, y, dy, nullptr, M, C, gamma, mean, invstd, dgamma, dbeta, accumulate,
                     relu, coef, partial, partial_blocks, xg, xg_slot, stream);
}

static int bn_bwd_impl(const void* dz_a, const void* dz_b, const void* z, const void* relu_mask, const void* y,
                       void* dy, void* dres, long long M, int C, const float* gamma, const float* mean,
                       const float* invstd, float* dgamma, float* dbeta, int accumulate, int relu, float* coef,
                       float* partial, int given_blocks, const b2_xgpu_ctx* xg, int xg_slot, void* stream, int qH,
                       int qW) {
  if (!bn_shape_ok(C)) return cudaErrorInvalidValue;
  cudaStream_t st = (cudaStream_t)stream;
  long long rpb;
  int nblk = plan_blocks(M, C, &rpb);
  const int RP = kBnThreads / (C / 8);
  const size_t sm = 2 * RP * C * sizeof(float);
  const __nv_bfloat16 *a = (const __nv_bfloat16*)dz_a, *b = (const __nv_bfloat16*)dz_b,
                      *zz = (const __nv_bfloat16*)z, *yy = (const __nv_bfloat16*)y;
  const bool dzb = dz_b != nullptr;
  const unsigned char* mk = (const unsigned char*)relu_mask;
  __nv_bfloat16 *dyp = (__nv_bfloat16*)dy, *drp = (__nv_bfloat16*)dres;
#define RED(R, T, ...)                                                                                              \
  do {                                                                                                              \
    if (dres) bn_bwd_reduce_kernel<R, T, true><<<nblk, kBnThreads, sm, st>>>(a, b, zz, mk, yy, mean, partial, drp, M, C, rpb); \
    else bn_bwd_reduce_kernel<R, T, false><<<nblk, kBnThreads, sm, st>>>(a, b, zz, mk, yy, mean, partial, drp, M, C, rpb);     \
  } while (0)
#define REDB(R, T, UU, BPS)                                                                                         \
  do {                                                                                                              \
    if (dres) bn_bwd_reduce_batched_kernel<R, T, true, UU, BPS><<<nblk, kBnThreads, sm, st>>>(a, b, mk, yy, mean, partial, drp, M, C, qH, qW); \
    else bn_bwd_reduce_batched_kernel<R, T, false, UU, BPS><<<nblk, kBnThreads, sm, st>>>(a, b, mk, yy, mean, partial, drp, M, C, qH, qW);     \
  } while (0)
#define REDV(UU, BPS)                                        \
  do {                                                       \
    if (relu && dzb && qH) REDB(true, 2, UU, BPS);           \
    else if (relu && dzb) REDB(true, 1, UU, BPS);            \
    else if (relu) REDB(true, 0, UU, BPS);                   \
    else if (dzb && qH) REDB(false, 2, UU, BPS);             \
    else if (dzb) REDB(false, 1, UU, BPS);                   \
    else REDB(false, 0, UU, BPS);                            \
  } while (0)
  // batched variants need the 1-bit ReLU mask (or no ReLU); the z-tensor form keeps the plain loop
  if (qH && ((relu && !mk) || !dres)) return cudaErrorInvalidValue;   // compact second input: mask form + dres only
  const int variant = (relu && !mk) ? 0 : ((qH && g_bn_reduce_variant == 0) ? 2 : g_bn_reduce_variant);
  if (given_blocks > 0) {
    nblk = given_blocks;
  } else if (variant == 0) {
    B2_BWD_DISPATCH(RED, 0);
  } else {
    long long cap = (long long)host_num_sms() * 2;   // two CTAs per SM
    if (nblk > cap) nblk = (int)cap;
    REDV(4, 2);
  }
#undef REDV
#undef REDB
#undef RED
  XgpuCtx ctx = xgpu_from_c(xg);
  bn_bwd_finalize_kernel<<<C / kFinCh, kFinThreads, 0, st>>>(partial, nblk, C, (float)M, gamma, mean, invstd, dgamma,
                                                          dbeta, accumulate, coef, ctx, xg_slot);
  const long long nvec = M * (C / 8);
  const int grid = stream_grid(nvec);
  // batched apply (one incoming gradient, mask form or no ReLU): same grid rule as the batched reduce pass
  const bool batched_apply = g_bn_apply_batched && (dres || (!dzb && (!relu || mk)));
  if (batched_apply) {
    long long gb = (long long)host_num_sms() * 2;
    long long need = (M + RP - 1) / RP;
    if (gb > need) gb = need < 1 ? 1 : need;
    if (dres || !relu)
      bn_bwd_apply_batched_kernel<false, 4, 2><<<(int)gb, kBnThreads, 0, st>>>(dres ? drp : a, nullptr, yy, coef, dyp, M, C);
    else
      bn_bwd_apply_batched_kernel<true, 4, 2><<<(int)gb, kBnThreads, 0, st>>>(a, mk, yy, coef, dyp, M, C);
    return cudaGetLastError();
  }
  if (dres) {
    // the reduce pass has already produced dz' (= dres): dy = k1*dres + k2*y + k3
    bn_bwd_apply_kernel<false, false, false><<<grid, 256, 0, st>>>(drp, nullptr, nullptr, nullptr, yy, coef, dyp,
                                                                   nullptr, nvec, C / 8);
    return cudaGetLastError();
  }
#define APP(R, T, ...) bn_bwd_apply_kernel<R, T, false><<<grid, 256, 0, st>>>(a, b, zz, mk, yy, coef, dyp, nullptr, nvec, C / 8)
  B2_BWD_DISPATCH(APP, 0);
#undef APP
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ no-allocation variant
// Training-mode BatchNorm forward for callers that must not allocate device memory: a latency-critical loop where
// cudaMalloc's implicit device synchronisation is not acceptable, a host application that hands this library a
// fixed memory budget and forbids side allocations, or a GPU so full that an allocation could fail mid-step.
// The per-CTA partial sums (the only scratch the BN forward needs) live in a file-scope device array instead of a
// caller-provided buffer.  The array is sized for the widest layer the library supports and exists once per GPU
// that loads the library, for the lifetime of the process.
constexpr int kNoAllocMaxChannels = B2_BN_NOALLOC_MAX_CHANNELS;
static __device__ __align__(256) float g_bn_noalloc_scratch[(size_t)b2::kMaxPartialBlocks * 2 * kNoAllocMaxChannels];

// Device address of the scratch array for the current device.  In host code the array's name is only a handle for
// the runtime, not something a kernel can dereference, so the address is looked up once per device (with that
// device current) and cached.
static cudaError_t bn_noalloc_scratch(float** out) {
  static std::mutex mu;
  static float* ptrs[64] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> g(mu);
  if (!ptrs[dev]) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_bn_noalloc_scratch);
    if (e != cudaSuccess) return e;
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return cudaSuccess;
}

extern "C" int b2_bn_fwd_train_noalloc(const void* y, const void* res, void* z, long long M, int C,
                                       const float* gamma, const float* beta, float* running_mean,
                                       float* running_var, long long* num_batches_tracked, float momentum, float eps,
                                       int relu, float* mean, float* invstd, float* scale_shift,
                                       const b2_xgpu_ctx* xg, int xg_slot, void* stream) {
  if (C > kNoAllocMaxChannels) {
    std::fprintf(stderr,
                 "b2_bn_fwd_train_noalloc: C=%d exceeds the built-in scratch (B2_BN_NOALLOC_MAX_CHANNELS=%d); use "
                 "b2_bn_fwd_train with a caller-provided partial buffer\n",
                 C, kNoAllocMaxChannels);
    return cudaErrorInvalidValue;
  }
  if (!bn_shape_ok(C) || M <= 0) {
    std::fprintf(stderr, "b2_bn_fwd_train_noalloc: unsupported shape M=%lld C=%d (C must divide 2048 and be >= 8)\n",
                 M, C);
    return cudaErrorInvalidValue;
  }
  float* scratch = nullptr;
  cudaError_t e = bn_noalloc_scratch(&scratch);
  if (e != cudaSuccess) return e;
  // Nothing in the scratch has to start at zero: the statistics kernel overwrites every partial the finalize
  // kernel reads, so whatever the previous call left behind is never observed.
  return b2_bn_fwd_train(y, res, z, /*relu_mask=*/nullptr, M, C, gamma, beta, running_mean, running_var, num_batches_tracked, momentum, eps,
                         relu, mean, invstd, scale_shift, scratch, xg, xg_slot, stream);
}
