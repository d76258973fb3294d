// Classifier head, fused: global average pool -> fully connected layer (tcgen05, fp32 logits) in ONE kernel, and the
// cross-entropy kernels that consume / produce its tensors without any cast in between.
//
// Forward (`head_fwd_kernel`): thread-block clusters of 8 CTAs. A cluster owns 16 images; CTA r of the cluster
//   1. pools 1/8 of the channels of those 16 images (coalesced 16-byte loads of the NHWC feature map, fp32 sums),
//      rounds the means to bf16 and scatters them through DISTRIBUTED SHARED MEMORY into the B-operand tile
//      (K-major, 128-byte swizzle) of all 8 CTAs - the pooled matrix never makes a round trip through HBM before
//      the GEMM (a copy is written out once for the backward pass);
//   2. streams its own 128 classes of the weight matrix through a TMA / mbarrier ring and runs
//      D[128 classes x 16 images] += W[128 x K] * pooled[16 x K]^T on the tensor cores (operands swapped so that
//      the 16-image side is the MMA's N), accumulator in TMEM;
//   3. reads the accumulator back (tcgen05.ld), adds the bias and stores fp32 logits, coalesced along the classes.
//   The weight TMA ring is primed before the pooling starts, so the weight traffic hides behind the pooling.
// Loss forward (`ce_fwd_f32_kernel`): log-sum-exp + NLL per row on the fp32 logits, mean over the batch by the last
//   CTA to finish (fixed summation order: deterministic), one launch.
// Loss backward (`ce_bwd_fused_kernel`): bf16 dlogits = (softmax - onehot) * g / rows AND the bias gradient
//   (column sums, fixed order) in one launch.
//
// Replaces aten::mean + cuBLAS addmm + _log_softmax + nll_loss (+ their backward) that the reference reaches through
// torchvision's avgpool/fc and nn.CrossEntropyLoss (reference resnet/main.py:76,102,121-123).
#include "b200ddp.h"
#include "ptx.cuh"
#include "tmap.h"

#include <cstdio>

namespace b2 {

constexpr int kHeadCluster = 8;        // CTAs per cluster = class tiles of 128 (covers up to 1024 classes)
constexpr int kHeadImgs = 16;          // images per cluster = MMA N
constexpr int kHeadThreads = 512;      // 16 warps: one image each while pooling; then warp 0 = TMA, 1 = MMA, 4-7 = epilogue
constexpr int kHeadStages = 6;         // weight ring: 6 x 16 KB
constexpr int kHeadStageBytes = 128 * 128;   // 128 classes x 64 bf16
constexpr int kHeadMaxC = 2048;

B2_DEVICE void st_shared_cluster_v4(uint32_t cluster_addr, uint4 v) {
  asm volatile("st.shared::cluster.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(cluster_addr), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}
B2_DEVICE void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
// 32 lanes x 16 columns of fp32: thread i of the warp gets TMEM lane i, columns [col, col+16).
B2_DEVICE void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}

struct HeadSmem {
  // B operand: for every 64-channel block a [16 images x 128 B] K-major tile with the 128-byte swizzle
  alignas(1024) uint8_t b[kHeadMaxC / 64 * kHeadImgs * 128];       // 64 KB
  alignas(1024) uint8_t a[kHeadStages][kHeadStageBytes];           // 96 KB
  uint64_t full[kHeadStages], empty[kHeadStages], accum;
  uint32_t tmem;
};

// x: [N][HW][C] bf16 (NHWC feature map); w via tmW: [ncls][C] bf16; pooled out: [N][C] bf16; logits: [N][ld] fp32.
// The weight tiles come in by plain (non-multicast) TMA loads of this GPU's own memory, one per CTA.
__global__ void __cluster_dims__(kHeadCluster, 1, 1) __launch_bounds__(kHeadThreads, 1)
    head_fwd_kernel(const __grid_constant__ CUtensorMap tmW, const __nv_bfloat16* __restrict__ x,
                    const float* __restrict__ bias, __nv_bfloat16* __restrict__ pooled, float* __restrict__ logits,
                    int N, int HW, int C, int ncls, int ld) {
  extern __shared__ uint8_t smem_raw[];
  HeadSmem& s = *reinterpret_cast<HeadSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t crank = cluster_ctarank();
  const int img0 = (blockIdx.x / kHeadCluster) * kHeadImgs;
  const int cls0 = (int)crank * 128;
  const int kblocks = C / 64;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmW);
    for (int i = 0; i < kHeadStages; ++i) {
      mbar_init(&s.full[i], 1);
      mbar_init(&s.empty[i], 1);
    }
    mbar_init(&s.accum, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<32>(&s.tmem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  cluster_sync_all();   // every CTA of the cluster is resident: its shared memory may be written remotely from here on

  // ---- prime the weight ring (overlaps the pooling below)
  const int primed = kblocks < kHeadStages ? kblocks : kHeadStages;
  if (warp == 0 && elect_one()) {
    for (int i = 0; i < primed; ++i) {
      mbar_arrive_expect_tx(&s.full[i], kHeadStageBytes);
      tma_load_2d(s.a[i], &tmW, &s.full[i], i * 64, cls0);
    }
  }
  __syncwarp();

  // ---- pooling: this CTA owns channels [crank * C/8, (crank+1) * C/8) of the cluster's 16 images
  {
    const int cs = C / kHeadCluster;            // channels of this CTA's slice (multiple of 8)
    const int L = cs / 8;                       // lanes per pixel (power of two, <= 32)
    const int P = 32 / L;                       // pixels per warp iteration
    const int c_lane = (int)crank * cs + (lane % L) * 8;
    const int psub = lane / L;
    const float inv = 1.f / (float)HW;
    const uint32_t b_base = smem_u32(s.b);
    for (int im = warp; im < kHeadImgs; im += kHeadThreads / 32) {
      const int img = img0 + im;
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = 0.f;
      if (img < N) {
        const __nv_bfloat16* src = x + ((long long)img * HW) * C + c_lane;
        int p = psub;
        for (; p + 13 * P < HW; p += 14 * P) {  // 14 independent 16-byte loads in flight per lane
          uint4 u[14];
#pragma unroll
          for (int q = 0; q < 14; ++q) u[q] = ld_nc_v4(src + (long long)(p + q * P) * C);
#pragma unroll
          for (int q = 0; q < 14; ++q) {
            const float2 a0 = unpack_bf16x2(u[q].x), a1 = unpack_bf16x2(u[q].y), a2 = unpack_bf16x2(u[q].z),
                         a3 = unpack_bf16x2(u[q].w);
            acc[0] += a0.x; acc[1] += a0.y; acc[2] += a1.x; acc[3] += a1.y;
            acc[4] += a2.x; acc[5] += a2.y; acc[6] += a3.x; acc[7] += a3.y;
          }
        }
        for (; p + 6 * P < HW; p += 7 * P) {    // 7 independent 16-byte loads in flight per lane
          uint4 u[7];
#pragma unroll
          for (int q = 0; q < 7; ++q) u[q] = ld_nc_v4(src + (long long)(p + q * P) * C);
#pragma unroll
          for (int q = 0; q < 7; ++q) {
            const float2 a0 = unpack_bf16x2(u[q].x), a1 = unpack_bf16x2(u[q].y), a2 = unpack_bf16x2(u[q].z),
                         a3 = unpack_bf16x2(u[q].w);
            acc[0] += a0.x; acc[1] += a0.y; acc[2] += a1.x; acc[3] += a1.y;
            acc[4] += a2.x; acc[5] += a2.y; acc[6] += a3.x; acc[7] += a3.y;
          }
        }
        for (; p < HW; p += P) {
          const uint4 u = ld_nc_v4(src + (long long)p * C);
          const float2 a0 = unpack_bf16x2(u.x), a1 = unpack_bf16x2(u.y), a2 = unpack_bf16x2(u.z),
                       a3 = unpack_bf16x2(u.w);
          acc[0] += a0.x; acc[1] += a0.y; acc[2] += a1.x; acc[3] += a1.y;
          acc[4] += a2.x; acc[5] += a2.y; acc[6] += a3.x; acc[7] += a3.y;
        }
      }
      // fold the P pixel groups of the warp together (fixed order: deterministic)
      for (int o = L; o < 32; o <<= 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
      }
      if (lane < L) {
        uint4 v;
        v.x = pack_bf16x2(acc[0] * inv, acc[1] * inv);
        v.y = pack_bf16x2(acc[2] * inv, acc[3] * inv);
        v.z = pack_bf16x2(acc[4] * inv, acc[5] * inv);
        v.w = pack_bf16x2(acc[6] * inv, acc[7] * inv);
        if (img < N) *reinterpret_cast<uint4*>(pooled + (long long)img * C + c_lane) = v;
        // position of (image im, channels c_lane..c_lane+7) in the swizzled K-major tile
        const int kb = c_lane >> 6, chunk = (c_lane & 63) >> 3;
        const uint32_t off = (uint32_t)kb * (kHeadImgs * 128) + (uint32_t)(im >> 3) * 1024 + (uint32_t)(im & 7) * 128 +
                             (uint32_t)((chunk ^ (im & 7)) << 4);
#pragma unroll
        for (uint32_t r = 0; r < (uint32_t)kHeadCluster; ++r) st_shared_cluster_v4(mapa_cluster(b_base + off, r), v);
      }
    }
  }
  fence_proxy_async_all();   // generic-proxy writes (local and remote) before the tensor core's async-proxy reads
  cluster_sync_all();        // every CTA's slice has landed in every CTA's B tile

  if (warp == 0) {
    // ---- TMA producer: the rest of the weight k-blocks, as ring slots are released by the MMA warp
    if (elect_one()) {
      for (int i = primed; i < kblocks; ++i) {
        const int st = i % kHeadStages;
        mbar_wait(&s.empty[st], ((i / kHeadStages) - 1) & 1);
        mbar_arrive_expect_tx(&s.full[st], kHeadStageBytes);
        tma_load_2d(s.a[st], &tmW, &s.full[st], i * 64, cls0);
      }
    }
  } else if (warp == 1) {
    // ---- MMA issuer
    fence_proxy_async_smem();
    constexpr uint32_t idesc = make_idesc_bf16(128, kHeadImgs, false, false);
    const uint32_t d_tmem = s.tmem;
    if (elect_one()) {
      for (int i = 0; i < kblocks; ++i) {
        const int st = i % kHeadStages;
        mbar_wait(&s.full[st], (i / kHeadStages) & 1);
        tc_fence_after();
        const uint32_t aBase = smem_u32(s.a[st]);
        const uint32_t bBase = smem_u32(s.b) + (uint32_t)i * (kHeadImgs * 128);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t da = make_smem_desc_sw128(aBase + k * 32, 16, 1024);
          const uint64_t db = make_smem_desc_sw128(bBase + k * 32, 16, 1024);
          umma_bf16(d_tmem, da, db, idesc, (i | k) != 0);
        }
        umma_commit(&s.empty[st]);
      }
      umma_commit(&s.accum);
    }
  } else if (warp >= 4 && warp < 8) {
    // ---- epilogue: TMEM lane = class, column = image (warps 4..7: TMEM lane quadrant = warp % 4)
    mbar_wait(&s.accum, 0);
    tc_fence_after();
    const int q = warp - 4;                       // TMEM lane quarter this warp may read
    uint32_t v[16];
    tmem_ld16(s.tmem + ((uint32_t)(q * 32) << 16), v);
    tmem_ld_wait();
    const int cls = cls0 + q * 32 + lane;
    if (cls < ncls) {
      const float bv = bias ? bias[cls] : 0.f;
#pragma unroll
      for (int j = 0; j < kHeadImgs; ++j)
        if (img0 + j < N) logits[(long long)(img0 + j) * ld + cls] = __uint_as_float(v[j]) + bv;
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) tmem_free<32>(s.tmem);
}

}  // namespace b2

namespace b2 {
// ------------------------------------------------------------------------------------------------ cross entropy
// One CTA per row on fp32 logits; the last CTA to finish adds the per-row losses up in index order and writes the
// mean (counter resets itself: safe for CUDA-graph replays).
__global__ void __launch_bounds__(128) ce_fwd_f32_kernel(const float* __restrict__ logits,
                                                         const long long* __restrict__ labels, float* lse,
                                                         float* loss_row, float* loss, unsigned int* counter,
                                                         int rows, int ncls, int ld) {
  __shared__ float red[4];
  __shared__ bool last;
  const int row = blockIdx.x;
  const float* lr = logits + (long long)row * ld;
  float mx = -INFINITY;
  for (int j = threadIdx.x; j < ncls; j += blockDim.x) mx = fmaxf(mx, lr[j]);
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  __syncthreads();
  float sum = 0.f;
  for (int j = threadIdx.x; j < ncls; j += blockDim.x) sum += __expf(lr[j] - mx);
  sum = warp_sum(sum);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    const float l = mx + logf(red[0] + red[1] + red[2] + red[3]);
    lse[row] = l;
    const long long lab = labels[row];
    loss_row[row] = (lab >= 0 && lab < ncls) ? l - lr[lab] : 0.f;
    __threadfence();
    last = atomicAdd(counter, 1u) == (unsigned)rows - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  float t = 0.f;
  for (int i = threadIdx.x; i < rows; i += blockDim.x) t += __ldcg(loss_row + i);
  t = warp_sum(t);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    *loss = (red[0] + red[1] + red[2] + red[3]) / (float)rows;
    *counter = 0;
  }
}

// dlogits (bf16) = (softmax - onehot) * g / rows and dbias[j] (+)= sum_rows dlogits (of the ROUNDED values, so that
// the bias gradient equals the column sum of the matrix the weight-gradient GEMM sees). One CTA = 32 columns x all
// rows: 8 warps stride the rows, partial sums are folded in warp order.
__global__ void __launch_bounds__(256) ce_bwd_fused_kernel(const float* __restrict__ logits,
                                                           const long long* __restrict__ labels,
                                                           const float* __restrict__ lse,
                                                           const float* __restrict__ gscale,
                                                           __nv_bfloat16* __restrict__ dlogits, float* dbias,
                                                           int accumulate_bias, int rows, int ncls, int ld, int ld_out) {
  __shared__ float part[8][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int col = blockIdx.x * 32 + lane;
  const float g = (gscale ? *gscale : 1.f) / (float)rows;
  float colsum = 0.f;
  constexpr int U = 8;
  for (int r0 = warp; r0 < rows; r0 += 8 * U) {
    float lv[U], ls[U];
    long long lb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int r = r0 + u * 8;
      const bool ok = r < rows && col < ncls;
      lv[u] = ok ? logits[(long long)r * ld + col] : 0.f;
      ls[u] = r < rows ? lse[r] : 0.f;
      lb[u] = r < rows ? labels[r] : -1;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int r = r0 + u * 8;
      if (r < rows && col < ncls) {
        const float p = __expf(lv[u] - ls[u]);
        const __nv_bfloat16 d = __float2bfloat16((p - (lb[u] == col ? 1.f : 0.f)) * g);
        dlogits[(long long)r * ld_out + col] = d;
        colsum += __bfloat162float(d);
      }
    }
  }
  part[warp][lane] = colsum;
  __syncthreads();
  if (warp == 0 && col < ncls && dbias) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += part[w][lane];
    dbias[col] = accumulate_bias ? dbias[col] + t : t;
  }
}

}  // namespace b2

using namespace b2;

extern "C" int b2_head_supported(int C, int ncls) {
  if (C < 64 || C > kHeadMaxC || (C % 64)) return 0;
  const int L = C / 64;                  // lanes per pixel of a CTA's channel slice (C/8 channels, 8 per lane)
  if (L & (L - 1)) return 0;
  return ncls >= 1 && ncls <= 128 * kHeadCluster;
}

extern "C" int b2_head_fwd(const void* x, const void* w_bf16, const float* bias, void* pooled, float* logits, int N,
                           int HW, int C, int ncls, int ld, void* stream) {
  if (!b2_head_supported(C, ncls) || N <= 0 || HW <= 0) return cudaErrorInvalidValue;
  CUtensorMap tmW;
  int e = make_tmap_2d_bf16(&tmW, w_bf16, (uint64_t)C, (uint64_t)ncls, 64, 128);
  if (e) return e;
  const size_t smem = sizeof(HeadSmem) + 1024;
  static bool configured = false;
  if (!configured) {
    e = cudaFuncSetAttribute(head_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e) return e;
    configured = true;
  }
  const int clusters = (N + kHeadImgs - 1) / kHeadImgs;
  head_fwd_kernel<<<clusters * kHeadCluster, kHeadThreads, smem, (cudaStream_t)stream>>>(
      tmW, (const __nv_bfloat16*)x, bias, (__nv_bfloat16*)pooled, logits, N, HW, C, ncls, ld);
  return cudaGetLastError();
}

extern "C" int b2_ce_fwd_f32(const float* logits, const long long* labels, float* lse, float* loss_row, float* loss,
                             unsigned int* counter, int rows, int ncls, int ld, void* stream) {
  if (rows <= 0 || !counter) return cudaErrorInvalidValue;
  ce_fwd_f32_kernel<<<rows, 128, 0, (cudaStream_t)stream>>>(logits, labels, lse, loss_row, loss, counter, rows, ncls,
                                                            ld);
  return cudaGetLastError();
}

extern "C" int b2_ce_bwd_fused(const float* logits, const long long* labels, const float* lse, const float* gscale,
                               void* dlogits_bf16, float* dbias, int accumulate_bias, int rows, int ncls, int ld,
                               int ld_out, void* stream) {
  if (rows <= 0 || ncls <= 0) return cudaErrorInvalidValue;
  ce_bwd_fused_kernel<<<(ncls + 31) / 32, 256, 0, (cudaStream_t)stream>>>(
      logits, labels, lse, gscale, (__nv_bfloat16*)dlogits_bf16, dbias, accumulate_bias, rows, ncls, ld, ld_out);
  return cudaGetLastError();
}
