// ResNet stem (7x7 / stride 2 / pad 3, 3 input channels, 64 output channels) on the tensor cores without ever
// materialising an im2col matrix.
//
// The input batch is repacked once per step into a zero-padded NHWC4 bf16 image (3 rows top/bottom, 4 pixels
// left, >= 4 right, channel 3 = 0; 8 bytes per pixel).  For an output pixel (oh, ow) and a filter row kh, the
// 8 pixels x 4 channels starting at padded pixel 2*ow of padded row 2*oh+kh are 64 contiguous, 16-byte aligned
// bytes: filter columns kw' = 0..7 with kw' = kw + 1 (kw' = 0 and channel 3 carry zero weights).  A tiled TMA
// descriptor whose "row" dimension is the output column ow with a 16-byte stride therefore reads the GEMM
// A-operand rows straight out of the image (overlapping windows): one box of [<=128 ow x 64 B] per filter row, 7
// boxes per output row, 64-byte swizzle.  K = 7 * 32 = 224.
//
//   fprop : y[(n,oh), ow, 64]  = A[(n,oh,ow), 224] * Wp[64, 224]^T     M = 128 (one output row), N = 64, K-major;
//           A is not even replicated in shared memory: one TMA box brings the 7 raw padded rows and the operand
//           descriptor itself (no swizzle, 16-byte row step) expresses the overlapping windows
//   wgrad : dWp^T[224, 64]     = A[pixels, 224]^T  * dy[pixels, 64]    M = 2 x 128 (k), N = 64, both MN-major,
//           accumulated in TMEM over all the output rows a CTA owns, one partial per CTA, fixed-order reduce.
//
// fprop also leaves the BatchNorm statistics partials of y behind (same scheme as csrc/conv_igemm.cu).
// Replaces the 7x7 convolution torchvision's ResNet reaches through cuDNN (reference resnet/main.py:76).
#include "b200ddp.h"
#include "ptx.cuh"
#include "tmap.h"

namespace b2 {

constexpr int kStemKH = 7;
constexpr int kStemK = 224;          // 7 filter rows x (8 columns x 4 channels)
constexpr int kStemCout = 64;
constexpr int kStemRowsMax = 112;    // output columns per tile (one output row); wider images use the im2col path
constexpr int kStemStages = 3;

// ---------------------------------------------------------------------------------------------- packing
// NCHW (fp32 or bf16) -> padded NHWC4 bf16 image, padding included (every element of xp is written).
// x_slot (optional): device memory holding the address of the batch; the kernel then reads the batch through it, so
// a captured CUDA graph can be pointed at a different input buffer before every replay (no copy into a static
// buffer; utils/graph.py).
template <typename T>
__global__ void stem_pack_input_kernel(const T* __restrict__ x, const void* const* __restrict__ x_slot,
                                       uint2* __restrict__ xp, int N, int H, int W, int Hp, int Wp) {
  if (x_slot) x = reinterpret_cast<const T*>(*x_slot);
  const long long total = (long long)N * Hp * Wp;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int wp = (int)(i % Wp);
    const long long r = i / Wp;
    const int hp = (int)(r % Hp);
    const int n = (int)(r / Hp);
    const int h = hp - 3, w = wp - 4;
    uint2 o = make_uint2(0u, 0u);
    if (h >= 0 && h < H && w >= 0 && w < W) {
      const T* src = x + ((long long)n * 3 * H + h) * W + w;
      const float c0 = (float)src[0], c1 = (float)src[(long long)H * W], c2 = (float)src[2LL * H * W];
      o.x = pack_bf16x2(c0, c1);
      o.y = pack_bf16x2(c2, 0.f);
    }
    xp[i] = o;
  }
}

// fp32 OHWI [64][7][7][3] -> bf16 [64][7][8][4] (kw' = kw + 1, zero elsewhere)
__global__ void stem_pack_weight_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wp) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= kStemCout * kStemK) return;
  const int c = i & 3, kwp = (i >> 2) & 7, kh = (i >> 5) % 7, co = i / kStemK;
  float v = 0.f;
  if (c < 3 && kwp >= 1) v = w[((co * 7 + kh) * 7 + (kwp - 1)) * 3 + c];
  wp[i] = __float2bfloat16(v);
}

// partial[cta][k 256][64] (k = kh*32 + kw'*4 + c) -> dW fp32 OHWI [64][7][7][3], summed over CTAs in index order
__global__ void stem_wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ dw, int parts,
                                         int accumulate) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= kStemCout * 147) return;
  const int c = i % 3, kw = (i / 3) % 7, kh = (i / 21) % 7, co = i / 147;
  const int k = kh * 32 + (kw + 1) * 4 + c;
  float s = 0.f;
  for (int p = 0; p < parts; ++p) s += ws[((size_t)p * 256 + k) * kStemCout + co];
  dw[i] = accumulate ? dw[i] + s : s;
}

// ---------------------------------------------------------------------------------------------- fprop
// warp 0: TMA producer, warp 1: MMA issuer + TMEM owner, warps 2-5: epilogue, warps 6-9: BatchNorm statistics
constexpr int kStemFThreads = 320;
// A stage = the 7 raw padded image rows an output row needs (<= 232 pixels x 8 B each), nothing replicated: the
// A-operand descriptor (K-major, no swizzle) walks them as a Toeplitz matrix - core matrix = 8 output columns x 16 B
// with rows 16 B apart (the stride-2 window step), next K chunk 16 B further (LBO), next 8 rows 128 B further (SBO).
constexpr int kStemFStages = 8;
constexpr int kStemAStage = 13312;                         // >= 6 * 1856 + 16 * 127 + 64 (last window of row 127)
constexpr int kStemBBytes = kStemKH * kStemCout * 64;      // 7 boxes of [64 rows x 64 B]
constexpr int kStemPanel = kStemRowsMax * 128;             // output panel: [112 rows x 128 B]
constexpr int kStemFSmem = kStemFStages * kStemAStage + kStemBBytes + 2 * kStemPanel + 1024 + 256;

__global__ void __launch_bounds__(kStemFThreads, 1)
    stem_fprop_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmC, float* __restrict__ stats, int Ho, int Wo, int pitch,
                      int num_items) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sB = smem + kStemFStages * kStemAStage;
  uint8_t* epi = sB + kStemBBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi + 2 * kStemPanel);
  uint64_t* full = bars;                       // [stages]
  uint64_t* empty = bars + kStemFStages;        // [stages]
  uint64_t* tfull = bars + 2 * kStemFStages;    // [2]
  uint64_t* tempty = tfull + 2;                // [2]
  uint64_t* bfull = tempty + 2;                // weights landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bfull + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStemFStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 4);
    }
    mbar_init(bfull, 1);
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
  }
  if (warp == 1) tmem_alloc<128>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(bfull, kStemBBytes);
      for (int kh = 0; kh < kStemKH; ++kh) tma_load_2d(sB + kh * (kStemCout * 64), &tmB, bfull, kh * 32, 0);
      uint32_t g = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++g) {
        const int n = item / Ho, oh = item - n * Ho;
        const int s = g % kStemFStages;
        mbar_wait(&empty[s], ((g / kStemFStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&full[s], (uint32_t)(kStemKH * pitch));
        tma_load_3d(smem + s * kStemAStage, &tmA, &full[s], 0, 2 * oh, n);   // 7 consecutive padded rows
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, kStemCout, false, false);
      mbar_wait(bfull, 0);
      uint32_t g = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++g) {
        const uint32_t acc = g & 1;
        mbar_wait(&tempty[acc], ((g >> 1) & 1) ^ 1);
        const int s = g % kStemFStages;
        mbar_wait(&full[s], (g / kStemFStages) & 1);
        tc_fence_after();
        const uint32_t aBase = smem_u32(smem + s * kStemAStage), bBase = smem_u32(sB);
#pragma unroll
        for (int kh = 0; kh < kStemKH; ++kh) {
#pragma unroll
          for (int ks = 0; ks < 2; ++ks) {
            const uint64_t da = make_smem_desc(aBase + kh * pitch + ks * 32, 16, 128, 0);
            const uint64_t db = make_smem_desc(bBase + kh * (kStemCout * 64) + ks * 32, 16, 512, 4);
            umma_bf16(tmem_base + acc * kStemCout, da, db, idesc, (kh | ks) != 0);
          }
        }
        umma_commit(&empty[s]);
        umma_commit(&tfull[acc]);
      }
    }
  } else if (warp >= 6) {
    // ------------------------------------------------ statistics warps (see csrc/conv_igemm.cu for the scheme)
    if (stats) {
      const int sw = warp - 6, g = lane & 7;
      const int rg = sw >> 1;                        // row group: panel rows [56*rg, 56*rg + 56)
      const int chunk = (sw & 1) * 4 + (lane >> 3);  // 16-byte chunk = 8 channels
      {
        float4* z = reinterpret_cast<float4*>(stats + (size_t)blockIdx.x * 2 * 2 * kStemCout);
        for (int i = threadIdx.x - 192; i < 2 * 2 * kStemCout / 4; i += 128) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      asm volatile("bar.arrive 6, 256;" ::: "memory");   // both panel buffers start out free
      asm volatile("bar.arrive 7, 256;" ::: "memory");
      float a[8], b[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) a[e] = b[e] = 0.f;
      uint32_t t = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++t) {
        const uint32_t bsel = t & 1;
        asm volatile("bar.sync %0, 256;" ::"r"(2 + bsel) : "memory");
        const uint32_t src = smem_u32(epi) + bsel * kStemPanel + (rg * 56 + g) * 128 + ((chunk ^ g) << 4);
        uint4 w[7];
#pragma unroll
        for (int i = 0; i < 7; ++i) w[i] = ld_shared_v4(src + i * 1024);
        __threadfence_block();
        asm volatile("bar.arrive %0, 256;" ::"r"(6 + bsel) : "memory");
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          const uint32_t ww[4] = {w[i].x, w[i].y, w[i].z, w[i].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float lo = __uint_as_float(ww[e] << 16), hi = __uint_as_float(ww[e] & 0xffff0000u);
            a[2 * e] += lo;
            b[2 * e] = fmaf(lo, lo, b[2 * e]);
            a[2 * e + 1] += hi;
            b[2 * e + 1] = fmaf(hi, hi, b[2 * e + 1]);
          }
        }
      }
#pragma unroll
      for (int w = 4; w >= 1; w >>= 1) {
        const bool up = (g & w) != 0;
#pragma unroll
        for (int j = 0; j < w; ++j) {
          const float sa = up ? a[j] : a[j + w], ka = up ? a[j + w] : a[j];
          const float sb = up ? b[j] : b[j + w], kb = up ? b[j + w] : b[j];
          a[j] = ka + __shfl_xor_sync(0xffffffffu, sa, w);
          b[j] = kb + __shfl_xor_sync(0xffffffffu, sb, w);
        }
      }
      float* dst = stats + (size_t)(blockIdx.x * 2 + rg) * 2 * kStemCout + chunk * 8 + g;
      dst[0] = a[0];
      dst[kStemCout] = b[0];
    }
  } else {
    // ------------------------------------------------ epilogue warps 2..5 (TMEM lane quadrant = warp % 4)
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const bool store_thread = (threadIdx.x == 64);
    const uint32_t swz = (row & 7) << 4;
    uint32_t t = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++t) {
      const uint32_t acc = t & 1, bsel = t & 1;
      mbar_wait(&tfull[acc], (t >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * kStemCout + (uint32_t(quad * 32) << 16);
      uint32_t v0[32], v1[32];
      tmem_ld32(taddr, v0);
      tmem_ld32(taddr + 32, v1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);   // accumulator stage drained into registers
      uint8_t* buf = epi + bsel * kStemPanel;
      if (store_thread) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      asm volatile("bar.sync 1, 128;" ::: "memory");   // the buffer's previous store has finished reading it
      if (stats) asm volatile("bar.sync %0, 256;" ::"r"(6 + bsel) : "memory");   // ... and the statistics warps
      if (row >= Wo) {   // rows past the image edge were computed from whatever follows in shared memory
#pragma unroll
        for (int e = 0; e < 32; ++e) v0[e] = v1[e] = 0u;
      }
      if (row < kStemRowsMax) {
        uint8_t* prow = buf + row * 128;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 o;
          o.x = pack_bf16x2(__uint_as_float(v0[8 * q + 0]), __uint_as_float(v0[8 * q + 1]));
          o.y = pack_bf16x2(__uint_as_float(v0[8 * q + 2]), __uint_as_float(v0[8 * q + 3]));
          o.z = pack_bf16x2(__uint_as_float(v0[8 * q + 4]), __uint_as_float(v0[8 * q + 5]));
          o.w = pack_bf16x2(__uint_as_float(v0[8 * q + 6]), __uint_as_float(v0[8 * q + 7]));
          *reinterpret_cast<uint4*>(prow + ((uint32_t(q << 4)) ^ swz)) = o;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          uint4 o;
          o.x = pack_bf16x2(__uint_as_float(v1[8 * q + 0]), __uint_as_float(v1[8 * q + 1]));
          o.y = pack_bf16x2(__uint_as_float(v1[8 * q + 2]), __uint_as_float(v1[8 * q + 3]));
          o.z = pack_bf16x2(__uint_as_float(v1[8 * q + 4]), __uint_as_float(v1[8 * q + 5]));
          o.w = pack_bf16x2(__uint_as_float(v1[8 * q + 6]), __uint_as_float(v1[8 * q + 7]));
          *reinterpret_cast<uint4*>(prow + ((uint32_t((4 + q) << 4)) ^ swz)) = o;
        }
      }
      fence_proxy_async_smem();
      if (stats) {
        __threadfence_block();
        asm volatile("bar.arrive %0, 256;" ::"r"(2 + bsel) : "memory");
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (store_thread) {
        tma_store_3d(buf, &tmC, 0, 0, item);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
  }

  if (threadIdx.x == 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_free<128>(tmem_base);
}

// ---------------------------------------------------------------------------------------------- wgrad
// dWp^T[k, cout] accumulated in TMEM over every output row this CTA owns (K = pixels, 16 per MMA):
//   A (M = k, MN-major, 64-byte swizzle): the same overlapping-window boxes as fprop, [rows x 64 B] per filter row;
//     m-tile 0 = filter rows 0..3, m-tile 1 = filter rows 4..6 plus one phantom panel whose products are discarded.
//   B (N = cout, MN-major, 128-byte swizzle): dy[rows x 128 B].
// warp 0: TMA producer, warp 1: MMA issuer + TMEM owner, warps 2-5: epilogue (once, at the end).
constexpr int kStemWThreads = 192;
constexpr int kStemWDy = kStemRowsMax * 128;                  // 14336
constexpr int kStemWStage = kStemWDy + kStemKH * kStemRowsMax * 64;   // + 7 x 7168
constexpr int kStemWSmem = kStemStages * kStemWStage + kStemRowsMax * 64 /*phantom panel*/ + 1024 + 256;

__global__ void __launch_bounds__(kStemWThreads, 1)
    stem_wgrad_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmDy,
                      float* __restrict__ ws, int Ho, int rows, int num_items) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStemStages * kStemWStage + kStemRowsMax * 64);
  uint64_t* full = bars;
  uint64_t* empty = bars + kStemStages;
  uint64_t* done = bars + 2 * kStemStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t panel = (uint32_t)rows * 64;   // bytes of one [rows x 64 B] window panel
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStemStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(done, 1);
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmDy);
  }
  if (warp == 1) tmem_alloc<128>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t g = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++g) {
        const int n = item / Ho, oh = item - n * Ho;
        const int s = g % kStemStages;
        mbar_wait(&empty[s], ((g / kStemStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&full[s], (uint32_t)rows * 128 + kStemKH * panel);
        uint8_t* st = smem + s * kStemWStage;
        tma_load_3d(st, &tmDy, &full[s], 0, 0, item);
        for (int kh = 0; kh < kStemKH; ++kh)
          tma_load_4d(st + kStemWDy + kh * panel, &tmX, &full[s], 0, 0, 2 * oh + kh, n);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, kStemCout, true, true);
      uint32_t g = 0;
      const int ksteps = rows >> 4;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++g) {
        const int s = g % kStemStages;
        mbar_wait(&full[s], (g / kStemStages) & 1);
        tc_fence_after();
        const uint32_t dyBase = smem_u32(smem + s * kStemWStage), xBase = dyBase + kStemWDy;
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint64_t db = make_smem_desc_sw128(dyBase + ks * 2048, 8192, 1024);
#pragma unroll
          for (int mt = 0; mt < 2; ++mt) {
            const uint64_t da = make_smem_desc(xBase + mt * 4 * panel + ks * 1024, panel, 512, 4);
            umma_bf16(tmem_base + mt * kStemCout, da, db, idesc, (g | (uint32_t)ks) != 0);
          }
        }
        umma_commit(&empty[s]);
      }
      umma_commit(done);
    }
  } else {
    const int quad = warp & 3;
    mbar_wait(done, 0);
    tc_fence_after();
#pragma unroll 1
    for (int mt = 0; mt < 2; ++mt) {
      const int k = mt * 128 + quad * 32 + lane;
      const uint32_t taddr = tmem_base + mt * kStemCout + (uint32_t(quad * 32) << 16);
      float* orow = ws + ((size_t)blockIdx.x * 256 + k) * kStemCout;
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld32(taddr + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 8; ++q)
          st_v4(orow + c * 32 + q * 4, make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_free<128>(tmem_base);
}

static int stem_num_sms() { return b2::host_num_sms(); }

struct StemGeom {
  int Ho, Wo, Hp, Wp;
};
static StemGeom stem_geom(int H, int W) {
  StemGeom g;
  g.Ho = (H + 6 - 7) / 2 + 1;
  g.Wo = (W + 6 - 7) / 2 + 1;
  g.Hp = H + 6;
  g.Wp = (W + 8 + 1) & ~1;   // even: the row pitch (8 B per pixel) must be a multiple of 16 B
  return g;
}

// Overlapping-window view of the padded image: (32 elements = 8 pixels | ow, 16-byte stride | padded row | image)
static int stem_window_map(CUtensorMap* tm, const void* xp, int N, const StemGeom& g, uint32_t box_rows) {
  const uint64_t dims[4] = {32, (uint64_t)g.Wo, (uint64_t)g.Hp, (uint64_t)N};
  const uint64_t strides[3] = {16, (uint64_t)g.Wp * 8, (uint64_t)g.Hp * g.Wp * 8};
  const uint32_t box[4] = {32, box_rows, 1, 1};
  return make_tmap_tiled_bf16(tm, xp, 4, dims, strides, box, 64);
}
// y / dy as (64 channels | ow | (n, oh)): one output row per box, rows past Wo clipped / zero filled
static int stem_rows_map(CUtensorMap* tm, const void* y, int N, const StemGeom& g, uint32_t box_rows) {
  const uint64_t dims[3] = {64, (uint64_t)g.Wo, (uint64_t)N * g.Ho};
  const uint64_t strides[2] = {128, (uint64_t)g.Wo * 128};
  const uint32_t box[3] = {64, box_rows, 1};
  return make_tmap_tiled_bf16(tm, y, 3, dims, strides, box, 128);
}

}  // namespace b2

using namespace b2;

extern "C" int b2_stem_supported(int H, int W) {
  const StemGeom g = stem_geom(H, W);
  return H >= 1 && W >= 1 && g.Wo >= 1 && g.Wo <= kStemRowsMax;
}

extern "C" long long b2_stem_padded_elems(int N, int H, int W) {
  const StemGeom g = stem_geom(H, W);
  return (long long)N * g.Hp * g.Wp * 4;
}

static int stem_pack_input_launch(const void* x_nchw, const void* const* x_slot, int x_is_bf16, void* xp, int N,
                                  int H, int W, void* stream) {
  const StemGeom g = stem_geom(H, W);
  const long long total = (long long)N * g.Hp * g.Wp;
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)stem_num_sms() * 16;
  if (blocks > cap) blocks = cap;
  if (x_is_bf16)
    stem_pack_input_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x_nchw, x_slot, (uint2*)xp, N, H, W, g.Hp, g.Wp);
  else
    stem_pack_input_kernel<float><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((const float*)x_nchw, x_slot,
                                                                                  (uint2*)xp, N, H, W, g.Hp, g.Wp);
  return cudaGetLastError();
}
extern "C" int b2_stem_pack_input(const void* x_nchw, int x_is_bf16, void* xp, int N, int H, int W, void* stream) {
  return stem_pack_input_launch(x_nchw, nullptr, x_is_bf16, xp, N, H, W, stream);
}
extern "C" int b2_stem_pack_input_indirect(const void* const* x_slot, int x_is_bf16, void* xp, int N, int H, int W,
                                           void* stream) {
  if (!x_slot) return cudaErrorInvalidValue;
  return stem_pack_input_launch(nullptr, x_slot, x_is_bf16, xp, N, H, W, stream);
}

extern "C" int b2_stem_pack_weight(const float* w_ohwi, void* wp, void* stream) {
  stem_pack_weight_kernel<<<(kStemCout * kStemK + 255) / 256, 256, 0, (cudaStream_t)stream>>>(w_ohwi,
                                                                                             (__nv_bfloat16*)wp);
  return cudaGetLastError();
}

extern "C" int b2_stem_fprop(const void* xp, const void* wp, void* y, int N, int H, int W, float* stats,
                             int* stats_blocks, void* stream) {
  if (stats_blocks) *stats_blocks = 0;
  if (!b2_stem_supported(H, W)) return cudaErrorInvalidValue;
  const StemGeom g = stem_geom(H, W);
  CUtensorMap tmA, tmB, tmC;
  int e;
  {  // raw padded rows, 8-byte pixels: (pixel | padded row | image), box = 7 whole rows, no swizzle
    const uint64_t dims[3] = {(uint64_t)g.Wp, (uint64_t)g.Hp, (uint64_t)N};
    const uint64_t strides[2] = {(uint64_t)g.Wp * 8, (uint64_t)g.Hp * g.Wp * 8};
    const uint32_t box[3] = {(uint32_t)g.Wp, (uint32_t)kStemKH, 1};
    if ((e = make_tmap_tiled_u64(&tmA, xp, 3, dims, strides, box))) return e;
  }
  {
    const uint64_t dims[2] = {(uint64_t)kStemK, (uint64_t)kStemCout};
    const uint64_t strides[1] = {(uint64_t)kStemK * 2};
    const uint32_t box[2] = {32, (uint32_t)kStemCout};
    if ((e = make_tmap_tiled_bf16(&tmB, wp, 2, dims, strides, box, 64))) return e;
  }
  if ((e = stem_rows_map(&tmC, y, N, g, kStemRowsMax))) return e;
  static bool configured = false;
  if (!configured) {
    cudaError_t ce = cudaFuncSetAttribute(stem_fprop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStemFSmem);
    if (ce != cudaSuccess) return ce;
    configured = true;
  }
  const int items = N * g.Ho;
  const int grid = items < stem_num_sms() ? items : stem_num_sms();
  const bool with_stats = stats && stats_blocks;
  stem_fprop_kernel<<<grid, kStemFThreads, kStemFSmem, (cudaStream_t)stream>>>(
      tmA, tmB, tmC, with_stats ? stats : nullptr, g.Ho, g.Wo, g.Wp * 8, items);
  if (with_stats) *stats_blocks = 2 * grid;
  return cudaGetLastError();
}

extern "C" long long b2_stem_wgrad_workspace_floats(void) { return (long long)stem_num_sms() * 256 * kStemCout; }

extern "C" int b2_stem_wgrad(const void* xp, const void* dy, float* dw_ohwi, float* workspace, int N, int H, int W,
                             int accumulate, void* stream) {
  if (!b2_stem_supported(H, W)) return cudaErrorInvalidValue;
  const StemGeom g = stem_geom(H, W);
  const int rows = (g.Wo + 15) & ~15;
  CUtensorMap tmX, tmDy;
  int e;
  if ((e = stem_window_map(&tmX, xp, N, g, (uint32_t)rows))) return e;
  if ((e = stem_rows_map(&tmDy, dy, N, g, (uint32_t)rows))) return e;
  static bool configured = false;
  if (!configured) {
    cudaError_t ce = cudaFuncSetAttribute(stem_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStemWSmem);
    if (ce != cudaSuccess) return ce;
    configured = true;
  }
  const int items = N * g.Ho;
  const int grid = items < stem_num_sms() ? items : stem_num_sms();
  cudaStream_t st = (cudaStream_t)stream;
  stem_wgrad_kernel<<<grid, kStemWThreads, kStemWSmem, st>>>(tmX, tmDy, workspace, g.Ho, rows, items);
  cudaError_t ce = cudaGetLastError();
  if (ce != cudaSuccess) return ce;
  stem_wgrad_reduce_kernel<<<(kStemCout * 147 + 255) / 256, 256, 0, st>>>(workspace, dw_ohwi, grid, accumulate);
  return cudaGetLastError();
}
