// 3x3 / stride 1 / pad 1 convolution (fprop and dgrad) for the narrow ResNet stages (64 or 128 reduction channels)
// as a WEIGHT-STATIONARY, HALO-REUSING implicit GEMM on tcgen05.
//
// The general kernel (csrc/conv_igemm.cu) fetches the activation operand once per filter tap with TMA im2col: nine
// L2 -> shared-memory reads of every input pixel, plus the weight tile again for every 128 pixels. For 64 -> 64 @ 56x56
// that is 1.35 GB of L2 traffic for 206 MB of unique data and the kernel sits at the L2 bandwidth limit (378 TFLOP/s;
// cuDNN's weight-stationary kernel: 1150, profiles/conv_vs_cudnn_bs256_r2a.md). Here instead
//   * a CTA keeps the weights of its 64 output channels for ALL nine taps resident in shared memory (72 / 144 KB),
//     loaded once per kernel;
//   * a tile is R whole rows of one image, laid out in a zero-padded frame of Wp = W + 2 pixels per row
//     (R = 128 / Wp rows, i.e. 128 MMA rows); its input is ONE TMA box of (R + 2) x Wp pixels x 64 channels (the halo
//     comes from the TMA's out-of-bounds zero fill);
//   * the nine taps are nine tcgen05.mma passes over the SAME slab: in the padded frame the tap offset
//     (r, s) is the constant row shift r * Wp + s, so the A-operand descriptor of a tap is the slab's descriptor with
//     its start address advanced by that many 128-byte rows (the swizzle phase goes into the descriptor's base-offset
//     field) - the generalisation of the stem kernel's overlapping-window descriptors;
//   * rows of the frame that are padding (w >= W, h >= H, or beyond R * Wp) produce junk accumulator rows: the
//     epilogue zeroes them, and the TMA store of the [R][Wp][64] panel clips them against the real tensor extent.
// L2 -> shared traffic per 116 output pixels drops from 216 KB to 30 KB (64 channels).
// The epilogue warps also produce the fused BatchNorm statistics (fprop: sum / sum of squares of the bf16 outputs;
// dgrad: the BatchNorm-backward sums of the unit the gradient flows into), as the general kernel's statistics warps do.
//
// Replaces cuDNN fprop / dgrad for torchvision's conv2 of every layer1 / layer2 block (reference resnet/main.py:76).
#include "b200ddp.h"
#include "ptx.cuh"
#include "tmap.h"

namespace b2 {

constexpr int kHaloThreads = 256;   // warp 0: slab TMA, 1: MMA + TMEM, 2: weight TMA, 3: -, 4..7: epilogue + statistics
constexpr int kHaloWTile = 64 * 128;   // one (tap, k-block) weight tile: 64 x 64 bf16

struct HaloParams {
  int N, H, W, Wp, R;
  int tiles_per_img, m_tiles, n_tiles;
  int KB;            // 64-channel k-blocks of the reduction axis (1 or 2)
  int Ncols;         // GEMM N over the whole layer (fprop: Cout, dgrad: Cin)
  int tapC;          // = Cin: column offset of a tap in the [Cout][9*Cin] weight matrix
  int NS;            // slab ring depth
  int slab_bytes;    // shared memory reserved per slab (128 + 2*Wp + 2 pixel rows, rounded to 1 KB)
  int slab_tx;       // bytes one slab load delivers
  float* stats;      // nullable: [gridDim.x * 4][2][Ncols]
  const __nv_bfloat16* bwd_y;      // dgrad statistics: pre-BN activations of the unit dx feeds [pixels][Ncols]
  const unsigned char* bwd_mask;   // its ReLU mask, 1 bit / element (nullable)
  const float* bwd_mean;
  int dbg;           // bit 0: leave the descriptor's base-offset field zero (hardware probe)
};

B2_DEVICE void tma_store_4d(const void* smem_src, const CUtensorMap* m, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
// K-major, 128-byte swizzle, start address on any 128-byte row of the swizzle pattern: the row phase goes into the
// matrix-descriptor base offset (bits 49..51).
B2_DEVICE uint64_t make_smem_desc_sw128_row(uint32_t addr, bool with_base_offset) {
  uint64_t d = make_smem_desc_sw128(addr, 16, 1024);
  if (with_base_offset) d |= uint64_t((addr >> 7) & 7) << 49;
  return d;
}

template <bool DGRAD, int KBT>
__global__ void __launch_bounds__(kHaloThreads, 1)
    conv3x3_halo_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                        const __grid_constant__ CUtensorMap tmC, const HaloParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* wres = smem;                                   // [9][KBT] weight tiles, resident
  uint8_t* slabs = wres + 9 * KBT * kHaloWTile;           // [NS] input slabs
  uint8_t* panels = slabs + p.NS * p.slab_bytes;          // [2] output panels, 128 rows x 128 B
  uint64_t* bars = reinterpret_cast<uint64_t*>(panels + 2 * 128 * 128);
  uint64_t* wfull = bars;
  uint64_t* sfull = bars + 1;
  uint64_t* sempty = sfull + 4;
  uint64_t* tfull = sempty + 4;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x % p.n_tiles;
  const int n0 = n_tile * 64;
  const int m_first = blockIdx.x / p.n_tiles, m_step = gridDim.x / p.n_tiles;

  if (threadIdx.x == 0) {
    mbar_init(wfull, 1);
    for (int i = 0; i < 4; ++i) {
      mbar_init(&sfull[i], 1);
      mbar_init(&sempty[i], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 4);   // one arrival per epilogue warp
    }
    fence_mbar_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
  }
  if (warp == 1) tmem_alloc<128>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 2) {
    // ===================================================== weights: all nine taps of this CTA's 64 columns, once
    if (lane == 0) {
      mbar_arrive_expect_tx(wfull, 9 * KBT * kHaloWTile);
      for (int tap = 0; tap < 9; ++tap)
        for (int cb = 0; cb < KBT; ++cb) {
          uint8_t* dst = wres + (tap * KBT + cb) * kHaloWTile;
          if (!DGRAD) tma_load_2d(dst, &tmB, wfull, tap * p.tapC + cb * 64, n0);        // [cout n0..][cin block]
          else tma_load_2d(dst, &tmB, wfull, tap * p.tapC + n0, cb * 64);               // [cout block][cin n0..]
        }
    }
  } else if (warp == 0) {
    // ===================================================== input slabs
    if (lane == 0) {
      uint32_t g = 0;
      for (int mt = m_first; mt < p.m_tiles; mt += m_step) {
        const int img = mt / p.tiles_per_img, h0 = (mt - img * p.tiles_per_img) * p.R;
        for (int cb = 0; cb < KBT; ++cb, ++g) {
          const int s = g % p.NS;
          mbar_wait(&sempty[s], ((g / p.NS) & 1) ^ 1);
          mbar_arrive_expect_tx(&sfull[s], p.slab_tx);
          tma_load_4d(slabs + s * p.slab_bytes, &tmA, &sfull[s], cb * 64, -1, h0 - 1, img);
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 64, false, DGRAD);
      const bool bo = !(p.dbg & 1);
      mbar_wait(wfull, 0);
      uint32_t g = 0, t = 0;
      for (int mt = m_first; mt < p.m_tiles; mt += m_step, ++t) {
        const uint32_t acc = t & 1;
        mbar_wait(&tempty[acc], ((t >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 64;
        for (int cb = 0; cb < KBT; ++cb, ++g) {
          const int s = g % p.NS;
          mbar_wait(&sfull[s], (g / p.NS) & 1);
          tc_fence_after();
          const uint32_t slab = smem_u32(slabs + s * p.slab_bytes);
#pragma unroll 1
          for (int tap = 0; tap < 9; ++tap) {
            const int r = tap / 3, c = tap - r * 3;
            const int shift = DGRAD ? (2 - r) * p.Wp + (2 - c) : r * p.Wp + c;
            const uint32_t aBase = slab + (uint32_t)shift * 128;
            const uint32_t bBase = smem_u32(wres + (tap * KBT + cb) * kHaloWTile);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t da = make_smem_desc_sw128_row(aBase + k * 32, bo);
              const uint64_t db = DGRAD ? make_smem_desc_sw128(bBase + k * 2048, 8192, 1024)
                                        : make_smem_desc_sw128(bBase + k * 32, 16, 1024);
              umma_bf16(d_tmem, da, db, idesc, (cb | tap | k) != 0);
            }
          }
          umma_commit(&sempty[s]);
        }
        umma_commit(&tfull[acc]);
      }
    }
  } else if (warp >= 4) {
    // ===================================================== epilogue + statistics (TMEM lane quadrant = warp - 4)
    const int quad = warp - 4;
    const int row = quad * 32 + lane;             // accumulator row = pixel index in the padded frame of the tile
    const int e = threadIdx.x - 128;              // statistics role: 16-byte chunk ch8 of rows rg, rg + 16, ...
    const int ch8 = e & 7, rg = e >> 3;
    const bool store_thread = (threadIdx.x == 128);
    const bool want_stats = p.stats != nullptr;
    float sa[8], sb[8], mu[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) sa[j] = sb[j] = mu[j] = 0.f;
    if (want_stats) {
      float* z = p.stats + (size_t)(blockIdx.x * 4 + quad) * 2 * p.Ncols;
      for (int i = lane; i < 2 * p.Ncols; i += 32) z[i] = 0.f;
      if (DGRAD) {
#pragma unroll
        for (int j = 0; j < 8; ++j) mu[j] = __ldg(p.bwd_mean + n0 + ch8 * 8 + j);
      }
      __syncwarp();
    }
    const int rr_row = row / p.Wp, w_row = row - rr_row * p.Wp;
    uint32_t t = 0;
    for (int mt = m_first; mt < p.m_tiles; mt += m_step, ++t) {
      const int img = mt / p.tiles_per_img, h0 = (mt - img * p.tiles_per_img) * p.R;
      const uint32_t acc = t & 1;
      // dgrad statistics: the companion values of this thread's 8 (row, chunk) cells, requested before the wait
      uint4 yv[8];
      uint32_t mk[8];
      if (DGRAD && want_stats) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int r = rg + 16 * j, rr = r / p.Wp, ww = r - rr * p.Wp;
          const bool ok = rr < p.R && ww < p.W && h0 + rr < p.H;
          const size_t off = ((size_t)(img * p.H + h0 + rr) * p.W + ww) * p.Ncols + n0 + ch8 * 8;
          yv[j] = ok ? ld_nc_v4(p.bwd_y + off) : make_uint4(0u, 0u, 0u, 0u);
          mk[j] = !ok ? 0u : (p.bwd_mask ? (uint32_t)__ldg(p.bwd_mask + (off >> 3)) : 0xffu);
        }
      }
      mbar_wait(&tfull[acc], (t >> 1) & 1);
      tc_fence_after();
      uint32_t v[64];
      {
        uint32_t lo[32], hi[32];
        const uint32_t taddr = tmem_base + acc * 64 + (uint32_t(quad * 32) << 16);
        tmem_ld32(taddr, lo);
        tmem_ld32(taddr + 32, hi);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          v[j] = lo[j];
          v[32 + j] = hi[j];
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);   // the accumulator is in registers: the MMA warp may reuse it
      const bool valid = rr_row < p.R && w_row < p.W && h0 + rr_row < p.H;
      uint4 o[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        o[q].x = valid ? pack_bf16x2(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1])) : 0u;
        o[q].y = valid ? pack_bf16x2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3])) : 0u;
        o[q].z = valid ? pack_bf16x2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5])) : 0u;
        o[q].w = valid ? pack_bf16x2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7])) : 0u;
      }
      uint8_t* panel = panels + (t & 1) * (128 * 128);
      // the panel buffer's previous TMA store (two tiles ago) has finished reading it
      if (store_thread) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      asm volatile("bar.sync 1, 128;" ::: "memory");
      {
        uint8_t* prow = panel + row * 128;
        const uint32_t swz = (row & 7) << 4;
#pragma unroll
        for (int q = 0; q < 8; ++q) *reinterpret_cast<uint4*>(prow + ((uint32_t(q << 4)) ^ swz)) = o[q];
      }
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (store_thread) {
        tma_store_4d(panel, &tmC, n0, 0, h0, img);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      if (want_stats) {
        const uint32_t src = smem_u32(panel) + rg * 128 + ((ch8 ^ (rg & 7)) << 4);   // (rg + 16 j) & 7 == rg & 7
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint4 w4 = ld_shared_v4(src + j * 16 * 128);
          const uint32_t ww[4] = {w4.x, w4.y, w4.z, w4.w};
          if (!DGRAD) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float lo = __uint_as_float(ww[q] << 16), hi = __uint_as_float(ww[q] & 0xffff0000u);
              sa[2 * q] += lo;
              sb[2 * q] = fmaf(lo, lo, sb[2 * q]);
              sa[2 * q + 1] += hi;
              sb[2 * q + 1] = fmaf(hi, hi, sb[2 * q + 1]);
            }
          } else {
            const uint32_t yy[4] = {yv[j].x, yv[j].y, yv[j].z, yv[j].w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float glo = ((mk[j] >> (2 * q)) & 1u) ? __uint_as_float(ww[q] << 16) : 0.f;
              const float ghi = ((mk[j] >> (2 * q + 1)) & 1u) ? __uint_as_float(ww[q] & 0xffff0000u) : 0.f;
              sa[2 * q] += glo;
              sb[2 * q] = fmaf(glo, __uint_as_float(yy[q] << 16) - mu[2 * q], sb[2 * q]);
              sa[2 * q + 1] += ghi;
              sb[2 * q + 1] = fmaf(ghi, __uint_as_float(yy[q] & 0xffff0000u) - mu[2 * q + 1], sb[2 * q + 1]);
            }
          }
        }
      }
    }
    if (want_stats) {
      // lanes with the same chunk (lane & 7) fold their row groups together, fixed order: deterministic
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        sa[j] += __shfl_xor_sync(0xffffffffu, sa[j], 8);
        sb[j] += __shfl_xor_sync(0xffffffffu, sb[j], 8);
        sa[j] += __shfl_xor_sync(0xffffffffu, sa[j], 16);
        sb[j] += __shfl_xor_sync(0xffffffffu, sb[j], 16);
      }
      if (lane < 8) {
        float* dst = p.stats + (size_t)(blockIdx.x * 4 + quad) * 2 * p.Ncols + n0 + lane * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          dst[j] = sa[j];
          dst[p.Ncols + j] = sb[j];
        }
      }
    }
    if (store_thread) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_free<128>(tmem_base);
}

}  // namespace b2
