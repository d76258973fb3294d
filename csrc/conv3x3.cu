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
//     its start address advanced by that many 128-byte rows. Measured on B200 (tools/halo_probe.py): the 128-byte
//     swizzle is a pure function of the shared-memory ADDRESS bits, for the tensor core's reads as for the TMA's
//     writes, so a start address on any row of the 1 KB pattern works with the descriptor's base-offset field left
//     zero (an early version that put the row phase there produced garbage; with it zero the output is bit-identical
//     to the general kernel's, and unaligned windows cost no time) - the swizzled generalisation of the stem kernel's
//     overlapping-window descriptors;
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

constexpr int kHaloThreads = 384;   // warp 0: slab TMA, 1: MMA + TMEM, 2: weight TMA, 3: -, 4..11: epilogue + statistics
constexpr int kHaloWTile = 64 * 128;   // one (tap, k-block) weight tile: 64 x 64 bf16
// Accumulator stages in TMEM (64 columns each). A tile is short (36 or 72 MMAs of N = 64), and the round trip
// tcgen05.commit -> epilogue wakes -> tcgen05.ld -> arrive -> MMA warp wakes measured ~2.9 us here: with two stages the
// tensor pipe idled 1.45 us per tile ("handshakes only" probe, tools/halo_probe.py); eight stages hide it.
constexpr int kHaloAcc = 8;

struct HaloParams {
  int N, H, W, Wp, R;
  int tiles_per_img, m_tiles, n_tiles;
  int KB;            // 64-channel k-blocks of the reduction axis (1 or 2)
  int Ncols;         // GEMM N over the whole layer (fprop: Cout, dgrad: Cin)
  int tapC;          // = Cin: column offset of a tap in the [Cout][9*Cin] weight matrix
  int NS;            // slab ring depth
  int slab_bytes;    // shared memory reserved per slab (128 + 2*Wp + 2 pixel rows, rounded to 1 KB)
  int slab_tx;       // bytes one slab load delivers
  float* stats;      // nullable: [gridDim.x][2][Ncols]
  const __nv_bfloat16* bwd_y;      // dgrad statistics: pre-BN activations of the unit dx feeds [pixels][Ncols]
  const unsigned char* bwd_mask;   // its ReLU mask, 1 bit / element (nullable)
  const float* bwd_mean;
  unsigned long long* prof;   // optional [gridDim.x][8] cycle counters (tools/halo_probe.py profile)
  int dbg;           // timing probes (results are wrong): bit 5 no slab loads once the ring is primed, bit 6 no stores,
                     // bit 7 epilogue reads the accumulator out and nothing else
};

B2_DEVICE void tma_store_4d(const void* smem_src, const CUtensorMap* m, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
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
  uint64_t* tempty = tfull + kHaloAcc;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + kHaloAcc);

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
    for (int a = 0; a < kHaloAcc; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 8);   // one arrival per epilogue warp
    }
    fence_mbar_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
  }
  if (warp == 1) tmem_alloc<64 * kHaloAcc>(tmem_slot);
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
      long long w_empty = 0;
      for (int mt = m_first; mt < p.m_tiles; mt += m_step) {
        const int img = mt / p.tiles_per_img, h0 = (mt - img * p.tiles_per_img) * p.R;
        for (int cb = 0; cb < KBT; ++cb, ++g) {
          const int s = g % p.NS;
          const long long c0 = clock64();
          mbar_wait(&sempty[s], ((g / p.NS) & 1) ^ 1);
          w_empty += clock64() - c0;
          if ((p.dbg & 32) && g >= (uint32_t)p.NS) {   // timing probe: no slab loads once the ring is primed
            mbar_arrive(&sfull[s]);
            continue;
          }
          mbar_arrive_expect_tx(&sfull[s], p.slab_tx);
          tma_load_4d(slabs + s * p.slab_bytes, &tmA, &sfull[s], cb * 64, -1, h0 - 1, img);
        }
      }
      if (p.prof) p.prof[blockIdx.x * 8 + 0] = (unsigned long long)w_empty;
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, 64, false, DGRAD);
      // constant parts of the shared-memory descriptors (128-byte swizzle, SBO = 1024; LBO = 16 for the K-major
      // operands, 8192 for the MN-major weight tile of dgrad) and the per-tap row shifts of the A window
      const uint32_t desc_hi = (uint32_t)(make_smem_desc_sw128(0, 16, 1024) >> 32);
      const uint32_t b_lo0 = ((smem_u32(wres) >> 4) & 0x3FFFu) | ((DGRAD ? (8192u >> 4) : 1u) << 16);
      uint32_t tap_off[9];
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const int r = tap / 3, c = tap - r * 3;
        const int shift = DGRAD ? (2 - r) * p.Wp + (2 - c) : r * p.Wp + c;
        tap_off[tap] = (uint32_t)shift * 8;            // rows of 128 bytes, in 16-byte units
      }
      long long w_acc = 0, w_slab = 0, w_issue = 0;
      const long long k0 = clock64();
      mbar_wait(wfull, 0);
      const long long k1 = clock64();
      uint32_t g = 0, t = 0;
      for (int mt = m_first; mt < p.m_tiles; mt += m_step, ++t) {
        const uint32_t acc = t % kHaloAcc;
        long long c0 = clock64();
        mbar_wait(&tempty[acc], ((t / kHaloAcc) & 1) ^ 1);
        w_acc += clock64() - c0;
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 64;
        for (int cb = 0; cb < KBT; ++cb, ++g) {
          const int s = g % p.NS;
          c0 = clock64();
          mbar_wait(&sfull[s], (g / p.NS) & 1);
          const long long c1 = clock64();
          w_slab += c1 - c0;
          tc_fence_after();
          // descriptors: high words are constant, low words = (address >> 4) | LBO field; a tap / k-step only adds a
          // small constant to the low word (no carry out of the 14-bit address field: shared memory is < 256 KB).
          // One thread issues every MMA of the CTA, so the issue loop has to be a handful of instructions per MMA:
          // with the descriptors rebuilt from scratch it took ~95 cycles per MMA and starved the tensor pipe.
          const uint32_t a_lo = ((smem_u32(slabs + s * p.slab_bytes) >> 4) & 0x3FFFu) | (1u << 16);
          const uint32_t b_lo = b_lo0 + (uint32_t)cb * (kHaloWTile >> 4);
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint32_t alo = a_lo + tap_off[tap] + k * 2;
              const uint32_t blo = b_lo + tap * KBT * (kHaloWTile >> 4) + (DGRAD ? k * 128 : k * 2);
              const uint64_t da = ((uint64_t)desc_hi << 32) | alo;
              const uint64_t db = ((uint64_t)desc_hi << 32) | blo;
              umma_bf16(d_tmem, da, db, idesc, (cb | tap | k) != 0);
            }
          }
          umma_commit(&sempty[s]);
          w_issue += clock64() - c1;
        }
        umma_commit(&tfull[acc]);
      }
      if (p.prof) {
        p.prof[blockIdx.x * 8 + 1] = (unsigned long long)w_acc;
        p.prof[blockIdx.x * 8 + 2] = (unsigned long long)w_slab;
        p.prof[blockIdx.x * 8 + 3] = (unsigned long long)w_issue;
        p.prof[blockIdx.x * 8 + 4] = (unsigned long long)(k1 - k0);
        p.prof[blockIdx.x * 8 + 5] = (unsigned long long)(clock64() - k1);
      }
    }
  } else if (warp >= 4) {
    // ===================================================== epilogue + statistics: two groups of four warps, group eg
    // converts columns [32 eg, 32 eg + 32) of the 64-column panel (TMEM lane quadrant = warp % 4)
    const int quad = warp & 3, eg = (warp - 4) >> 2;
    const int row = quad * 32 + lane;             // accumulator row = pixel index in the padded frame of the tile
    const int e = threadIdx.x - 128;              // statistics role: 16-byte chunk ch8 of rows rg, rg + 32, ...
    const int ch8 = e & 7, rg = e >> 3;           // rg in [0, 32)
    const bool store_thread = (threadIdx.x == 128);
    const bool want_stats = p.stats != nullptr;
    float sa[8], sb[8], mu[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) sa[j] = sb[j] = mu[j] = 0.f;
    if (want_stats && DGRAD) {
#pragma unroll
      for (int j = 0; j < 8; ++j) mu[j] = __ldg(p.bwd_mean + n0 + ch8 * 8 + j);
    }
    const int rr_row = row / p.Wp, w_row = row - rr_row * p.Wp;
    long long w_full = 0;
    const long long ep0 = clock64();
    uint32_t t = 0;
    for (int mt = m_first; mt < p.m_tiles; mt += m_step, ++t) {
      const int img = mt / p.tiles_per_img, h0 = (mt - img * p.tiles_per_img) * p.R;
      const uint32_t acc = t % kHaloAcc;
      // dgrad statistics: the companion values of this thread's 4 (row, chunk) cells, requested before the wait
      uint4 yv[4];
      uint32_t mk[4];
      if (DGRAD && want_stats) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int r = rg + 32 * j, rr = r / p.Wp, ww = r - rr * p.Wp;
          const bool ok = rr < p.R && ww < p.W && h0 + rr < p.H;
          const size_t off = ((size_t)(img * p.H + h0 + rr) * p.W + ww) * p.Ncols + n0 + ch8 * 8;
          yv[j] = ok ? ld_nc_v4(p.bwd_y + off) : make_uint4(0u, 0u, 0u, 0u);
          mk[j] = !ok ? 0u : (p.bwd_mask ? (uint32_t)__ldg(p.bwd_mask + (off >> 3)) : 0xffu);
        }
      }
      const long long e0 = clock64();
      mbar_wait(&tfull[acc], (t / kHaloAcc) & 1);
      w_full += clock64() - e0;
      tc_fence_after();
      uint32_t v[32];
      tmem_ld32(tmem_base + acc * 64 + eg * 32 + (uint32_t(quad * 32) << 16), v);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);   // the accumulator is in registers: the MMA warp may reuse it
      if (p.dbg & 128) continue;                    // timing probe: accumulator read-out only
      const bool valid = rr_row < p.R && w_row < p.W && h0 + rr_row < p.H;
      uint4 o[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        o[q].x = valid ? pack_bf16x2(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1])) : 0u;
        o[q].y = valid ? pack_bf16x2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3])) : 0u;
        o[q].z = valid ? pack_bf16x2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5])) : 0u;
        o[q].w = valid ? pack_bf16x2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7])) : 0u;
      }
      uint8_t* panel = panels + (t & 1) * (128 * 128);
      // the panel buffer's previous TMA store (two tiles ago) has finished reading it
      if (store_thread) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      asm volatile("bar.sync 1, 256;" ::: "memory");
      {
        uint8_t* prow = panel + row * 128;
        const uint32_t swz = (row & 7) << 4;
#pragma unroll
        for (int q = 0; q < 4; ++q) *reinterpret_cast<uint4*>(prow + ((uint32_t((eg * 4 + q) << 4)) ^ swz)) = o[q];
      }
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (store_thread && !(p.dbg & 64)) {          // (bit 6: timing probe without the stores)
        tma_store_4d(panel, &tmC, n0, 0, h0, img);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      if (want_stats) {
        const uint32_t src = smem_u32(panel) + rg * 128 + ((ch8 ^ (rg & 7)) << 4);   // (rg + 32 j) & 7 == rg & 7
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint4 w4 = ld_shared_v4(src + j * 32 * 128);
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
    if (store_thread && p.prof) {
      p.prof[blockIdx.x * 8 + 6] = (unsigned long long)w_full;
      p.prof[blockIdx.x * 8 + 7] = (unsigned long long)(clock64() - ep0);
    }
    if (store_thread) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    if (want_stats) {
      // lanes with the same chunk (lane & 7) fold their row groups together, then the eight warps are folded in warp
      // order through shared memory (panel 0 is free by now): fixed order, deterministic; ONE partial row per CTA
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        sa[j] += __shfl_xor_sync(0xffffffffu, sa[j], 8);
        sb[j] += __shfl_xor_sync(0xffffffffu, sb[j], 8);
        sa[j] += __shfl_xor_sync(0xffffffffu, sa[j], 16);
        sb[j] += __shfl_xor_sync(0xffffffffu, sb[j], 16);
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");      // every thread is done with the panels (and the last store)
      float* red = reinterpret_cast<float*>(panels);       // [8 warps][2][64]
      if (lane < 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          red[((warp - 4) * 2 + 0) * 64 + lane * 8 + j] = sa[j];
          red[((warp - 4) * 2 + 1) * 64 + lane * 8 + j] = sb[j];
        }
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      float* dst = p.stats + (size_t)blockIdx.x * 2 * p.Ncols;
      for (int i = e; i < 2 * p.Ncols; i += 256) {
        const int which = i / p.Ncols, col = i - which * p.Ncols;
        float tot = 0.f;
        if (col >= n0 && col < n0 + 64) {
#pragma unroll
          for (int wv = 0; wv < 8; ++wv) tot += red[(wv * 2 + which) * 64 + (col - n0)];
        }
        dst[i] = tot;                                      // columns of other CTAs' tiles: zero
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_free<64 * kHaloAcc>(tmem_base);
}

}  // namespace b2

using namespace b2;

static int g_halo_enabled = 1;   // b2_conv_set_halo(0) routes every 3x3 layer through the general kernel again
static int g_halo_dbg = 0;
static unsigned long long* g_halo_prof = nullptr;
extern "C" void b2_conv_set_halo_profile(void* dev_buf) { g_halo_prof = (unsigned long long*)dev_buf; }
extern "C" void b2_conv_set_halo(int on) { g_halo_enabled = on; }
extern "C" void b2_conv_set_halo_debug(int bits) { g_halo_dbg = bits; }

// Shapes this kernel takes: 3x3 / stride 1 / pad 1, reduction channels 64 or 128 (the nine taps of 64 output columns
// then fit in shared memory next to the slab ring), output columns a multiple of 64, rows of at most 126 pixels, and a
// padded frame that fills at least three quarters of the 128 MMA rows.
extern "C" int b2_conv_halo_supported(int N, int H, int W, int K, int Ncols) {
  if (!g_halo_enabled || N <= 0) return 0;
  // 64 reduction channels: 2x the general kernel (0.084 vs 0.157 ms, 64 -> 64 @ 56x56, batch 256). With 128 the two
  // kernels tie (0.079 vs 0.081 ms: the N = 64 MMAs are bound by shared-memory operand reads, 63 cycles each), so
  // those layers stay on the general kernel unless b2_conv_set_halo(2) asks for this one.
  if (K != 64 && !(K == 128 && g_halo_enabled >= 2)) return 0;
  if (Ncols % 64 || Ncols <= 0) return 0;
  const int Wp = W + 2;
  if (Wp > 128) return 0;
  int R = 128 / Wp;
  if (R > H) R = H;
  return R * Wp >= 96;
}

// dgrad = false: src = x [N][H][W][Cin], dst = y [N][H][W][Cout], K = Cin, Ncols = Cout.
// dgrad = true : src = dy [N][H][W][Cout], dst = dx [N][H][W][Cin], K = Cout, Ncols = Cin.
// w is always the OHWI weight [Cout][3][3][Cin] (bf16). stats / stats_blocks as in b2_conv_fprop_stats /
// b2_conv_dgrad_bnstats: one partial row per CTA.
extern "C" int b2_conv3x3_halo(int dgrad, const void* src, const void* w, void* dst, int N, int H, int W, int Cin,
                               int Cout, float* stats, int* stats_blocks, const void* bwd_y, const void* bwd_mask,
                               const float* bwd_mean, void* stream) {
  if (stats_blocks) *stats_blocks = 0;
  const int K = dgrad ? Cout : Cin, Ncols = dgrad ? Cin : Cout;
  if (!b2_conv_halo_supported(N, H, W, K, Ncols)) return cudaErrorInvalidValue;
  HaloParams p{};
  p.N = N; p.H = H; p.W = W; p.Wp = W + 2;
  p.R = 128 / p.Wp;
  if (p.R > H) p.R = H;
  p.tiles_per_img = (H + p.R - 1) / p.R;
  p.m_tiles = N * p.tiles_per_img;
  p.n_tiles = Ncols / 64;
  p.KB = K / 64;
  p.Ncols = Ncols;
  p.tapC = Cin;
  const int slab_px = (128 + 2 * p.Wp + 2 + 7) / 8 * 8;
  p.slab_bytes = slab_px * 128;
  p.slab_tx = 64 * p.Wp * (p.R + 2) * 2;
  const int fixed = 9 * p.KB * kHaloWTile + 2 * 128 * 128 + 256 + 1024;
  int ns = (227 * 1024 - fixed) / p.slab_bytes;
  if (ns > 4) ns = 4;
  if (ns < 2) return cudaErrorInvalidValue;
  p.NS = ns;
  p.dbg = g_halo_dbg;
  p.prof = g_halo_prof;
  const size_t smem = (size_t)fixed + (size_t)ns * p.slab_bytes;

  CUtensorMap tmA, tmB, tmC;
  int e;
  {  // source slab: (channels | w | h | n), box = 64 channels x Wp pixels x (R + 2) rows, zero fill outside the image
    const uint64_t dims[4] = {(uint64_t)K, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    const uint64_t strides[3] = {(uint64_t)K * 2, (uint64_t)W * K * 2, (uint64_t)H * W * K * 2};
    const uint32_t box[4] = {64, (uint32_t)p.Wp, (uint32_t)(p.R + 2), 1};
    if ((e = make_tmap_tiled_bf16(&tmA, src, 4, dims, strides, box, 128))) return e;
  }
  if ((e = make_tmap_2d_bf16(&tmB, w, (uint64_t)9 * Cin, (uint64_t)Cout, 64, 64))) return e;
  {  // destination: the [R][Wp][64] panel; the columns w >= W and rows h >= H of the padded frame are clipped
    const uint64_t dims[4] = {(uint64_t)Ncols, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    const uint64_t strides[3] = {(uint64_t)Ncols * 2, (uint64_t)W * Ncols * 2, (uint64_t)H * W * Ncols * 2};
    const uint32_t box[4] = {64, (uint32_t)p.Wp, (uint32_t)p.R, 1};
    if ((e = make_tmap_tiled_bf16(&tmC, dst, 4, dims, strides, box, 128))) return e;
  }
  int grid = host_num_sms() / p.n_tiles * p.n_tiles;
  if (grid < p.n_tiles) return cudaErrorInvalidValue;
  if (grid > p.m_tiles * p.n_tiles) grid = p.m_tiles * p.n_tiles;
  if (stats && stats_blocks) {
    p.stats = stats;
    p.bwd_y = (const __nv_bfloat16*)bwd_y;
    p.bwd_mask = (const unsigned char*)bwd_mask;
    p.bwd_mean = bwd_mean;
    *stats_blocks = grid;
  }
  cudaStream_t st = (cudaStream_t)stream;
#define B2_HALO_LAUNCH(DG, KBT)                                                                                    \
  do {                                                                                                             \
    static bool configured = false;                                                                                \
    if (!configured) {                                                                                             \
      cudaError_t ce = cudaFuncSetAttribute(conv3x3_halo_kernel<DG, KBT>,                                          \
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);              \
      if (ce != cudaSuccess) return ce;                                                                            \
      configured = true;                                                                                           \
    }                                                                                                              \
    conv3x3_halo_kernel<DG, KBT><<<grid, kHaloThreads, smem, st>>>(tmA, tmB, tmC, p);                               \
  } while (0)
  if (!dgrad && p.KB == 1) B2_HALO_LAUNCH(false, 1);
  else if (!dgrad) B2_HALO_LAUNCH(false, 2);
  else if (p.KB == 1) B2_HALO_LAUNCH(true, 1);
  else B2_HALO_LAUNCH(true, 2);
#undef B2_HALO_LAUNCH
  return cudaGetLastError();
}
