// Implicit-GEMM convolution for sm_100a: tcgen05.mma with TMEM accumulators, weights (and 1x1 activations)
// staged by TMA, im2col rows of k>1 / strided convolutions gathered by producer warps with zero-filling
// cp.async straight into the 128B-swizzled UMMA layout.  One kernel template serves
//   fprop : y[N*Ho*Wo, Cout]   = im2col(x)[.., taps*Cin]      * W[Cout, taps*Cin]^T     (A K-major, B K-major)
//   dgrad : dx[N*H*W, Cin]     = im2col_T(dy)[.., taps*Cout]  * W[(tap,Cout), Cin]      (A K-major, B MN-major)
//   wgrad : dW[Cout, tap, Cin] = dy[pixels, Cout]^T           * im2col(x)[pixels, Cin]  (A MN-major, B MN-major)
// Layouts: activations NHWC bf16, weights OHWI bf16 (== torch channels_last of an OIHW tensor), dW fp32 OHWI.
//
// Replaces cuDNN fprop/dgrad/wgrad used by the reference through torchvision's ResNet
// (reference: resnet/main.py:76,121,123 -> aten::convolution / convolution_backward).
#include "b200ddp.h"
#include "ptx.cuh"
#include "tmap.h"

namespace b2 {

enum ConvKind { FPROP = 0, DGRAD = 1, WGRAD = 2 };

struct ConvParams {
  // Gathered activation tensor (x for fprop/wgrad, dy for dgrad), NHWC bf16.
  const __nv_bfloat16* src;
  int srcH, srcW, srcC;
  // Pixel grid that GEMM rows (fprop/dgrad: M rows; wgrad: K rows) enumerate: [N, rowH, rowW].
  int rowH, rowW;
  int KH, KW, stride, pad;
  int M;        // GEMM M  (fprop/dgrad: pixels; wgrad: Cout)
  int Ncols;    // GEMM N  (fprop: Cout; dgrad: Cin; wgrad: Cin)
  int Kblocks;  // number of 64-wide k-blocks in total (wgrad: over all pixels)
  int cpb;      // 64-channel blocks per tap of the K axis (fprop: Cin/64, dgrad: Cout/64); wgrad: unused
  int Kpix;     // wgrad: number of pixels (K extent)
  int splits;   // wgrad: split-K factor
  int kb_per_split;
  int im2col_pad;  // TMA im2col: padding of the (possibly transposed) convolution the descriptor encodes
  int tapC;     // channels per tap in the weight's K axis (= Cin): column offset of a tap in W2d
  const float* bias;  // fprop only, optional: added per output column (classifier)
  void* out;    // fprop/dgrad: bf16 [M, Ncols]; wgrad: fp32 workspace [splits][Cout][taps*Cin]
  long long out_ld;
  long long split_stride;  // wgrad: elements between split slabs
  // ---- persistent kernel (v2)
  int n_tiles, m_tiles, taps;      // work-item grid: item -> (n tile fastest, m tile, tap, split)
  int im2col_pad_h, im2col_pad_w;  // = -lower corner of the im2col bounding box
  // dgrad tap walk: kh = kh0 + kstep*th (th < nth), same for w. Strided dgrad runs once per output-pixel parity
  // class (ph, pw): rows enumerate the class grid [N, rowH, rowW] and land on pixel (cls*i+ph, cls*j+pw) of dx.
  int kh0, kw0, kstep, nth, ntw;
  int cls, ph, pw, outH, outW;
  // wgrad, k>1: the N axis is the flattened (tap, cin) axis of dW; every 64-column box of a tile belongs to one tap,
  // so one dy tile is reused for up to four taps (flatN = 1).
  int flatN;
  // fprop (v2): per-CTA partial BatchNorm statistics of the bf16 output, [gridDim][2][Ncols] fp32 (sum, sum of
  // squares), accumulated tile by tile in a fixed order -> deterministic. nullptr = not requested.
  float* stats;
  // dgrad fan-in fusion (v2, TMA-store path): dx = relu_mask ? (dgrad + add) : 0 — the gradient fan-in of a residual
  // fork and the ReLU backward of the tensor being differentiated, done by the statistics warps on the bf16 panel in
  // shared memory (coalesced companion loads), together with the BatchNorm-backward sums of that tensor.
  // dgrad + BatchNorm-backward statistics (v2, TMA-store path): the statistics warps also accumulate, per channel of
  // dx, sum(g) and sum(g * (y - mean)) with g = relu_mask ? dx : 0 into `stats` - the reduce pass of the BatchNorm
  // whose output gradient this dgrad produces.
  const __nv_bfloat16* bwd_y;       // [M][Ncols] pre-BN activations of that BatchNorm
  const unsigned char* bwd_mask;    // [M][Ncols/8] ReLU mask (nullable: no ReLU)
  const float* bwd_mean;            // [Ncols]
  int fanin;                    // 1: the statistics warps add `add`, apply `amask`, write the panel back and store it
  const __nv_bfloat16* add;     // [M][Ncols] bf16, optional
  const unsigned char* amask;   // [M][Ncols/8] bits, optional (only with add)
  int panels;  // v2, MN-major operands fetched as one multi-panel TMA box: bit 0 = A (wgrad), bit 1 = B (dgrad/wgrad)
  int dbg;  // timing experiments only (b2_conv_set_debug): 1 = skip the TMA loads once the ring is primed, 2 = skip the MMAs
  // wgrad (v2): 0 = every split stores its fp32 slab (wgrad_reduce_kernel sums them in a fixed order afterwards);
  // 1 = `out` is dW itself and every split adds its tile with red.global.add.v4.f32 (the sum forms in L2, no slab
  // round trip, no second kernel; the order of the fp32 additions is whatever the memory system makes it).
  int red_out;
};

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kGatherThreads = 128;
constexpr int kThreads = 192;  // warps 0-3 gather + epilogue, warp 4 TMA, warp 5 MMA/TMEM

template <int BN>
struct Cfg {
  static constexpr int kStages = (BN == 256) ? 4 : (BN == 128 ? 3 : 4);
  // Gather warps publish a stage this many k-blocks after issuing its cp.asyncs (loads stay in flight).
  static constexpr int kLag = kStages - 2;
  static constexpr int kABytes = BM * 128;  // 16 KB either major-ness
  static constexpr int kBBytes = BN * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kSmem = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

// Decode a pixel index into (n, h, w) for a [*, PH, PW] grid.
B2_DEVICE void decode_pix(int p, int PH, int PW, int& n, int& h, int& w) {
  int hw = PH * PW;
  n = p / hw;
  int r = p - n * hw;
  h = r / PW;
  w = r - h * PW;
}

// AMODE: how the activation operand is staged. 0 = tiled TMA (pure GEMM), 1 = cp.async gather by the producer
// warps, 2 = TMA im2col (the hardware walks the output pixels and applies the filter-tap offset).
enum { ACT_TILED = 0, ACT_GATHER = 1, ACT_IM2COL = 2 };

template <int KIND, int AMODE, int BN>
__global__ void __launch_bounds__(kThreads, (BN <= 128) ? 2 : 1)
    igemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const ConvParams p) {
  using C = Cfg<BN>;
  constexpr int kLag = C::kLag;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + C::kStages;
  uint64_t* tmem_full = bars + 2 * C::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * C::kStages + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // Tile coordinates. x = N tile (fastest, so CTAs that share an A tile are co-scheduled), y = M tile.
  const int n0 = blockIdx.x * BN;
  const int m0 = blockIdx.y * BM;
  int tap = 0, split = 0, kb_begin = 0, kb_end = p.Kblocks;
  if (KIND == WGRAD) {
    tap = blockIdx.z / p.splits;
    split = blockIdx.z - tap * p.splits;
    kb_begin = split * p.kb_per_split;
    kb_end = min(p.Kblocks, kb_begin + p.kb_per_split);
  }
  const int nkb = kb_end - kb_begin;

  // Which operand (if any) the gather warps fill: fprop/dgrad -> A, wgrad -> B.
  constexpr bool GATHER = (AMODE == ACT_GATHER);
  constexpr bool IM2COL = (AMODE == ACT_IM2COL);
  constexpr bool kGatherA = GATHER && (KIND != WGRAD);
  constexpr bool kGatherB = GATHER && (KIND == WGRAD);
  constexpr uint32_t kTmaBytes = (kGatherA ? 0 : C::kABytes) + (kGatherB ? 0 : C::kBBytes);

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(&full[s], 1 + (GATHER ? kGatherThreads : 0));
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 5) tmem_alloc<BN>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 4) {
    // ===================================================== TMA producer (one lane)
    if (lane == 0) {
      // im2col: base coordinates of this tile's first output pixel in the source tensor (fprop/dgrad)
      int iw0 = 0, ih0 = 0, in0 = 0;
      if (IM2COL && KIND != WGRAD) {
        int n, h, w;
        decode_pix(m0, p.rowH, p.rowW, n, h, w);
        in0 = n;
        ih0 = h * p.stride - p.im2col_pad;
        iw0 = w * p.stride - p.im2col_pad;
      }
      for (int i = 0; i < nkb; ++i) {
        const int kb = kb_begin + i;
        const int s = i % C::kStages;
        const uint32_t ph = (i / C::kStages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        uint8_t* sA = smem + s * C::kStageBytes;
        uint8_t* sB = sA + C::kABytes;
        mbar_arrive_expect_tx(&full[s], kTmaBytes);
        if (KIND == FPROP) {
          if (IM2COL) {
            const int tp = kb / p.cpb, cb = kb - tp * p.cpb;
            const int kh = tp / p.KW, kw = tp - kh * p.KW;
            tma_load_im2col(sA, &tmA, &full[s], cb * 64, iw0, ih0, in0, (uint16_t)kw, (uint16_t)kh);
          } else if (!kGatherA) {
            tma_load_2d(sA, &tmA, &full[s], kb * BK, m0);  // x2d [M][Cin], box {64,128}
          }
          tma_load_2d(sB, &tmB, &full[s], kb * BK, n0);                 // W2d [Cout][K], box {64,BN}
        } else if (KIND == DGRAD) {
          const int t = kb / p.cpb, cob = kb - t * p.cpb;
          if (IM2COL) {  // stride-1 dgrad == correlation of dy with the flipped filter, pad' = K-1-pad
            const int kh = t / p.KW, kw = t - kh * p.KW;
            tma_load_im2col(sA, &tmA, &full[s], cob * 64, iw0, ih0, in0, (uint16_t)(p.KW - 1 - kw),
                            (uint16_t)(p.KH - 1 - kh));
          } else if (!kGatherA) {
            tma_load_2d(sA, &tmA, &full[s], kb * BK, m0);  // dy2d [M][Cout], box {64,128}
          }
#pragma unroll
          for (int j = 0; j < BN / 64; ++j)  // W2d [Cout][taps*Cin], box {64 cin, 64 cout} -> MN-major B
            tma_load_2d(sB + j * 8192, &tmB, &full[s], t * p.tapC + n0 + 64 * j, cob * 64);
        } else {
#pragma unroll
          for (int j = 0; j < 2; ++j)  // dy2d [pixels][Cout], box {64 cout, 64 pixels} -> MN-major A
            tma_load_2d(sA + j * 8192, &tmA, &full[s], m0 + 64 * j, kb * BK);
          if (IM2COL) {  // x rows for this tap: 64 output pixels starting at kb*64, 64 channels per box
            int n, h, w;
            decode_pix(kb * BK, p.rowH, p.rowW, n, h, w);
            const int kh = tap / p.KW, kw = tap - kh * p.KW;
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_im2col(sB + j * 8192, &tmB, &full[s], n0 + 64 * j, w * p.stride - p.im2col_pad,
                              h * p.stride - p.im2col_pad, n, (uint16_t)kw, (uint16_t)kh);
          } else if (!kGatherB) {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)  // x2d [pixels][Cin]
              tma_load_2d(sB + j * 8192, &tmB, &full[s], n0 + 64 * j, kb * BK);
          }
        }
      }
    }
  } else if (warp == 5) {
    // ===================================================== MMA issuer (one lane)
    if (lane == 0) {
      constexpr bool kAmn = (KIND == WGRAD);
      constexpr bool kBmn = (KIND != FPROP);
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, kAmn, kBmn);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % C::kStages;
        const uint32_t ph = (i / C::kStages) & 1;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t aBase = smem_u32(smem + s * C::kStageBytes);
        const uint32_t bBase = aBase + C::kABytes;
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          // K-major: step 16 elements (32 B) inside the swizzle atom. MN-major: step 16 rows (2 KB).
          const uint64_t da = kAmn ? make_smem_desc_sw128(aBase + k * 2048, 8192, 1024)
                                   : make_smem_desc_sw128(aBase + k * 32, 16, 1024);
          const uint64_t db = kBmn ? make_smem_desc_sw128(bBase + k * 2048, 8192, 1024)
                                   : make_smem_desc_sw128(bBase + k * 32, 16, 1024);
          umma_bf16(tmem_base, da, db, idesc, (i | k) != 0);
        }
        umma_commit(&empty[s]);  // smem slot reusable once these MMAs retire
      }
      umma_commit(tmem_full);
    }
  } else {
    // ===================================================== gather producers (warps 0-3)
    if (GATHER) {
      const int t = threadIdx.x;  // 0..127
      const __nv_bfloat16* __restrict__ src = p.src;
      if (KIND != WGRAD) {
        // Row t of the A tile = output pixel m0+t for the whole tile.
        const int m = m0 + t;
        const bool row_ok = m < p.M;
        int n, h, w;
        decode_pix(row_ok ? m : 0, p.rowH, p.rowW, n, h, w);
        const uint32_t swz = (t & 7);
        int pending = 0;
        for (int i = 0; i < nkb; ++i) {
          const int kb = kb_begin + i;
          const int s = i % C::kStages;
          const uint32_t ph = (i / C::kStages) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          const int tp = kb / p.cpb, cb = kb - tp * p.cpb;
          const int kh = tp / p.KW, kw = tp - kh * p.KW;
          int sh, sw;
          bool ok = row_ok;
          if (KIND == FPROP) {
            sh = h * p.stride - p.pad + kh;
            sw = w * p.stride - p.pad + kw;
          } else {  // DGRAD: dy pixel that tap (kh,kw) maps onto dx pixel (h,w), if any
            const int th = h + p.pad - kh, tw = w + p.pad - kw;
            ok = ok && th >= 0 && tw >= 0;
            if (p.stride == 1) {
              sh = th; sw = tw;
            } else {
              sh = th / p.stride; sw = tw / p.stride;
              ok = ok && (sh * p.stride == th) && (sw * p.stride == tw);
            }
          }
          ok = ok && sh >= 0 && sw >= 0 && sh < p.srcH && sw < p.srcW;
          const __nv_bfloat16* g =
              ok ? src + ((size_t)((n * p.srcH + sh) * p.srcW + sw) * p.srcC + cb * 64) : src;
          const uint32_t dst = smem_u32(smem + s * C::kStageBytes) + t * 128;
          const uint32_t nbytes = ok ? 16u : 0u;
#pragma unroll
          for (int j = 0; j < 8; ++j) cp_async16(dst + ((j ^ swz) << 4), g + j * 8, nbytes);
          cp_async_commit();
          ++pending;
          if (pending > kLag) {
            cp_async_wait<kLag>();
            fence_proxy_async_smem();
            mbar_arrive(&full[(i - kLag) % C::kStages]);
            --pending;
          }
        }
        cp_async_wait<0>();
        fence_proxy_async_smem();
        for (int i = nkb - pending; i < nkb; ++i) mbar_arrive(&full[i % C::kStages]);
      } else {
        // WGRAD: B tile rows are K (64 pixels per k-block); BN/64 channel blocks of 64.
        constexpr int kItems = 64 * (BN / 64);
        constexpr int kPer = (kItems + kGatherThreads - 1) / kGatherThreads;
        const int kh = tap / p.KW, kw = tap - kh * p.KW;
        int pending = 0;
        for (int i = 0; i < nkb; ++i) {
          const int kb = kb_begin + i;
          const int s = i % C::kStages;
          const uint32_t ph = (i / C::kStages) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          const uint32_t sB = smem_u32(smem + s * C::kStageBytes + C::kABytes);
#pragma unroll
          for (int q = 0; q < kPer; ++q) {
            const int item = t + q * kGatherThreads;
            if (item < kItems) {
              const int r = item & 63, j = item >> 6;
              const int pix = kb * BK + r;
              bool ok = pix < p.Kpix;
              int n, h, w;
              decode_pix(ok ? pix : 0, p.rowH, p.rowW, n, h, w);
              const int sh = h * p.stride - p.pad + kh, sw = w * p.stride - p.pad + kw;
              ok = ok && sh >= 0 && sw >= 0 && sh < p.srcH && sw < p.srcW;
              const __nv_bfloat16* g =
                  ok ? src + ((size_t)((n * p.srcH + sh) * p.srcW + sw) * p.srcC + n0 + j * 64) : src;
              const uint32_t dst = sB + j * 8192 + r * 128;
              const uint32_t nbytes = ok ? 16u : 0u;
              const uint32_t swz = r & 7;
#pragma unroll
              for (int c = 0; c < 8; ++c) cp_async16(dst + ((c ^ swz) << 4), g + c * 8, nbytes);
            }
          }
          cp_async_commit();
          ++pending;
          if (pending > kLag) {
            cp_async_wait<kLag>();
            fence_proxy_async_smem();
            mbar_arrive(&full[(i - kLag) % C::kStages]);
            --pending;
          }
        }
        cp_async_wait<0>();
        fence_proxy_async_smem();
        for (int i = nkb - pending; i < nkb; ++i) mbar_arrive(&full[i % C::kStages]);
      }
    }

    // ===================================================== epilogue (warps 0-3): TMEM -> regs -> global
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int row = warp * 32 + lane;  // accumulator row == TMEM lane
    const uint32_t taddr = tmem_base + (uint32_t(warp * 32) << 16);
    if (KIND != WGRAD) {
      __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.out);
      const bool ok = (m0 + row) < p.M;
      __nv_bfloat16* orow = out + (size_t)(m0 + row) * p.out_ld + n0;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t v[32];
        tmem_ld32(taddr + c * 32, v);
        tmem_ld_wait();
        if (ok) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int col = n0 + c * 32 + q * 8;
            if (col + 8 > p.Ncols) break;
            if (p.bias) {
#pragma unroll
              for (int e = 0; e < 8; ++e)
                v[8 * q + e] = __float_as_uint(__uint_as_float(v[8 * q + e]) + __ldg(p.bias + col + e));
            }
            uint4 o;
            o.x = pack_bf16x2(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1]));
            o.y = pack_bf16x2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
            o.z = pack_bf16x2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
            o.w = pack_bf16x2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
            st_v4(orow + c * 32 + q * 8, o);
          }
        }
      }
    } else {
      float* out = reinterpret_cast<float*>(p.out) + (size_t)split * p.split_stride;
      const bool ok = (m0 + row) < p.M;
      float* orow = out + (size_t)(m0 + row) * p.out_ld + (size_t)tap * p.tapC + n0;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t v[32];
        tmem_ld32(taddr + c * 32, v);
        tmem_ld_wait();
        if (ok) {
#pragma unroll
          for (int q = 0; q < 8; ++q)
            st_v4(orow + c * 32 + q * 4, make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_free<BN>(tmem_base);
}

// dW[i] = (accumulate ? dW[i] : 0) + sum_s ws[s][i]   (fixed order -> deterministic)
__global__ void wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ dw, long long n, int splits,
                                    long long split_stride, int accumulate) {
  long long i = (blockIdx.x * (long long)blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  float4 acc = accumulate ? *reinterpret_cast<const float4*>(dw + i) : make_float4(0.f, 0.f, 0.f, 0.f);
  // eight partial slabs requested at a time, added in slab order (the sum stays bit-reproducible): with one load in
  // flight per thread the kernel was pure latency (up to 148 dependent round trips for the small weight tensors)
  int s = 0;
  for (; s + 8 <= splits; s += 8) {
    float4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = __ldcs(reinterpret_cast<const float4*>(ws + (long long)(s + u) * split_stride + i));
#pragma unroll
    for (int u = 0; u < 8; ++u) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
  }
  for (; s < splits; ++s) {
    const float4 v = __ldcs(reinterpret_cast<const float4*>(ws + (long long)s * split_stride + i));
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  }
  *reinterpret_cast<float4*>(dw + i) = acc;
}

template <int KIND, int AMODE, int BN>
static cudaError_t launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvParams& p, dim3 grid,
                          cudaStream_t st) {
  auto kern = igemm_kernel<KIND, AMODE, BN>;
  static bool configured = false;  // per instantiation
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::kSmem);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  kern<<<grid, kThreads, Cfg<BN>::kSmem, st>>>(tmA, tmB, p);
  return cudaGetLastError();
}

template <int KIND, int AMODE>
static cudaError_t launch_bn(int bn, const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvParams& p, dim3 grid,
                             cudaStream_t st) {
  switch (bn) {
    case 64: return launch<KIND, AMODE, 64>(tmA, tmB, p, grid, st);
    case 128: return launch<KIND, AMODE, 128>(tmA, tmB, p, grid, st);
    case 256: return launch<KIND, AMODE, 256>(tmA, tmB, p, grid, st);
  }
  return cudaErrorInvalidValue;
}

static int pick_bn(int ncols, int force) {
  if (force == 64 || force == 128 || force == 256) return (ncols % force == 0) ? force : 64;
  if (ncols % 128 == 0) return 128;
  return 64;
}

static int g_force_bn = 0;
static int g_dbg = 0;
static int g_mt2 = 1;     // v2 wgrad: 256 x 256 tile per CTA where Cout allows (b2_conv_set_wgrad_mt2)
static int g_wgrad_waves = 1;   // v2 wgrad split-K target: work items per SM (1: half the partial-sum traffic of 2)
static int g_wgrad_red = 1;  // v2 wgrad: split-K sums form in L2 (red.add.v4.f32 into dW) instead of slabs + a reduce kernel
static int g_panels = 1;  // v2: fetch MN-major operands as one multi-panel TMA box (b2_conv_set_panels)
static int g_act_mode = -1;  // -1 auto, else force ACT_GATHER / ACT_IM2COL for k>1 or strided convs

template <int KIND>
static cudaError_t launch_mode(int mode, int bn, const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvParams& p,
                               dim3 grid, cudaStream_t st) {
  switch (mode) {
    case ACT_TILED: return launch_bn<KIND, ACT_TILED>(bn, tmA, tmB, p, grid, st);
    case ACT_GATHER: return launch_bn<KIND, ACT_GATHER>(bn, tmA, tmB, p, grid, st);
    case ACT_IM2COL: return launch_bn<KIND, ACT_IM2COL>(bn, tmA, tmB, p, grid, st);
  }
  return cudaErrorInvalidValue;
}

// Staging mode for the activation operand of a (KH,KW,stride,pad) convolution.
static int pick_mode(int kind, int KH, int KW, int stride, int pad) {
  if (KH == 1 && KW == 1 && stride == 1 && pad == 0) return ACT_TILED;
  if (g_act_mode == ACT_GATHER) return ACT_GATHER;
  if (kind == DGRAD && stride != 1) return ACT_GATHER;  // strided dgrad has holes the im2col walk cannot skip
  return ACT_IM2COL;
}


// =====================================================================================================
// v2: persistent, TMEM double-buffered.  One CTA per SM loops over work items; while the four epilogue warps drain
// accumulator stage t&1 (TMEM -> registers -> bf16 -> global), the MMA warp is already accumulating the next item
// into the other stage and the TMA warp is several k-blocks ahead.  Activation staging is TMA only (tiled for
// pure GEMMs, im2col otherwise); strided dgrad is decomposed into output-pixel parity classes so that the im2col
// walk never multiplies structural zeros.
// =====================================================================================================
// warp 0 TMA (A), warp 1 MMA + TMEM alloc, warps 2-5 and 7-10 epilogue, warp 6 TMA (B)
constexpr int kThreads2 = 352;
constexpr int kEpiThreads2 = 256;
constexpr int kThreads2W = 224;  // wgrad (fp32 stores straight from registers) runs with the first epilogue group only
// fprop with fused BatchNorm statistics: 4 or 8 more warps (11..) read the bf16 output panels back from shared memory
template <int BN>
constexpr int kStatWarps2 = (BN == 64) ? 4 : 8;

// MT = 128-row accumulators per CTA. MT = 2 (wgrad only, BN = 256): one 256 x 256 tile in all 512 TMEM columns - no
// accumulator double buffering (a wgrad item is one long K loop, there is nothing to overlap), and a third less
// operand traffic per FLOP, which is what bounds wgrad (its operands come out of L2, not out of a shared weight tile).
template <int BN, int MT = 1>
struct Cfg2 {
  static constexpr int kStageBytes = MT * BM * 128 + BN * 128;
  static constexpr int kStages = (MT == 2) ? 3 : ((BN == 256) ? 4 : (BN == 128 ? 6 : 8));
  static constexpr int kEpiBytes = 2 * BM * 128;  // two 128-row x 64-column bf16 panels for the TMA-store epilogue
  static constexpr int kSmem = kStages * kStageBytes + kEpiBytes + 1024 + 256;
};

struct Item {
  int n0, m0, tap, split, kb_begin, nkb;
};
B2_DEVICE Item decode_item(const ConvParams& p, int item, int kind, int bn, int bm = BM) {
  Item it;
  const int nt = item % p.n_tiles;
  int r = item / p.n_tiles;
  const int mt = r % p.m_tiles;
  it.n0 = nt * bn;
  it.m0 = mt * bm;
  it.tap = 0; it.split = 0; it.kb_begin = 0; it.nkb = p.Kblocks;
  if (kind == WGRAD) {
    const int z = r / p.m_tiles;
    it.tap = z / p.splits;
    it.split = z - it.tap * p.splits;
    it.kb_begin = it.split * p.kb_per_split;
    it.nkb = min(p.Kblocks, it.kb_begin + p.kb_per_split) - it.kb_begin;
  }
  return it;
}

// FAN = 1 (dgrad only): the fan-in variant - the statistics warps finish the panel (skip add, ReLU mask) and a store
// warp writes it out; a separate instantiation so that the default kernels do not carry its code and registers.
template <int KIND, int AMODE, int BN, int MT = 1, int FAN = 0>
__global__ void __launch_bounds__(KIND == WGRAD ? kThreads2W : kThreads2 + 32 * kStatWarps2<BN> + (FAN ? 32 : 0), 1)
    igemm2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmC, const ConvParams p, int num_items) {
  using C = Cfg2<BN, MT>;
  static_assert(MT == 1 || (KIND == WGRAD && BN == 256), "two accumulators per CTA: wgrad, 256-wide tiles only");
  constexpr bool IM2COL = (AMODE == ACT_IM2COL);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_smem = smem + C::kStages * C::kStageBytes;  // 1024-aligned: stage sizes are multiples of 1 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_smem + C::kEpiBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + C::kStages;
  uint64_t* tfull = bars + 2 * C::kStages;       // [2] accumulator stage ready for the epilogue
  uint64_t* tempty = bars + 2 * C::kStages + 2;  // [2] accumulator stage drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * C::kStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr uint32_t kABytes = MT * BM * 128, kBBytes = BN * 128;

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(&full[s], 2);  // one arrive.expect_tx from each of the two TMA producer threads
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], KIND == WGRAD ? 4 : kEpiThreads2 / 32);  // one arrival per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) tmem_alloc<2 * BN>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 || warp == 6) {
    // ===================================================== TMA producers: warp 0 stages A, warp 6 stages B
    // (two issuing threads: the MN-major operands of dgrad/wgrad need one TMA per 64-column box)
    const bool doA = (warp == 0), doB = (warp == 6);
    if (lane == 0) {
      uint32_t g = 0;  // global k-block counter (smem ring position carries over from item to item)
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        const Item it = decode_item(p, item, KIND, BN, MT * BM);
        int iw0 = 0, ih0 = 0, in0 = 0;
        if (IM2COL && KIND != WGRAD) {
          int n, h, w;
          decode_pix(it.m0, p.rowH, p.rowW, n, h, w);
          in0 = n;
          ih0 = h * p.stride - p.im2col_pad_h;
          iw0 = w * p.stride - p.im2col_pad_w;
        }
        // wgrad over the flattened (tap, cin) axis: which tap / channel block each 64-column box of this tile is
        int box_ci[BN / 64], box_kh[BN / 64], box_kw[BN / 64], box_valid = BN / 64;
        if (KIND == WGRAD && IM2COL && p.flatN) {
          box_valid = min(BN / 64, (p.Ncols - it.n0 + 63) / 64);  // boxes past the last tap are not loaded
#pragma unroll
          for (int j = 0; j < BN / 64; ++j) {
            const int col = it.n0 + 64 * j;
            const int tp = col / p.tapC;
            box_ci[j] = col - tp * p.tapC;
            box_kh[j] = tp / p.KW;
            box_kw[j] = tp - box_kh[j] * p.KW;
          }
        }
        for (int i = 0; i < it.nkb; ++i, ++g) {
          const int kb = it.kb_begin + i;
          const int s = g % C::kStages;
          const uint32_t ph = (g / C::kStages) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          if ((p.dbg & 1) && g >= (uint32_t)C::kStages) {
            mbar_arrive(&full[s]);
            continue;
          }
          uint8_t* sA = smem + s * C::kStageBytes;
          uint8_t* sB = sA + kABytes;
          uint32_t txB = kBBytes;
          if (KIND == WGRAD && IM2COL && p.flatN) txB = box_valid * 8192;
          mbar_arrive_expect_tx(&full[s], doA ? kABytes : txB);
          if (KIND == FPROP) {
            if (doA) {
              if (IM2COL) {
                const int tp = kb / p.cpb, cb = kb - tp * p.cpb;
                const int kh = tp / p.KW, kw = tp - kh * p.KW;
                tma_load_im2col(sA, &tmA, &full[s], cb * 64, iw0, ih0, in0, (uint16_t)kw, (uint16_t)kh);
              } else {
                tma_load_2d(sA, &tmA, &full[s], kb * BK, it.m0);
              }
            } else {
              tma_load_2d(sB, &tmB, &full[s], kb * BK, it.n0);
            }
          } else if (KIND == DGRAD) {
            const int tp = kb / p.cpb, cob = kb - tp * p.cpb;
            int tapidx = 0;
            if (IM2COL) {
              const int th = tp / p.ntw, tw = tp - th * p.ntw;
              tapidx = (p.kh0 + p.kstep * th) * p.KW + (p.kw0 + p.kstep * tw);
              if (doA)
                tma_load_im2col(sA, &tmA, &full[s], cob * 64, iw0, ih0, in0, (uint16_t)(p.ntw - 1 - tw),
                                (uint16_t)(p.nth - 1 - th));
            } else if (doA) {
              tma_load_2d(sA, &tmA, &full[s], kb * BK, it.m0);
            }
            if (doB) {
              if (p.panels & 2) {
                tma_load_3d(sB, &tmB, &full[s], 0, cob * 64, (tapidx * p.tapC + it.n0) >> 6);
              } else {
#pragma unroll
                for (int j = 0; j < BN / 64; ++j)
                  tma_load_2d(sB + j * 8192, &tmB, &full[s], tapidx * p.tapC + it.n0 + 64 * j, cob * 64);
              }
            }
          } else if (doA) {
            if (p.panels & 1) {
              tma_load_3d(sA, &tmA, &full[s], 0, kb * BK, it.m0 >> 6);
            } else {
#pragma unroll
              for (int j = 0; j < 2 * MT; ++j) tma_load_2d(sA + j * 8192, &tmA, &full[s], it.m0 + 64 * j, kb * BK);
            }
          } else {
            if (IM2COL) {
              int n, h, w;
              decode_pix(kb * BK, p.rowH, p.rowW, n, h, w);
              const int bw = w * p.stride - p.im2col_pad_w, bh = h * p.stride - p.im2col_pad_h;
              if (p.flatN) {
#pragma unroll
                for (int j = 0; j < BN / 64; ++j)
                  if (j < box_valid)
                    tma_load_im2col(sB + j * 8192, &tmB, &full[s], box_ci[j], bw, bh, n, (uint16_t)box_kw[j],
                                    (uint16_t)box_kh[j]);
              } else {
                const int kh = it.tap / p.KW, kw = it.tap - kh * p.KW;
#pragma unroll
                for (int j = 0; j < BN / 64; ++j)
                  tma_load_im2col(sB + j * 8192, &tmB, &full[s], it.n0 + 64 * j, bw, bh, n, (uint16_t)kw, (uint16_t)kh);
              }
            } else if (p.panels & 2) {
              tma_load_3d(sB, &tmB, &full[s], 0, kb * BK, it.n0 >> 6);
            } else {
#pragma unroll
              for (int j = 0; j < BN / 64; ++j)
                tma_load_2d(sB + j * 8192, &tmB, &full[s], it.n0 + 64 * j, kb * BK);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    if (lane == 0) {
      constexpr bool kAmn = (KIND == WGRAD);
      constexpr bool kBmn = (KIND != FPROP);
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, kAmn, kBmn);
      uint32_t g = 0, t = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++t) {
        const Item it = decode_item(p, item, KIND, BN, MT * BM);
        const uint32_t acc = (MT == 2) ? 0u : (t & 1);
        mbar_wait(&tempty[acc], (((MT == 2) ? t : (t >> 1)) & 1) ^ 1);  // epilogue has drained the accumulator(s)
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int i = 0; i < it.nkb; ++i, ++g) {
          const int s = g % C::kStages;
          const uint32_t ph = (g / C::kStages) & 1;
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint32_t aBase = smem_u32(smem + s * C::kStageBytes);
          const uint32_t bBase = aBase + kABytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = kAmn ? make_smem_desc_sw128(aBase + k * 2048, 8192, 1024)
                                     : make_smem_desc_sw128(aBase + k * 32, 16, 1024);
            const uint64_t db = kBmn ? make_smem_desc_sw128(bBase + k * 2048, 8192, 1024)
                                     : make_smem_desc_sw128(bBase + k * 32, 16, 1024);
            if (!(p.dbg & 2)) umma_bf16(d_tmem, da, db, idesc, (i | k) != 0);
            if (MT == 2) {   // second 128 rows of the 256-row tile: panels 2,3 of the A stage, columns BN.. of TMEM
              const uint64_t da2 = make_smem_desc_sw128(aBase + 16384 + k * 2048, 8192, 1024);
              if (!(p.dbg & 2)) umma_bf16(d_tmem + BN, da2, db, idesc, (i | k) != 0);
            }
          }
          umma_commit(&empty[s]);
        }
        umma_commit(&tfull[acc]);
      }
    }
  } else if (warp >= 11) {
    // ===================================================== statistics warps 11.. (fprop with p.stats only)
    // BatchNorm statistics of exactly the bf16 values being stored, read back from the epilogue's shared-memory
    // panels while the epilogue warps convert the next panel. A warp pair owns one 64-column panel index (and one
    // of R row groups of it): each thread sums one 16-byte chunk (8 channels) over every 8th row; lanes g = 0..7
    // read rows = g (mod 8), eight different swizzle phases, so the 128-bit loads are conflict free (rows past M
    // are zero: their A rows were zero-filled). The sums stay in registers until the CTA moves to another n-tile;
    // a transposing butterfly over the 8 lanes then leaves lane g with channel chunk*8+g, which goes to
    // stats[R*cta + row group][2][Ncols].
    if (KIND == DGRAD && FAN) {
      // ---- fan-in mode: ALL statistics warps work on every panel (4 or 8 rows per thread). They add the skip-path
      // gradient, apply the ReLU mask of the tensor being differentiated, write the finished bf16 panel back, one
      // warp issues the TMA store (and alone hands the buffer back once the store has read it), and the
      // BatchNorm-backward sums sum(g), sum(g*y) are butterfly-reduced per panel. The companion tensors of the NEXT
      // panel are loaded into registers before waiting for it, so their latency is hidden behind the pipeline.
      constexpr int P = BN / 64, NSW = kStatWarps2<BN>, RQ = NSW / 2, NROW = BM / RQ / 8, T = NSW * 32;
      const int sw = warp - 11, g = lane & 7, q = sw >> 1;
      const int chunk = (sw & 1) * 4 + (lane >> 3);
      if (sw == NSW) {
        // ---- store warp: waits until a panel is final, issues its TMA store and hands the buffer back to the
        // epilogue as soon as the store has read it; the statistics warps never block on a store
        asm volatile("bar.arrive %0, %1;" ::"r"(P == 1 ? 6 : 6 + P - 2), "n"(kEpiThreads2 + 32) : "memory");
        asm volatile("bar.arrive %0, %1;" ::"r"(P == 1 ? 7 : 6 + P - 1), "n"(kEpiThreads2 + 32) : "memory");
        uint32_t t = 0;
        for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++t) {
          const Item it = decode_item(p, item, KIND, BN, MT * BM);
#pragma unroll 1
          for (int pn = 0; pn < P; ++pn) {
            const uint32_t bsel = (P == 1) ? (t & 1) : (pn & 1);
            const uint32_t slot = (P == 1) ? bsel : pn;
            asm volatile("bar.sync %0, %1;" ::"r"(11 + (int)bsel), "n"(T + 32) : "memory");   // the panel is final
            if (lane == 0) {
              tma_store_2d(epi_smem + bsel * (BM * 128), &tmC, it.n0 + pn * 64, it.m0);
              asm volatile("cp.async.bulk.commit_group;" ::: "memory");
              asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
            __syncwarp();
            asm volatile("bar.arrive %0, %1;" ::"r"(6 + slot), "n"(kEpiThreads2 + 32) : "memory");   // buffer free
          }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // stores have landed
      } else {
      const bool sums = p.stats != nullptr && p.bwd_y != nullptr;   // also produce the BatchNorm-backward sums
      if (sums) {
        float4* z = reinterpret_cast<float4*>(p.stats + (size_t)blockIdx.x * RQ * 2 * p.Ncols);
        for (int i = threadIdx.x - kThreads2; i < RQ * 2 * p.Ncols / 4; i += T) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      asm volatile("bar.sync 10, %0;" ::"n"(T) : "memory");
      const bool rmw = (gridDim.x % p.n_tiles) != 0;
      uint4 sk[NROW], yv[NROW];
      uint32_t mk[NROW];
      auto load_companions = [&](int item_, int pn_) {
        const Item i2 = decode_item(p, item_, KIND, BN, MT * BM);
        const int col = i2.n0 + pn_ * 64 + chunk * 8;
        const int row0 = i2.m0 + q * (BM / RQ) + g;
#pragma unroll
        for (int i = 0; i < NROW; ++i) {
          const int r = row0 + 8 * i;
          const bool ok = r < p.M;
          const size_t off = (size_t)r * p.Ncols + col;
          sk[i] = ok ? ld_nc_v4(p.add + off) : make_uint4(0u, 0u, 0u, 0u);
          if (sums) yv[i] = ok ? ld_nc_v4(p.bwd_y + off) : make_uint4(0u, 0u, 0u, 0u);
          mk[i] = !ok ? 0u : (p.amask ? (uint32_t)__ldg(p.amask + (off >> 3)) : 0xffu);
        }
      };
      float st_a[P], st_b[P];
#pragma unroll
      for (int j = 0; j < P; ++j) st_a[j] = st_b[j] = 0.f;
      int st_n0 = -1;
      auto flush = [&]() {
        if (st_n0 < 0) return;
        float* dst = p.stats + (size_t)(blockIdx.x * RQ + q) * 2 * p.Ncols;
#pragma unroll
        for (int j = 0; j < P; ++j) {
          const int gc = st_n0 + j * 64 + chunk * 8 + g;
          if (gc < p.Ncols) {
            const float s0 = st_a[j], s1 = st_b[j] - __ldg(p.bwd_mean + gc) * st_a[j];   // sum g*(y - mean)
            if (rmw) {
              dst[gc] += s0;
              dst[p.Ncols + gc] += s1;
            } else {
              dst[gc] = s0;
              dst[p.Ncols + gc] = s1;
            }
          }
          st_a[j] = st_b[j] = 0.f;
        }
      };
      int item = blockIdx.x;
      if (item < num_items) load_companions(item, 0);
      uint32_t t = 0;
      for (; item < num_items; item += gridDim.x, ++t) {
        const Item it = decode_item(p, item, KIND, BN, MT * BM);
        if (sums && it.n0 != st_n0) {
          flush();
          st_n0 = it.n0;
        }
        if (item + (int)gridDim.x < num_items) {
          // pull the companion lines of the NEXT tile into L2 now (a whole tile ahead): the register loads issued
          // one panel ahead then only pay the L2 latency. 8 threads share a row: 4 take the skip-gradient lines, 4
          // the activation lines (one 128-byte line per 64 columns).
          const Item nx = decode_item(p, item + (int)gridDim.x, KIND, BN, MT * BM);
          if ((chunk & 3) < P && (chunk < 4 || sums)) {
            const __nv_bfloat16* base = (chunk < 4 ? p.add : p.bwd_y) + nx.n0 + (chunk & 3) * 64;
#pragma unroll
            for (int i = 0; i < NROW; ++i) {
              const int r = nx.m0 + q * (BM / RQ) + g + 8 * i;
              if (r < p.M) {
                asm volatile("prefetch.global.L2 [%0];" ::"l"(base + (size_t)r * p.Ncols));
                if (chunk == 0 && p.amask)
                  asm volatile("prefetch.global.L2 [%0];" ::"l"(p.amask + (((size_t)r * p.Ncols + nx.n0) >> 3)));
              }
            }
          }
        }
#pragma unroll
        for (int pn = 0; pn < P; ++pn) {
          const uint32_t bsel = (P == 1) ? (t & 1) : (pn & 1);
          const uint32_t slot = (P == 1) ? bsel : pn;
          asm volatile("bar.sync %0, %1;" ::"r"(2 + slot), "n"(kEpiThreads2 + T) : "memory");   // panel written
          uint8_t* pbase = epi_smem + bsel * (BM * 128) + (q * (BM / RQ) + g) * 128 + ((chunk ^ g) << 4);
          float a[8], b[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) a[e] = b[e] = 0.f;
#pragma unroll
          for (int i = 0; i < NROW; ++i) {
            uint4* cell = reinterpret_cast<uint4*>(pbase + i * 1024);
            const uint4 w4 = *cell;
            const uint32_t ww[4] = {w4.x, w4.y, w4.z, w4.w};
            const uint32_t ss[4] = {sk[i].x, sk[i].y, sk[i].z, sk[i].w};
            const uint32_t yy[4] = {yv[i].x, yv[i].y, yv[i].z, yv[i].w};
            uint32_t oo[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              // packed bf16x2 add (one rounding, as fp32 add + round), then the two mask bits pick bytes or zero
              __nv_bfloat162 wv = *reinterpret_cast<const __nv_bfloat162*>(&ww[e]);
              __nv_bfloat162 sv = *reinterpret_cast<const __nv_bfloat162*>(&ss[e]);
              const __nv_bfloat162 sum2 = __hadd2(wv, sv);
              const uint32_t sel = (uint32_t)(0x3210324444104444ull >> (16 * ((mk[i] >> (2 * e)) & 3u))) & 0xffffu;
              oo[e] = __byte_perm(*reinterpret_cast<const uint32_t*>(&sum2), 0u, sel);
              if (sums) {
                const float glo = __uint_as_float(oo[e] << 16), ghi = __uint_as_float(oo[e] & 0xffff0000u);
                a[2 * e] += glo;
                b[2 * e] = fmaf(glo, __uint_as_float(yy[e] << 16), b[2 * e]);
                a[2 * e + 1] += ghi;
                b[2 * e + 1] = fmaf(ghi, __uint_as_float(yy[e] & 0xffff0000u), b[2 * e + 1]);
              }
            }
            *cell = make_uint4(oo[0], oo[1], oo[2], oo[3]);
          }
          fence_proxy_async_smem();
          // the companion registers are free: fetch the next panel's while this one is stored
          if (pn + 1 < P) {
            load_companions(item, pn + 1);
          } else if (item + (int)gridDim.x < num_items) {
            load_companions(item + (int)gridDim.x, 0);
          }
          // this thread's part of the panel is final: tell the store warp (one barrier per panel buffer)
          asm volatile("bar.arrive %0, %1;" ::"r"(11 + (int)bsel), "n"(T + 32) : "memory");
          (void)slot;
          if (sums) {
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
            st_a[pn] += a[0];
            st_b[pn] += b[0];
          }
        }
      }
      if (sums) flush();
      }
    } else if ((KIND == FPROP || KIND == DGRAD) && p.stats) {
      constexpr int P = BN / 64, NSW = kStatWarps2<BN>, R = NSW / (2 * P), NLD = (BM / R) / 8;
      constexpr int CNT = kEpiThreads2 + 64 * R;
      const int sw = warp - 11, g = lane & 7;
      const int pair = sw >> 1, pn = pair % P, rg = pair / P;
      const int chunk = (sw & 1) * 4 + (lane >> 3);
      {
        float4* z = reinterpret_cast<float4*>(p.stats + (size_t)blockIdx.x * R * 2 * p.Ncols);
        for (int i = threadIdx.x - kThreads2; i < R * 2 * p.Ncols / 4; i += NSW * 32)
          z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      asm volatile("bar.sync 10, %0;" ::"n"(NSW * 32) : "memory");
      // the panel buffers start out free: one arrival on each "free" barrier the epilogue waits on first
      if (P == 1) {
        asm volatile("bar.arrive 6, %0;" ::"n"(CNT) : "memory");
        asm volatile("bar.arrive 7, %0;" ::"n"(CNT) : "memory");
      } else if (pn >= P - 2) {
        asm volatile("bar.arrive %0, %1;" ::"r"(6 + pn), "n"(CNT) : "memory");
      }
      const bool rmw = (gridDim.x % p.n_tiles) != 0;   // this CTA comes back to an n-tile it has already flushed
      float a[8], b[8], mu[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) a[e] = b[e] = mu[e] = 0.f;
      int st_n0 = -1;
      auto flush = [&]() {
        if (st_n0 < 0) return;
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
        const int gc = st_n0 + pn * 64 + chunk * 8 + g;
        if (gc < p.Ncols) {
          float* dst = p.stats + (size_t)(blockIdx.x * R + rg) * 2 * p.Ncols + gc;
          if (rmw) {
            dst[0] += a[0];
            dst[p.Ncols] += b[0];
          } else {
            dst[0] = a[0];
            dst[p.Ncols] = b[0];
          }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) a[e] = b[e] = 0.f;
      };
      uint32_t t = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++t) {
        const Item it = decode_item(p, item, KIND, BN, MT * BM);
        if (it.n0 != st_n0) {
          flush();
          st_n0 = it.n0;
          if (KIND == DGRAD) {
#pragma unroll
            for (int e = 0; e < 8; ++e) mu[e] = __ldg(p.bwd_mean + it.n0 + pn * 64 + chunk * 8 + e);
          }
        }
        const uint32_t bsel = (P == 1) ? (t & 1) : (pn & 1);
        const uint32_t slot = (P == 1) ? bsel : pn;
        const uint32_t src =
            smem_u32(epi_smem) + bsel * (BM * 128) + (rg * (BM / R) + g) * 128 + ((chunk ^ g) << 4);
        if (KIND == DGRAD) {
          // BatchNorm-backward sums: CH rows at a time; the companion loads of the first CH rows are issued before
          // waiting for the panel, the panel buffer is handed back after its last shared-memory read.
          constexpr int CH = (BN >= 128) ? 4 : 8;   // 4 keeps the wide variants (96-register cap) out of local memory
          const int col = it.n0 + pn * 64 + chunk * 8;
          const int row0 = it.m0 + rg * (BM / R) + g;
#pragma unroll
          for (int c8 = 0; c8 < NLD / CH; ++c8) {
            uint4 yv[CH];
            uint32_t mk[CH];
#pragma unroll
            for (int i = 0; i < CH; ++i) {
              const int r = row0 + 8 * (c8 * CH + i);
              const bool ok = r < p.M;
              yv[i] = ok ? ld_nc_v4(p.bwd_y + (size_t)r * p.Ncols + col) : make_uint4(0u, 0u, 0u, 0u);
              mk[i] = !ok ? 0u : (p.bwd_mask ? (uint32_t)__ldg(p.bwd_mask + (size_t)r * (p.Ncols >> 3) + (col >> 3))
                                             : 0xffu);
            }
            if (c8 == 0) asm volatile("bar.sync %0, %1;" ::"r"(2 + slot), "n"(CNT) : "memory");   // panel written
            uint4 w[CH];
#pragma unroll
            for (int i = 0; i < CH; ++i) w[i] = ld_shared_v4(src + (c8 * CH + i) * 1024);
            if (c8 == NLD / CH - 1) {
              __threadfence_block();
              asm volatile("bar.arrive %0, %1;" ::"r"(6 + slot), "n"(CNT) : "memory");   // panel buffer free again
            }
#pragma unroll
            for (int i = 0; i < CH; ++i) {
              const uint32_t ww[4] = {w[i].x, w[i].y, w[i].z, w[i].w};
              const uint32_t yy[4] = {yv[i].x, yv[i].y, yv[i].z, yv[i].w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float glo = ((mk[i] >> (2 * e)) & 1u) ? __uint_as_float(ww[e] << 16) : 0.f;
                const float ghi = ((mk[i] >> (2 * e + 1)) & 1u) ? __uint_as_float(ww[e] & 0xffff0000u) : 0.f;
                a[2 * e] += glo;
                b[2 * e] = fmaf(glo, __uint_as_float(yy[e] << 16) - mu[2 * e], b[2 * e]);
                a[2 * e + 1] += ghi;
                b[2 * e + 1] = fmaf(ghi, __uint_as_float(yy[e] & 0xffff0000u) - mu[2 * e + 1], b[2 * e + 1]);
              }
            }
          }
          continue;
        }
        asm volatile("bar.sync %0, %1;" ::"r"(2 + slot), "n"(CNT) : "memory");   // panel written
        uint4 w[NLD];
#pragma unroll
        for (int i = 0; i < NLD; ++i) w[i] = ld_shared_v4(src + i * 1024);
        __threadfence_block();                                                    // reads done before the hand-back
        asm volatile("bar.arrive %0, %1;" ::"r"(6 + slot), "n"(CNT) : "memory");   // panel buffer free again
#pragma unroll
        for (int i = 0; i < NLD; ++i) {
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
      flush();
    }
  } else {
    // ===================================================== epilogue warps 2..5 and 7..10 (TMEM lane quadrant =
    // warp % 4). The second group doubles the drain rate: for bf16 outputs group eg converts columns
    // [32*eg, 32*eg+32) of every 64-column panel (wgrad is launched with the first group only).
    const int eg = warp >= 7 ? 1 : 0;
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const bool store_thread = (threadIdx.x == 64);  // first epilogue thread issues the TMA stores
    const bool tma_epi = (KIND != WGRAD) && p.cls == 1;
    const bool with_stats = (KIND != WGRAD) && (p.stats != nullptr || (KIND == DGRAD && FAN != 0));
    constexpr int P = BN / 64;
    constexpr int CNT = kEpiThreads2 + 64 * (kStatWarps2<BN> / (2 * P));
    constexpr bool fan = (KIND == DGRAD) && FAN != 0;
    const int free_cnt = fan ? kEpiThreads2 + 32 : CNT;   // fan-in mode: only the storing warp hands a buffer back
    const int full_cnt = fan ? kEpiThreads2 + 32 * kStatWarps2<BN> : CNT;   // ... and every statistics warp takes it
    uint32_t t = 0, panel_seq = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++t) {
      const Item it = decode_item(p, item, KIND, BN, MT * BM);
      const uint32_t acc = (MT == 2) ? 0u : (t & 1);
      mbar_wait(&tfull[acc], ((MT == 2) ? t : (t >> 1)) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + (uint32_t(quad * 32) << 16);
      const int m = it.m0 + row;
      const bool ok = m < p.M;
      if (KIND != WGRAD && tma_epi) {
        // TMEM -> registers -> bf16 -> 128B-swizzled shared panel -> TMA store (full-line writes; the store clips
        // rows >= M and columns >= Ncols). Panels are double buffered.
        const uint32_t swz = (row & 7) << 4;
#pragma unroll 1
        for (int pn = 0; pn < P; ++pn, ++panel_seq) {
          if (it.n0 + pn * 64 >= p.Ncols) break;
          const uint32_t bsel = panel_seq & 1;
          uint8_t* buf = epi_smem + bsel * (BM * 128);
          uint32_t v[32];
          tmem_ld32(taddr + pn * 64 + eg * 32, v);
          tmem_ld_wait();
          if (p.bias) {
            const int col = it.n0 + pn * 64 + eg * 32;
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (col + e < p.Ncols) v[e] = __float_as_uint(__uint_as_float(v[e]) + __ldg(p.bias + col + e));
          }
          uint4 o[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            o[q].x = pack_bf16x2(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1]));
            o[q].y = pack_bf16x2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
            o[q].z = pack_bf16x2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
            o[q].w = pack_bf16x2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
          }
          // the panel buffer is needed only now: its previous TMA store has finished reading it ...
          if (!fan) {
            if (store_thread) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads2) : "memory");
          }
          if (KIND != WGRAD && with_stats)   // ... and so have the statistics warps (fan-in mode: they own the store)
            asm volatile("bar.sync %0, %1;" ::"r"(6 + (P == 1 ? bsel : (pn + P - 2) % P)), "r"(free_cnt) : "memory");
          uint8_t* prow = buf + row * 128;
#pragma unroll
          for (int q = 0; q < 4; ++q)  // 16-byte chunk eg*4+q of the 128-byte panel row, XOR-swizzled by the row phase
            *reinterpret_cast<uint4*>(prow + ((uint32_t((eg * 4 + q) << 4)) ^ swz)) = o[q];
          fence_proxy_async_smem();
          if (with_stats) {                                  // panel complete: hand it to the statistics warps
            __threadfence_block();
            asm volatile("bar.arrive %0, %1;" ::"r"(2 + (P == 1 ? bsel : pn)), "r"(full_cnt) : "memory");
          }
          if (!fan) {
            asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads2) : "memory");
            if (store_thread) {
              tma_store_2d(buf, &tmC, it.n0 + pn * 64, it.m0);
              asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
          }
        }
      } else if (KIND != WGRAD) {
        // Class-strided dgrad: GEMM row m is pixel (n, i, j) of the parity class grid and lands on pixel
        // (cls*i + ph, cls*j + pw) of dx, so the tile is not a TMA box. Each warp transposes 32 rows x 64 columns
        // (128 bytes per row, 16-byte pieces XOR-swizzled by the row) through its own 4 KB of the panel buffers and
        // stores 4 rows x 128 contiguous bytes per instruction; straight from the registers an instruction touches 32
        // rows with 16 bytes each, which runs at half the store rate (tools/red_probe.cu) and bounds these kernels.
        int pix = -1;
        if (ok) {
          int n, i, j;
          decode_pix(m, p.rowH, p.rowW, n, i, j);
          pix = (n * p.outH + (i * p.cls + p.ph)) * p.outW + (j * p.cls + p.pw);
        }
        uint8_t* stage = epi_smem + ((eg << 2) | quad) * 4096;
        const int rsub = lane >> 3, piece = lane & 7;
        __nv_bfloat16* obase = reinterpret_cast<__nv_bfloat16*>(p.out) + it.n0 + piece * 8;
#pragma unroll 1
        for (int cc = eg * 2; cc < BN / 32; cc += 4) {
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            uint32_t v[32];
            tmem_ld32(taddr + (cc + h) * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              uint4 o;
              o.x = pack_bf16x2(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1]));
              o.y = pack_bf16x2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
              o.z = pack_bf16x2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
              o.w = pack_bf16x2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
              *reinterpret_cast<uint4*>(stage + lane * 128 + (((h * 4 + q) ^ (lane & 7)) << 4)) = o;
            }
          }
          __syncwarp();
          const bool col_ok = it.n0 + cc * 32 + piece * 8 + 8 <= p.Ncols;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int r = 4 * j + rsub;
            const uint4 o = *reinterpret_cast<const uint4*>(stage + r * 128 + ((piece ^ (r & 7)) << 4));
            const int pj = __shfl_sync(0xffffffffu, pix, r);
            if (pj >= 0 && col_ok) st_v4(obase + (size_t)pj * p.out_ld + cc * 32, o);
          }
          __syncwarp();
        }
      } else {
        // fp32 tile -> dW (or a split-K slab). Straight from the registers every warp instruction would touch 32
        // rows with 16 bytes each (measured: half the store rate, tools/red_probe.cu), so each 32 x 32 block is
        // transposed through a per-warp staging buffer (row pitch 144 B: conflict-free both ways) and leaves as
        // 4 rows x 128 contiguous bytes per instruction.
        float* out = reinterpret_cast<float*>(p.out) + (size_t)it.split * p.split_stride;
        const bool red = p.red_out != 0;
        uint8_t* stage = epi_smem + quad * (32 * 144);
        const int rsub = lane >> 3, piece = lane & 7;
#pragma unroll 1
        for (int mt = 0; mt < MT; ++mt) {
          const int mm0 = it.m0 + mt * BM + quad * 32;   // first of this warp's 32 rows
          float* obase = out + (size_t)(mm0 + rsub) * p.out_ld + (size_t)it.tap * p.tapC + it.n0 + piece * 4;
#pragma unroll 1
          for (int c = 0; c < BN / 32; ++c) {
            uint32_t v[32];
            tmem_ld32(taddr + mt * BN + c * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int q = 0; q < 8; ++q)
              *reinterpret_cast<uint4*>(stage + lane * 144 + q * 16) =
                  make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
            __syncwarp();
            const bool col_ok = !p.flatN || it.n0 + c * 32 + piece * 4 + 4 <= p.Ncols;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint4 o = *reinterpret_cast<const uint4*>(stage + (4 * j + rsub) * 144 + piece * 16);
              if (col_ok && mm0 + 4 * j + rsub < p.M) {
                float* dst = obase + (size_t)(4 * j) * p.out_ld + c * 32;
                if (red) red_add_v4(dst, o.x, o.y, o.z, o.w);
                else st_v4(dst, o);
              }
            }
            __syncwarp();
          }
        }
      }
      // release the accumulator stage: all of this warp's TMEM reads have completed (wait::ld above)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);
    }
  }

  if (threadIdx.x == 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // stores have landed
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_free<2 * BN>(tmem_base);
}

// =====================================================================================================
// v3: CTA pairs (cta_group::2).  Two CTAs of a 2-cluster (the two SMs of a TPC) compute one 256 x BN tile: each
// stages its own 128 rows of A and BN/2 columns of B, the leader issues tcgen05.mma.cta_group::2 which reads
// both CTAs' shared memory, and each CTA ends up with its 128 accumulator rows in its own TMEM.  Per SM the
// tensor core does the same work as in v2 but reads one third less shared memory per MMA (4 KB + 4 KB instead of
// 4 KB + 8 KB for BN = 256), which is what bounds v2 on the K-heavy layers.  Same persistent structure otherwise.
// =====================================================================================================
constexpr int kThreads3 = 192;  // warp 0 TMA, warp 1 MMA (leader CTA only) + TMEM alloc, warps 2-5 epilogue

template <int BN>
struct Cfg3 {
  static constexpr int kABytes = BM * 128;
  static constexpr int kBBytes = (BN / 2) * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN == 256) ? 6 : 8;
  static constexpr int kEpiBytes = 2 * BM * 128;
  static constexpr int kSmem = kStages * kStageBytes + kEpiBytes + 1024 + 256;
};

template <int KIND, int AMODE, int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads3, 1)
    igemm3_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmC, const ConvParams p, int num_items) {
  using C = Cfg3<BN>;
  constexpr bool IM2COL = (AMODE == ACT_IM2COL);
  constexpr int BNH = BN / 2;  // B columns staged by each CTA
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_smem = smem + C::kStages * C::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_smem + C::kEpiBytes);
  uint64_t* full = bars;                          // used in the leader CTA (both CTAs' loads count there)
  uint64_t* empty = bars + C::kStages;            // per CTA, signalled by the multicast commit
  uint64_t* tfull = bars + 2 * C::kStages;        // [2] per CTA
  uint64_t* tempty = bars + 2 * C::kStages + 2;   // [2] used in the leader CTA: 8 epilogue warps arrive
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * C::kStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();  // 0 = leader
  const int pair = blockIdx.x >> 1;
  const int npairs = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(&full[s], 2);   // one arrive.expect_tx per CTA of the pair
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 8);
    }
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  cluster_sync_all();
  if (warp == 1) tmem_alloc_pair<2 * BN>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================================================== TMA producer (both CTAs)
    if (lane == 0) {
      uint32_t g = 0;
      for (int item = pair; item < num_items; item += npairs) {
        Item it = decode_item(p, item, KIND, BN);
        // this CTA's share of the 256-row tile
        const int m0 = (KIND == WGRAD ? it.m0 * 2 : it.m0 * 2) + 128 * (int)rank;  // items index 256-row tiles
        const int nB = it.n0 + BNH * (int)rank;                                       // this CTA's B columns
        int iw0 = 0, ih0 = 0, in0 = 0;
        if (IM2COL && KIND != WGRAD) {
          int n, h, w;
          decode_pix(m0, p.rowH, p.rowW, n, h, w);
          in0 = n;
          ih0 = h * p.stride - p.im2col_pad_h;
          iw0 = w * p.stride - p.im2col_pad_w;
        }
        for (int i = 0; i < it.nkb; ++i, ++g) {
          const int kb = it.kb_begin + i;
          const int s = g % C::kStages;
          const uint32_t ph = (g / C::kStages) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* sA = smem + s * C::kStageBytes;
          uint8_t* sB = sA + C::kABytes;
          const uint32_t fb = mapa_cluster(smem_u32(&full[s]), 0);  // the leader's full barrier
          mbar_arrive_expect_tx_cluster(fb, C::kABytes + C::kBBytes);
          if (KIND == FPROP) {
            if (IM2COL) {
              const int tp = kb / p.cpb, cb = kb - tp * p.cpb;
              const int kh = tp / p.KW, kw = tp - kh * p.KW;
              tma_load_im2col_pair(sA, &tmA, fb, cb * 64, iw0, ih0, in0, (uint16_t)kw, (uint16_t)kh);
            } else {
              tma_load_2d_pair(sA, &tmA, fb, kb * BK, m0);
            }
            tma_load_2d_pair(sB, &tmB, fb, kb * BK, nB);   // W2d rows [nB, nB + BN/2): box {64, BN/2}
          } else if (KIND == DGRAD) {
            const int tp = kb / p.cpb, cob = kb - tp * p.cpb;
            int tapidx = 0;
            if (IM2COL) {
              const int th = tp / p.ntw, tw = tp - th * p.ntw;
              tapidx = (p.kh0 + p.kstep * th) * p.KW + (p.kw0 + p.kstep * tw);
              tma_load_im2col_pair(sA, &tmA, fb, cob * 64, iw0, ih0, in0, (uint16_t)(p.ntw - 1 - tw),
                                   (uint16_t)(p.nth - 1 - th));
            } else {
              tma_load_2d_pair(sA, &tmA, fb, kb * BK, m0);
            }
#pragma unroll
            for (int j = 0; j < BNH / 64; ++j)
              tma_load_2d_pair(sB + j * 8192, &tmB, fb, tapidx * p.tapC + nB + 64 * j, cob * 64);
          } else {
#pragma unroll
            for (int j = 0; j < 2; ++j) tma_load_2d_pair(sA + j * 8192, &tmA, fb, m0 + 64 * j, kb * BK);
            if (IM2COL) {
              int n, h, w;
              decode_pix(kb * BK, p.rowH, p.rowW, n, h, w);
              const int bw = w * p.stride - p.im2col_pad_w, bh = h * p.stride - p.im2col_pad_h;
#pragma unroll
              for (int j = 0; j < BNH / 64; ++j) {
                // flattened (tap, cin) column of this box; columns past the last tap read tap 0 (never stored)
                int col = nB + 64 * j;
                if (col >= p.Ncols) col = 0;
                const int tp = col / p.tapC, ci0 = col - tp * p.tapC;
                const int kh = tp / p.KW, kw = tp - kh * p.KW;
                tma_load_im2col_pair(sB + j * 8192, &tmB, fb, ci0, bw, bh, n, (uint16_t)kw, (uint16_t)kh);
              }
            } else {
#pragma unroll
              for (int j = 0; j < BNH / 64; ++j) tma_load_2d_pair(sB + j * 8192, &tmB, fb, nB + 64 * j, kb * BK);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer (leader CTA only)
    if (lane == 0 && rank == 0) {
      constexpr bool kAmn = (KIND == WGRAD);
      constexpr bool kBmn = (KIND != FPROP);
      constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BN, kAmn, kBmn);
      uint32_t g = 0, t = 0;
      for (int item = pair; item < num_items; item += npairs, ++t) {
        const Item it = decode_item(p, item, KIND, BN);
        const uint32_t acc = t & 1;
        mbar_wait(&tempty[acc], ((t >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int i = 0; i < it.nkb; ++i, ++g) {
          const int s = g % C::kStages;
          const uint32_t ph = (g / C::kStages) & 1;
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint32_t aBase = smem_u32(smem + s * C::kStageBytes);
          const uint32_t bBase = aBase + C::kABytes;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = kAmn ? make_smem_desc_sw128(aBase + k * 2048, 8192, 1024)
                                     : make_smem_desc_sw128(aBase + k * 32, 16, 1024);
            const uint64_t db = kBmn ? make_smem_desc_sw128(bBase + k * 2048, 8192, 1024)
                                     : make_smem_desc_sw128(bBase + k * 32, 16, 1024);
            umma_bf16_pair(d_tmem, da, db, idesc, (i | k) != 0);
          }
          umma_commit_pair(&empty[s]);   // frees the stage in both CTAs
        }
        umma_commit_pair(&tfull[acc]);   // wakes the epilogue warps of both CTAs
      }
    }
  } else {
    // ===================================================== epilogue (both CTAs, own 128 rows)
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const bool store_thread = (threadIdx.x == 64);
    const bool tma_epi = (KIND != WGRAD) && p.cls == 1;
    uint32_t t = 0, panel_seq = 0;
    for (int item = pair; item < num_items; item += npairs, ++t) {
      const Item it = decode_item(p, item, KIND, BN);
      const uint32_t acc = t & 1;
      mbar_wait(&tfull[acc], (t >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + (uint32_t(quad * 32) << 16);
      const int m0 = it.m0 * 2 + 128 * (int)rank;
      const int m = m0 + row;
      const bool ok = m < p.M;
      if (KIND != WGRAD && tma_epi) {
#pragma unroll 1
        for (int pn = 0; pn < BN / 64; ++pn, ++panel_seq) {
          if (it.n0 + pn * 64 >= p.Ncols) break;
          uint8_t* buf = epi_smem + (panel_seq & 1) * (BM * 128);
          if (store_thread) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            uint32_t v[32];
            tmem_ld32(taddr + pn * 64 + h * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              if (p.bias) {
                const int col = it.n0 + pn * 64 + h * 32 + q * 8;
#pragma unroll
                for (int e = 0; e < 8; ++e)
                  if (col + e < p.Ncols)
                    v[8 * q + e] = __float_as_uint(__uint_as_float(v[8 * q + e]) + __ldg(p.bias + col + e));
              }
              uint4 o;
              o.x = pack_bf16x2(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1]));
              o.y = pack_bf16x2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
              o.z = pack_bf16x2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
              o.w = pack_bf16x2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
              const int chunk = h * 4 + q;
              *reinterpret_cast<uint4*>(buf + row * 128 + ((chunk ^ (row & 7)) << 4)) = o;
            }
          }
          fence_proxy_async_smem();
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (store_thread) {
            tma_store_2d(buf, &tmC, it.n0 + pn * 64, m0);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        }
      } else if (KIND != WGRAD) {
        long long orow_idx = m;
        if (KIND == DGRAD && p.cls > 1 && ok) {
          int n, i, j;
          decode_pix(m, p.rowH, p.rowW, n, i, j);
          orow_idx = ((long long)n * p.outH + (i * p.cls + p.ph)) * p.outW + (j * p.cls + p.pw);
        }
        __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(p.out) + orow_idx * p.out_ld + it.n0;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t v[32];
          tmem_ld32(taddr + c * 32, v);
          tmem_ld_wait();
          if (ok) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const int col = it.n0 + c * 32 + q * 8;
              if (col + 8 > p.Ncols) break;
              uint4 o;
              o.x = pack_bf16x2(__uint_as_float(v[8 * q + 0]), __uint_as_float(v[8 * q + 1]));
              o.y = pack_bf16x2(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
              o.z = pack_bf16x2(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
              o.w = pack_bf16x2(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
              st_v4(orow + c * 32 + q * 8, o);
            }
          }
        }
      } else {
        float* out = reinterpret_cast<float*>(p.out) + (size_t)it.split * p.split_stride;
        float* orow = out + (size_t)m * p.out_ld + it.n0;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t v[32];
          tmem_ld32(taddr + c * 32, v);
          tmem_ld_wait();
          if (ok) {
#pragma unroll
            for (int q = 0; q < 8; ++q)
              if (it.n0 + c * 32 + q * 4 + 4 <= p.Ncols)
                st_v4(orow + c * 32 + q * 4, make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]));
          }
        }
      }
      // release the accumulator stage: the leader's MMA warp counts the epilogue warps of both CTAs
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_cluster(smem_u32(&tempty[acc]), 0));
    }
  }

  if (threadIdx.x == 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  tc_fence_before();
  cluster_sync_all();   // neither CTA may free TMEM or exit while its partner still uses its memory / barriers
  if (warp == 1) tmem_free_pair<2 * BN>(tmem_base);
}

static int num_sms() { return b2::host_num_sms(); }

template <int KIND, int AMODE, int BN, int MT = 1, int FAN = 0>
static cudaError_t launch2(const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvParams& p_in, int items,
                           cudaStream_t st) {
  ConvParams p = p_in;
  p.dbg = g_dbg;
  CUtensorMap tmC = tmB;  // wgrad and class-strided dgrad never dereference it
  if (KIND != WGRAD && p.cls == 1) {
    int e = make_tmap_2d_bf16(&tmC, p.out, (uint64_t)p.Ncols, (uint64_t)p.M, 64, 128);
    if (e) return (cudaError_t)e;
  }
  auto kern = igemm2_kernel<KIND, AMODE, BN, MT, FAN>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg2<BN, MT>::kSmem);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  const int grid = items < num_sms() ? items : num_sms();
  const int threads = (KIND != WGRAD && (p.stats || FAN)) ? kThreads2 + 32 * kStatWarps2<BN> + (FAN ? 32 : 0)
                                                  : (KIND == WGRAD ? kThreads2W : kThreads2);
  kern<<<grid, threads, Cfg2<BN, MT>::kSmem, st>>>(tmA, tmB, tmC, p, items);
  return cudaGetLastError();
}
template <int KIND, int AMODE>
static cudaError_t launch2_bn(int bn, const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvParams& p, int items,
                              cudaStream_t st) {
  switch (bn) {
    case 64: return launch2<KIND, AMODE, 64>(tmA, tmB, p, items, st);
    case 128: return launch2<KIND, AMODE, 128>(tmA, tmB, p, items, st);
    case 256: return launch2<KIND, AMODE, 256>(tmA, tmB, p, items, st);
  }
  return cudaErrorInvalidValue;
}
template <int KIND>
static cudaError_t launch2_mode(int mode, int bn, const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvParams& p,
                                int items, cudaStream_t st) {
  if (mode == ACT_TILED) return launch2_bn<KIND, ACT_TILED>(bn, tmA, tmB, p, items, st);
  if (mode == ACT_IM2COL) return launch2_bn<KIND, ACT_IM2COL>(bn, tmA, tmB, p, items, st);
  return cudaErrorInvalidValue;
}

template <int KIND, int AMODE, int BN>
static cudaError_t launch3(const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvParams& p, int items,
                           cudaStream_t st) {
  auto kern = igemm3_kernel<KIND, AMODE, BN>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg3<BN>::kSmem);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  CUtensorMap tmC = tmB;
  if (KIND != WGRAD && p.cls == 1) {
    int e = make_tmap_2d_bf16(&tmC, p.out, (uint64_t)p.Ncols, (uint64_t)p.M, 64, 128);
    if (e) return (cudaError_t)e;
  }
  const int pairs = num_sms() / 2;
  const int grid = 2 * (items < pairs ? items : pairs);
  kern<<<grid, kThreads3, Cfg3<BN>::kSmem, st>>>(tmA, tmB, tmC, p, items);
  return cudaGetLastError();
}
template <int KIND>
static cudaError_t launch3_mode(int mode, int bn, const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvParams& p,
                                int items, cudaStream_t st) {
  if (mode == ACT_TILED) {
    if (bn == 256) return launch3<KIND, ACT_TILED, 256>(tmA, tmB, p, items, st);
    if (bn == 128) return launch3<KIND, ACT_TILED, 128>(tmA, tmB, p, items, st);
  } else if (mode == ACT_IM2COL) {
    if (bn == 256) return launch3<KIND, ACT_IM2COL, 256>(tmA, tmB, p, items, st);
    if (bn == 128) return launch3<KIND, ACT_IM2COL, 128>(tmA, tmB, p, items, st);
  }
  return cudaErrorInvalidValue;
}

static int g_fuse_stats = 1;  // fprop epilogue also produces the BatchNorm statistics partials
static int g_impl = 2;  // 1 = one tile per CTA (v1), 2 = persistent (v2), 3 = persistent CTA pairs where they fit
// CTA pairs need a 256-row M tile and at least a 128-wide N tile.
static bool use_pairs(int bn, long long m_extent) { return g_impl == 3 && (bn == 128 || bn == 256) && m_extent >= 256; }
// Default build: the CTA-pair kernel (cta_group::2, each CTA stages half of the B tile) takes the plain 1x1 GEMMs
// with 256-wide tiles and K >= 256 - the evaluation-mode forward and the dgrads that produce no BatchNorm sums.
// Measured 4-10 % faster than the persistent kernel on every such ResNet-50 shape (profiles/impl_sweep_bs256_r2.md);
// it has no statistics warps and its wgrad is slower, so everything else stays on v2.
static int g_pairs_auto = 1;
extern "C" void b2_conv_set_pairs_auto(int on) { g_pairs_auto = on; }
static bool auto_pairs(int bn, long long m_extent, int KH, int KW, int kdepth, bool plain) {
  return g_pairs_auto && g_impl == 2 && g_act_mode != ACT_GATHER && !g_force_bn && plain && KH == 1 && KW == 1 &&
         bn == 256 && kdepth >= 256 && m_extent >= 256;
}
}  // namespace b2

using namespace b2;

extern "C" void b2_conv_set_tile_n(int bn) { g_force_bn = bn; }
extern "C" void b2_conv_set_act_mode(int mode) { g_act_mode = mode; }
extern "C" void b2_conv_set_debug(int bits) { g_dbg = bits; }
extern "C" void b2_conv_set_panels(int on) { g_panels = on; }
extern "C" void b2_conv_set_wgrad_mt2(int on) { g_mt2 = on; }
extern "C" void b2_conv_set_wgrad_waves(int n) { g_wgrad_waves = n; }
extern "C" void b2_conv_set_wgrad_red(int on) { g_wgrad_red = on; }

extern "C" void b2_conv_set_impl(int impl) { g_impl = impl; }

// N-tile policy (measured on B200, batch 256, tools/conv_check.py): the persistent kernel prefers the widest tile
// the channel count allows; wgrad only when Cout fills at least two M tiles.
static int bn_for(int ncols, int kind = FPROP, int m_extent = 1 << 30) {
  if (g_force_bn) return pick_bn(ncols, g_force_bn);
  if (g_impl >= 2) {
    const bool wide_ok = (kind != WGRAD) || m_extent >= 256;
    if (ncols % 256 == 0 && wide_ok) return 256;
  }
  return pick_bn(ncols, 0);
}

// stats / stats_blocks: optional. When the persistent kernel runs, it also leaves per-CTA partial BatchNorm
// statistics of y in stats ([*stats_blocks][2][Cout] fp32); *stats_blocks == 0 means "not produced, run bn_stats".
static int conv_fprop_impl(const void* x, const void* w, const float* bias, void* y, int N, int H, int W, int Cin,
                           int Cout, int KH, int KW, int stride, int pad, float* stats, int* stats_blocks,
                           void* stream) {
  if (stats_blocks) *stats_blocks = 0;
  if (Cin % 64 || Cout % 8) return cudaErrorInvalidValue;
  // narrow 3x3 / stride 1 layers: weight-stationary kernel that loads every input pixel once (csrc/conv3x3.cu)
  if (KH == 3 && KW == 3 && stride == 1 && pad == 1 && !bias && g_impl == 2 && g_act_mode != ACT_GATHER &&
      b2_conv_halo_supported(N, H, W, Cin, Cout))
    return b2_conv3x3_halo(0, x, w, y, N, H, W, Cin, Cout, g_fuse_stats ? stats : nullptr, stats_blocks, nullptr,
                           nullptr, nullptr, stream);
  const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
  const int M = N * Ho * Wo, K = KH * KW * Cin;
  const int bn = bn_for(Cout);
  const int mode = pick_mode(FPROP, KH, KW, stride, pad);
  CUtensorMap tmA, tmB;
  int e;
  // A: x2d [M][Cin] on the pure-GEMM path, im2col view of x otherwise; B: W2d [Cout][K], box {64, bn}
  if (mode == ACT_IM2COL) {
    if ((e = make_tmap_im2col_bf16(&tmA, x, N, H, W, Cin, -pad, -pad, pad - (KW - 1), pad - (KH - 1), 64, 128,
                                   stride)))
      return e;
  } else if ((e = make_tmap_2d_bf16(&tmA, x, (uint64_t)Cin, (uint64_t)N * H * W, 64, 128))) {
    return e;
  }
  const bool pairs = (use_pairs(bn, M) || auto_pairs(bn, M, KH, KW, K, !stats && !bias)) && mode != ACT_GATHER &&
                     Cout % bn == 0;
  if ((e = make_tmap_2d_bf16(&tmB, w, (uint64_t)K, (uint64_t)Cout, 64, pairs ? bn / 2 : bn))) return e;
  ConvParams p{};
  p.src = (const __nv_bfloat16*)x; p.srcH = H; p.srcW = W; p.srcC = Cin;
  p.rowH = Ho; p.rowW = Wo; p.KH = KH; p.KW = KW; p.stride = stride; p.pad = pad;
  p.M = M; p.Ncols = Cout; p.Kblocks = K / 64; p.cpb = Cin / 64; p.tapC = Cin;
  p.out = y; p.out_ld = Cout; p.splits = 1; p.bias = bias; p.im2col_pad = pad;
  p.im2col_pad_h = pad; p.im2col_pad_w = pad; p.cls = 1; p.taps = KH * KW;
  p.n_tiles = (Cout + bn - 1) / bn; p.m_tiles = (M + BM - 1) / BM;
  cudaStream_t st = (cudaStream_t)stream;
  if (pairs) {
    p.m_tiles = (M + 2 * BM - 1) / (2 * BM);
    return launch3_mode<FPROP>(mode, bn, tmA, tmB, p, p.n_tiles * p.m_tiles, st);
  }
  if (g_impl >= 2 && mode != ACT_GATHER) {
    const int items = p.n_tiles * p.m_tiles;
    if (stats && stats_blocks && g_fuse_stats && Cout % bn == 0) {   // every 64-column panel of every tile exists
      const int grid = items < num_sms() ? items : num_sms();   // must match launch2()
      // one (256-wide tile) or two partial rows per CTA, zeroed by the kernel itself
      p.stats = stats;
      *stats_blocks = (bn == 256 ? 1 : 2) * grid;
    }
    return launch2_mode<FPROP>(mode, bn, tmA, tmB, p, items, st);
  }
  dim3 grid(p.n_tiles, p.m_tiles, 1);
  return launch_mode<FPROP>(mode, bn, tmA, tmB, p, grid, st);
}

extern "C" int b2_conv_fprop(const void* x, const void* w, const float* bias, void* y, int N, int H, int W, int Cin,
                             int Cout, int KH, int KW, int stride, int pad, void* stream) {
  return conv_fprop_impl(x, w, bias, y, N, H, W, Cin, Cout, KH, KW, stride, pad, nullptr, nullptr, stream);
}
extern "C" int b2_conv_fprop_stats(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout,
                                   int KH, int KW, int stride, int pad, float* stats, int* stats_blocks,
                                   void* stream) {
  return conv_fprop_impl(x, w, nullptr, y, N, H, W, Cin, Cout, KH, KW, stride, pad, stats, stats_blocks, stream);
}
extern "C" void b2_conv_set_fuse_stats(int on) { g_fuse_stats = on; }

// Strided dgrad, one output-pixel parity class at a time (persistent kernel + TMA im2col over dy).
static int dgrad_strided_v2(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int KH,
                            int KW, int stride, int pad, int Ho, int Wo, int bn, cudaStream_t st) {
  CUtensorMap tmB;
  int e;
  const bool panelsB = g_panels && g_impl == 2;   // every parity class goes through the persistent kernel
  if (panelsB) {
    if ((e = make_tmap_panels_bf16(&tmB, w, (uint64_t)KH * KW * Cin, (uint64_t)Cout, 64, bn / 64))) return e;
  } else if ((e = make_tmap_2d_bf16(&tmB, w, (uint64_t)KH * KW * Cin, (uint64_t)Cout, 64, 64))) {
    return e;
  }
  const int cpb = (Cout + 63) / 64;
  bool need_zero = false;
  for (int ph = 0; ph < stride; ++ph)
    for (int pw = 0; pw < stride; ++pw) {
      const int kh0 = (ph + pad) % stride, kw0 = (pw + pad) % stride;
      if (kh0 >= KH || kw0 >= KW || ph >= H || pw >= W) need_zero = true;
    }
  if (need_zero) {
    cudaError_t ce = cudaMemsetAsync(dx, 0, (size_t)N * H * W * Cin * 2, st);
    if (ce != cudaSuccess) return ce;
  }
  for (int ph = 0; ph < stride && ph < H; ++ph) {
    for (int pw = 0; pw < stride && pw < W; ++pw) {
      const int kh0 = (ph + pad) % stride, kw0 = (pw + pad) % stride;
      if (kh0 >= KH || kw0 >= KW) continue;  // no tap reaches this class: dx stays zero
      const int nth = (KH - kh0 + stride - 1) / stride, ntw = (KW - kw0 + stride - 1) / stride;
      const int d0h = (ph + pad - kh0) / stride, d0w = (pw + pad - kw0) / stride;
      const int Hc = (H - ph + stride - 1) / stride, Wc = (W - pw + stride - 1) / stride;
      const int lower_h = d0h - (nth - 1), lower_w = d0w - (ntw - 1);
      const int upper_h = Hc - Ho + lower_h, upper_w = Wc - Wo + lower_w;
      CUtensorMap tmA;
      if ((e = make_tmap_im2col_bf16(&tmA, dy, N, Ho, Wo, Cout, lower_w, lower_h, upper_w, upper_h, 64, 128, 1)))
        return e;
      ConvParams p{};
      p.src = (const __nv_bfloat16*)dy; p.srcH = Ho; p.srcW = Wo; p.srcC = Cout;
      p.rowH = Hc; p.rowW = Wc; p.KH = KH; p.KW = KW; p.stride = 1; p.pad = pad;
      p.M = N * Hc * Wc; p.Ncols = Cin; p.cpb = cpb; p.Kblocks = nth * ntw * cpb; p.tapC = Cin;
      p.out = dx; p.out_ld = Cin; p.splits = 1;
      p.im2col_pad_h = -lower_h; p.im2col_pad_w = -lower_w;
      p.kh0 = kh0; p.kw0 = kw0; p.kstep = stride; p.nth = nth; p.ntw = ntw;
      p.cls = stride; p.ph = ph; p.pw = pw; p.outH = H; p.outW = W; p.taps = KH * KW;
      p.panels = panelsB ? 2 : 0;
      p.n_tiles = Cin / bn; p.m_tiles = (p.M + BM - 1) / BM;
      cudaError_t ce;
      if (use_pairs(bn, p.M)) {
        p.m_tiles = (p.M + 2 * BM - 1) / (2 * BM);
        ce = launch3_mode<DGRAD>(ACT_IM2COL, bn, tmA, tmB, p, p.n_tiles * p.m_tiles, st);
      } else {
        ce = launch2_mode<DGRAD>(ACT_IM2COL, bn, tmA, tmB, p, p.n_tiles * p.m_tiles, st);
      }
      if (ce != cudaSuccess) return ce;
    }
  }
  return 0;
}

struct DgradBn {   // optional BatchNorm-backward statistics request (see ConvParams::bwd_y)
  const void* y;
  const void* mask;
  const float* mean;
  float* partial;
  int* blocks;
};
static int conv_dgrad_impl(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int KH,
                           int KW, int stride, int pad, const void* add, const void* mask, void* stream,
                           const DgradBn* bn = nullptr);

extern "C" int b2_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int KH,
                             int KW, int stride, int pad, void* stream) {
  return conv_dgrad_impl(dy, w, dx, N, H, W, Cin, Cout, KH, KW, stride, pad, nullptr, nullptr, stream);
}

// Residual fan-in fused into dgrad: dx = mask ? (dgrad + add) : 0 (add: bf16 NHWC like dx; mask: 1 bit per element,
// nullable), computed by the statistics warps on the bf16 panel in shared memory, together with the BatchNorm-backward
// sums of dx (as b2_conv_dgrad_bnstats, with g = dx already masked). Returns -1 (nothing launched) when the active
// kernel variant cannot do it: the caller then runs the plain dgrad and the two-input BatchNorm backward.
extern "C" int b2_conv_dgrad_fused(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                                   int KH, int KW, int stride, int pad, const void* add, const void* mask,
                                   const void* bn_y, const float* bn_mean, float* partial, int* blocks,
                                   void* stream) {
  if (!add || !blocks || (bn_y && (!bn_mean || !partial))) return cudaErrorInvalidValue;
  *blocks = 0;
  const int bn = bn_for(Cin);
  const bool ok = g_impl == 2 && g_fuse_stats && stride == 1 && Cin % 64 == 0 && Cin % bn == 0 &&
                  pick_mode(DGRAD, KH, KW, stride, pad) != ACT_GATHER;
  if (!ok) return -1;
  DgradBn req{bn_y, nullptr, bn_mean, partial, blocks};
  return conv_dgrad_impl(dy, w, dx, N, H, W, Cin, Cout, KH, KW, stride, pad, add, mask, stream, &req);
}

// dgrad that also leaves the partial sums of the BatchNorm backward of dx behind (dx is the gradient of that
// BatchNorm's ReLU output): partial receives [*blocks][2][Cin] rows of (sum g, sum g*(y-mean)), g = mask ? dx : 0,
// to be handed to b2_bn_bwd_partial. *blocks == 0: not produced by the active kernel variant (run the reduce pass).
extern "C" int b2_conv_dgrad_bnstats(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout,
                                     int KH, int KW, int stride, int pad, const void* bn_y, const void* bn_mask,
                                     const float* bn_mean, float* partial, int* blocks, void* stream) {
  if (!blocks || !partial || !bn_y || !bn_mean) return cudaErrorInvalidValue;
  *blocks = 0;
  DgradBn bn{bn_y, bn_mask, bn_mean, partial, blocks};
  return conv_dgrad_impl(dy, w, dx, N, H, W, Cin, Cout, KH, KW, stride, pad, nullptr, nullptr, stream, &bn);
}

static int conv_dgrad_impl(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int KH,
                           int KW, int stride, int pad, const void* add, const void* mask, void* stream,
                           const DgradBn* bnreq) {
  if (Cin % 64 || Cout % 8) return cudaErrorInvalidValue;
  if (KH == 3 && KW == 3 && stride == 1 && pad == 1 && !add && !mask && g_impl == 2 && g_act_mode != ACT_GATHER &&
      b2_conv_halo_supported(N, H, W, Cout, Cin)) {
    const bool st = bnreq && g_fuse_stats;
    return b2_conv3x3_halo(1, dy, w, dx, N, H, W, Cin, Cout, st ? bnreq->partial : nullptr, st ? bnreq->blocks : nullptr,
                           st ? bnreq->y : nullptr, st ? bnreq->mask : nullptr, st ? bnreq->mean : nullptr, stream);
  }
  const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
  const int M = N * H * W;
  const int cpb = (Cout + 63) / 64;
  const int bn = bn_for(Cin);
  cudaStream_t st = (cudaStream_t)stream;
  if (stride > 1 && g_impl >= 2 && g_act_mode != ACT_GATHER && Cout % 64 == 0)
    return dgrad_strided_v2(dy, w, dx, N, H, W, Cin, Cout, KH, KW, stride, pad, Ho, Wo, bn, st);
  const int mode = pick_mode(DGRAD, KH, KW, stride, pad);
  CUtensorMap tmA, tmB;
  int e;
  if (mode == ACT_IM2COL) {  // dy viewed through the transposed convolution: pad' = K-1-pad, stride 1
    const int pw = KW - 1 - pad, phh = KH - 1 - pad;
    if ((e = make_tmap_im2col_bf16(&tmA, dy, N, Ho, Wo, Cout, -pw, -phh, pw - (KW - 1), phh - (KH - 1), 64, 128, 1)))
      return e;
  } else if ((e = make_tmap_2d_bf16(&tmA, dy, (uint64_t)Cout, (uint64_t)N * Ho * Wo, 64, 128))) {
    return e;
  }
  const bool pairs_d = (use_pairs(bn, M) || auto_pairs(bn, M, KH, KW, Cout, !bnreq && !add && !mask && stride == 1 &&
                                                                            Cin % bn == 0)) &&
                       mode != ACT_GATHER;
  const bool panelsB = g_panels && g_impl == 2 && mode != ACT_GATHER && !pairs_d;   // the persistent kernel runs
  if (panelsB) {
    if ((e = make_tmap_panels_bf16(&tmB, w, (uint64_t)KH * KW * Cin, (uint64_t)Cout, 64, bn / 64))) return e;
  } else if ((e = make_tmap_2d_bf16(&tmB, w, (uint64_t)KH * KW * Cin, (uint64_t)Cout, 64, 64))) {
    return e;
  }
  ConvParams p{};
  p.panels = panelsB ? 2 : 0;
  p.add = (const __nv_bfloat16*)add;
  p.amask = (const unsigned char*)mask;
  p.src = (const __nv_bfloat16*)dy; p.srcH = Ho; p.srcW = Wo; p.srcC = Cout;
  p.rowH = H; p.rowW = W; p.KH = KH; p.KW = KW; p.stride = stride; p.pad = pad;
  p.M = M; p.Ncols = Cin; p.Kblocks = KH * KW * cpb; p.cpb = cpb; p.tapC = Cin;
  p.out = dx; p.out_ld = Cin; p.splits = 1; p.im2col_pad = KW - 1 - pad;
  p.im2col_pad_h = KH - 1 - pad; p.im2col_pad_w = KW - 1 - pad;
  p.kh0 = 0; p.kw0 = 0; p.kstep = 1; p.nth = KH; p.ntw = KW; p.cls = 1; p.outH = H; p.outW = W; p.taps = KH * KW;
  if (mode == ACT_IM2COL) p.stride = 1;
  p.n_tiles = Cin / bn; p.m_tiles = (M + BM - 1) / BM;
  if (pairs_d) {
    p.m_tiles = (M + 2 * BM - 1) / (2 * BM);
    return launch3_mode<DGRAD>(mode, bn, tmA, tmB, p, p.n_tiles * p.m_tiles, st);
  }
  if (g_impl >= 2 && mode != ACT_GATHER) {
    const int items = p.n_tiles * p.m_tiles;
    if (bnreq && g_fuse_stats && g_impl == 2 && Cin % bn == 0) {   // TMA-store epilogue, every panel exists
      const int grid = items < num_sms() ? items : num_sms();       // must match launch2()
      p.stats = bnreq->partial;
      p.bwd_y = (const __nv_bfloat16*)bnreq->y;
      p.bwd_mask = (const unsigned char*)bnreq->mask;
      p.bwd_mean = bnreq->mean;
      p.fanin = add ? 1 : 0;
      if (!bnreq->y) p.stats = nullptr;    // fan-in only, no BatchNorm-backward sums
      *bnreq->blocks = !bnreq->y ? 0 : (add ? (bn == 64 ? 2 : 4) : (bn == 256 ? 1 : 2)) * grid;
    }
    if (p.fanin) {
      if (mode == ACT_TILED) {
        if (bn == 256) return launch2<DGRAD, ACT_TILED, 256, 1, 1>(tmA, tmB, p, items, st);
        if (bn == 128) return launch2<DGRAD, ACT_TILED, 128, 1, 1>(tmA, tmB, p, items, st);
        return launch2<DGRAD, ACT_TILED, 64, 1, 1>(tmA, tmB, p, items, st);
      }
      if (bn == 256) return launch2<DGRAD, ACT_IM2COL, 256, 1, 1>(tmA, tmB, p, items, st);
      if (bn == 128) return launch2<DGRAD, ACT_IM2COL, 128, 1, 1>(tmA, tmB, p, items, st);
      return launch2<DGRAD, ACT_IM2COL, 64, 1, 1>(tmA, tmB, p, items, st);
    }
    return launch2_mode<DGRAD>(mode, bn, tmA, tmB, p, items, st);
  }
  dim3 grid(p.n_tiles, p.m_tiles, 1);
  return launch_mode<DGRAD>(mode, bn, tmA, tmB, p, grid, st);
}

static bool wgrad_mt2(int bn, int Cout) {
  return g_mt2 && g_impl == 2 && g_act_mode != ACT_GATHER && bn == 256 && Cout % 256 == 0;
}

// Split-K plan for wgrad: enough work items for ~2-4 per SM, bounded by the workspace.
extern "C" int b2_conv_wgrad_splits(int N, int Ho, int Wo, int Cin, int Cout, int taps, long long ws_bytes) {
  const long long kpix = (long long)N * Ho * Wo;
  const int kblocks = (int)((kpix + 63) / 64);
  const bool flat = (g_impl >= 2 && taps > 1 && g_act_mode != ACT_GATHER && !g_force_bn);
  const int bn = flat ? 256 : bn_for(Cin, WGRAD, Cout);
  const long long tiles = flat ? (long long)((Cout + BM - 1) / BM) * ((taps * Cin + 255) / 256)
                               : (long long)((Cout + BM - 1) / BM) * (Cin / bn) * taps;
  const int sms = num_sms();
  long long want = ((g_impl >= 2 ? 2 : 4) * sms + tiles - 1) / tiles;
  if (g_impl == 2 && g_wgrad_waves == 1 && tiles <= sms) want = sms / tiles;   // one item per SM, single wave
  if (wgrad_mt2(bn, Cout)) {   // half as many (256-row) tiles, one item per SM in a single wave
    const long long t2 = tiles / 2;
    want = sms / t2;
  }
  const long long wbytes = (long long)Cout * taps * Cin * 4;
  long long cap = ws_bytes / wbytes;
  if (want > cap) want = cap;
  if (want > kblocks) want = kblocks;
  if (want < 1) want = 1;
  const int per = (int)((kblocks + want - 1) / want);
  return (kblocks + per - 1) / per;
}

extern "C" int b2_conv_wgrad(const void* x, const void* dy, float* dw, float* workspace, long long ws_bytes, int N,
                             int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad, int accumulate,
                             void* stream) {
  if (Cin % 64 || Cout % 8) return cudaErrorInvalidValue;
  const bool fixed_order = (accumulate & 2) != 0;   // caller asks for run-to-run identical sums
  accumulate &= 1;
  const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
  const int taps = KH * KW;
  const long long kpix = (long long)N * Ho * Wo;
  const int kblocks = (int)((kpix + 63) / 64);
  const int mode = pick_mode(WGRAD, KH, KW, stride, pad);
  const bool flat = (g_impl >= 2 && taps > 1 && mode == ACT_IM2COL && !g_force_bn);
  const int bn = flat ? 256 : bn_for(Cin, WGRAD, Cout);
  const long long wcount = (long long)Cout * taps * Cin;
  if (ws_bytes < wcount * 4) return cudaErrorInvalidValue;
  const int splits = b2_conv_wgrad_splits(N, Ho, Wo, Cin, Cout, taps, ws_bytes);
  const int per = (kblocks + splits - 1) / splits;
  CUtensorMap tmA, tmB;
  int e;
  const bool v2 = g_impl == 2 && mode != ACT_GATHER;   // the persistent (non-pair) kernel runs
  const bool panelsA = g_panels && v2 && Cout % 64 == 0;
  const bool panelsB = g_panels && v2 && mode != ACT_IM2COL;
  const bool mt2 = v2 && wgrad_mt2(bn, Cout);
  if (panelsA) {
    if ((e = make_tmap_panels_bf16(&tmA, dy, (uint64_t)Cout, (uint64_t)kpix, 64, (mt2 ? 2 : 1) * BM / 64))) return e;
  } else if ((e = make_tmap_2d_bf16(&tmA, dy, (uint64_t)Cout, (uint64_t)kpix, 64, 64))) {
    return e;
  }
  if (mode == ACT_IM2COL) {
    if ((e = make_tmap_im2col_bf16(&tmB, x, N, H, W, Cin, -pad, -pad, pad - (KW - 1), pad - (KH - 1), 64, 64, stride)))
      return e;
  } else if (panelsB) {
    if ((e = make_tmap_panels_bf16(&tmB, x, (uint64_t)Cin, (uint64_t)N * H * W, 64, bn / 64))) return e;
  } else if ((e = make_tmap_2d_bf16(&tmB, x, (uint64_t)Cin, (uint64_t)N * H * W, 64, 64))) {
    return e;
  }
  ConvParams p{};
  p.panels = (panelsA ? 1 : 0) | (panelsB ? 2 : 0);
  p.src = (const __nv_bfloat16*)x; p.srcH = H; p.srcW = W; p.srcC = Cin;
  p.rowH = Ho; p.rowW = Wo; p.KH = KH; p.KW = KW; p.stride = stride; p.pad = pad; p.im2col_pad = pad;
  p.im2col_pad_h = pad; p.im2col_pad_w = pad; p.cls = 1; p.taps = taps;
  p.M = Cout; p.Ncols = Cin; p.Kblocks = kblocks; p.Kpix = (int)kpix; p.tapC = Cin;
  p.splits = splits; p.kb_per_split = per;
  p.out = workspace; p.out_ld = (long long)taps * Cin; p.split_stride = wcount;
  p.n_tiles = Cin / bn; p.m_tiles = (Cout + BM - 1) / BM;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t ce;
  const bool pairs = use_pairs(bn, Cout) && Cout % 256 == 0 && mode != ACT_GATHER && (flat || taps == 1);
  // Split-K sums formed in L2 by the epilogue's vector reductions, straight into dW: the persistent kernel only.
  // A single split that overwrites needs neither the reductions nor the slab.
  const bool direct = g_wgrad_red && !fixed_order && !pairs && (flat || (g_impl >= 2 && mode != ACT_GATHER));
  if (direct) {
    p.out = dw; p.split_stride = 0;
    p.red_out = (splits > 1 || accumulate) ? 1 : 0;
    if (p.red_out && !accumulate) {
      ce = cudaMemsetAsync(dw, 0, (size_t)wcount * 4, st);
      if (ce != cudaSuccess) return ce;
    }
  }
  if (pairs) {   // 256 rows of Cout per CTA pair; N is the flattened (tap, cin) axis
    p.flatN = 1;
    p.Ncols = taps * Cin;
    p.n_tiles = (p.Ncols + bn - 1) / bn;
    p.m_tiles = Cout / 256;
    ce = launch3_mode<WGRAD>(mode, bn, tmA, tmB, p, p.n_tiles * p.m_tiles * splits, st);
  } else if (flat) {
    p.flatN = 1;
    p.Ncols = taps * Cin;
    p.n_tiles = (p.Ncols + 255) / 256;
    if (mt2) {
      p.m_tiles = Cout / 256;
      ce = launch2<WGRAD, ACT_IM2COL, 256, 2>(tmA, tmB, p, p.n_tiles * p.m_tiles * splits, st);
    } else {
      ce = launch2_mode<WGRAD>(mode, bn, tmA, tmB, p, p.n_tiles * p.m_tiles * splits, st);
    }
  } else if (g_impl >= 2 && mode != ACT_GATHER) {
    if (mt2) {
      p.m_tiles = Cout / 256;
      const int items = p.n_tiles * p.m_tiles * taps * splits;
      ce = mode == ACT_TILED ? launch2<WGRAD, ACT_TILED, 256, 2>(tmA, tmB, p, items, st)
                             : launch2<WGRAD, ACT_IM2COL, 256, 2>(tmA, tmB, p, items, st);
    } else {
      ce = launch2_mode<WGRAD>(mode, bn, tmA, tmB, p, p.n_tiles * p.m_tiles * taps * splits, st);
    }
  } else {
    dim3 grid(p.n_tiles, p.m_tiles, taps * splits);
    ce = launch_mode<WGRAD>(mode, bn, tmA, tmB, p, grid, st);
  }
  if (ce != cudaSuccess) return ce;
  if (direct) return cudaSuccess;
  const int thr = 256;
  const long long nvec = (wcount + 3) / 4;
  wgrad_reduce_kernel<<<(unsigned)((nvec + thr - 1) / thr), thr, 0, st>>>(workspace, dw, wcount, splits, wcount,
                                                                          accumulate);
  return cudaGetLastError();
}
