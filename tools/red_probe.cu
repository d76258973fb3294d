// Micro-benchmark: is accumulating the wgrad split-K partials in L2 (red.global.add.v4.f32 from the epilogue threads,
// or a bulk cp.reduce.async from shared memory) cheaper than writing fp32 slabs and reducing them in a second kernel?
// Access pattern of the igemm2 wgrad epilogue: 148 CTAs, 128 threads, one accumulator row per thread, 2 x 256 columns
// of fp32 per thread in 16-byte pieces; `splits` CTAs target the same 256 x 256 tile of dw.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/red_probe tools/red_probe.cu && build/red_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int kRows = 256, kCols = 256;   // one tile: 256 x 256 fp32 = 256 KB

__device__ __forceinline__ void red_v4(float* p, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// mode 0: plain stores into slab `split`; mode 1: red.v4 into the shared tile
template <int MODE>
__global__ void __launch_bounds__(128) epi_kernel(float* out, int tiles, int splits, long long ld, long long slab) {
  const int item = blockIdx.x;
  if (item >= tiles * splits) return;
  const int tile = item / splits, split = item % splits;
  float* base = out + (MODE == 0 ? (long long)split * slab : 0) + (long long)tile * kCols;
  for (int mt = 0; mt < 2; ++mt) {
    float* row = base + (long long)(mt * 128 + threadIdx.x) * ld;
#pragma unroll 1
    for (int c = 0; c < kCols / 32; ++c) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float4 v = make_float4(1.f, 2.f, 3.f, (float)split);
        if (MODE == 0) *reinterpret_cast<float4*>(row + c * 32 + q * 4) = v;
        else red_v4(row + c * 32 + q * 4, v);
      }
    }
  }
}

// Same bytes, but a warp instruction covers RPW rows x (512 / RPW) contiguous bytes (RPW = 32 is the row-per-thread
// pattern above, RPW = 1 one 512-byte run): what a transposing epilogue would issue.
template <int MODE, int RPW>
__global__ void __launch_bounds__(128) epi_coal_kernel(float* out, int tiles, int splits, long long ld, long long slab) {
  const int item = blockIdx.x;
  if (item >= tiles * splits) return;
  const int tile = item / splits, split = item % splits;
  float* base = out + (MODE == 0 ? (long long)split * slab : 0) + (long long)tile * kCols;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  constexpr int LPR = 32 / RPW;          // lanes per row
  constexpr int CPI = LPR * 4;           // columns per instruction and row
  const int r_in = lane / LPR, piece = lane % LPR;
  // warp w owns rows [w * 64, w * 64 + 64) of the 256-row tile
  for (int r0 = 0; r0 < 64; r0 += RPW) {
    float* row = base + (long long)(warp * 64 + r0 + r_in) * ld;
#pragma unroll 4
    for (int c = 0; c < kCols; c += CPI) {
      const float4 v = make_float4(1.f, 2.f, 3.f, (float)split);
      if (MODE == 0) *reinterpret_cast<float4*>(row + c + piece * 4) = v;
      else red_v4(row + c + piece * 4, v);
    }
  }
}

// mode 2: the tile goes through shared memory (row-major 256 x 32 column chunks) and one thread per row issues a bulk
// reduce of 128 bytes; measures the TMA reduction path with the same addresses
__global__ void __launch_bounds__(128) epi_bulk_kernel(float* out, int tiles, int splits, long long ld) {
  extern __shared__ __align__(128) float sm[];
  const int item = blockIdx.x;
  if (item >= tiles * splits) return;
  const int tile = item / splits, split = item % splits;
  float* base = out + (long long)tile * kCols;
  for (int mt = 0; mt < 2; ++mt) {
    float* srow = sm + threadIdx.x * kCols;
    for (int c = 0; c < kCols; c += 4) *reinterpret_cast<float4*>(srow + c) = make_float4(1.f, 2.f, 3.f, (float)split);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    float* row = base + (long long)(mt * 128 + threadIdx.x) * ld;
    const unsigned saddr = (unsigned)__cvta_generic_to_shared(srow);
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(row), "r"(saddr),
                 "r"(kCols * 4)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

__global__ void reduce_kernel(const float* ws, float* dw, long long n, int splits, long long slab) {
  const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= n) return;
  float4 a = make_float4(0, 0, 0, 0);
  for (int s = 0; s < splits; ++s) {
    const float4 v = __ldcs(reinterpret_cast<const float4*>(ws + (long long)s * slab + i));
    a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
  }
  *reinterpret_cast<float4*>(dw + i) = a;
}

int main() {
  struct Cfg { int tiles, splits; const char* name; };
  const Cfg cfgs[] = {{9, 16, "3x3 256->256 (9 tiles x 16)"}, {36, 4, "3x3 512->512 (36 x 4)"},
                      {4, 37, "1x1 256->1024 (4 x 37)"},      {1, 148, "1x1 64->256 (1 x 148)"}};
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float *ws, *dw, *flush;
  const long long maxslab = 36LL * kRows * kCols;
  CK(cudaMalloc(&ws, 148LL * kRows * kCols * 4 * 2));
  CK(cudaMalloc(&dw, maxslab * 4));
  CK(cudaMalloc(&flush, 256 << 20));
  CK(cudaFuncSetAttribute(epi_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * kCols * 4));
  for (const Cfg& c : cfgs) {
    const long long ld = (long long)c.tiles * kCols, slab = (long long)kRows * ld;
    const int items = c.tiles * c.splits;
    float t[4] = {0, 0, 0, 0};
    const int reps = 20;
    for (int mode = 0; mode < 4; ++mode) {
      for (int r = -3; r < reps; ++r) {
        CK(cudaMemsetAsync(flush, r & 1, 256 << 20));
        CK(cudaEventRecord(e0));
        if (mode == 0) {
          epi_kernel<0><<<items, 128>>>(ws, c.tiles, c.splits, ld, slab);
          reduce_kernel<<<(unsigned)((slab / 4 + 255) / 256), 256>>>(ws, dw, slab, c.splits, slab);
        } else if (mode == 1) {
          CK(cudaMemsetAsync(dw, 0, slab * 4));
          epi_kernel<1><<<items, 128>>>(dw, c.tiles, c.splits, ld, slab);
        } else if (mode == 2) {
          CK(cudaMemsetAsync(dw, 0, slab * 4));
          epi_bulk_kernel<<<items, 128, 128 * kCols * 4>>>(dw, c.tiles, c.splits, ld);
        } else {
          epi_kernel<0><<<items, 128>>>(ws, c.tiles, c.splits, ld, slab);   // stores alone
        }
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (r >= 0) t[mode] += ms;
      }
    }
    {
      float tc[6] = {0, 0, 0, 0, 0, 0};
      for (int mode = 0; mode < 6; ++mode)
        for (int r = -3; r < reps; ++r) {
          CK(cudaMemsetAsync(flush, r & 1, 256 << 20));
          if (mode >= 3) CK(cudaMemsetAsync(dw, 0, slab * 4));
          CK(cudaEventRecord(e0));
          switch (mode) {
            case 0: epi_coal_kernel<0, 1><<<items, 128>>>(ws, c.tiles, c.splits, ld, slab); break;
            case 1: epi_coal_kernel<0, 4><<<items, 128>>>(ws, c.tiles, c.splits, ld, slab); break;
            case 2: epi_coal_kernel<0, 32><<<items, 128>>>(ws, c.tiles, c.splits, ld, slab); break;
            case 3: epi_coal_kernel<1, 1><<<items, 128>>>(dw, c.tiles, c.splits, ld, slab); break;
            case 4: epi_coal_kernel<1, 4><<<items, 128>>>(dw, c.tiles, c.splits, ld, slab); break;
            case 5: epi_coal_kernel<1, 32><<<items, 128>>>(dw, c.tiles, c.splits, ld, slab); break;
          }
          CK(cudaEventRecord(e1));
          CK(cudaEventSynchronize(e1));
          float ms;
          CK(cudaEventElapsedTime(&ms, e0, e1));
          if (r >= 0) tc[mode] += ms;
        }
      printf("   kernel alone, warp instruction = 1 / 4 / 32 rows: stores %6.2f %6.2f %6.2f us | red.v4 %6.2f %6.2f %6.2f us\n",
             tc[0] / reps * 1e3, tc[1] / reps * 1e3, tc[2] / reps * 1e3, tc[3] / reps * 1e3, tc[4] / reps * 1e3,
             tc[5] / reps * 1e3);
    }
    // check the red result: every element of the first row = sum over splits of (1,2,3,split)
    float h[4];
    CK(cudaMemsetAsync(dw, 0, slab * 4));
    epi_kernel<1><<<items, 128>>>(dw, c.tiles, c.splits, ld, slab);
    CK(cudaMemcpy(h, dw, 16, cudaMemcpyDeviceToHost));
    printf("%-30s partials %6.1f MB | store+reduce %7.2f us | memset+red.v4 %7.2f us | memset+bulk reduce %7.2f us | "
           "stores alone %7.2f us | check %g %g %g %g (expect %d %d %d %d)\n",
           c.name, items * 256.0 / 1024, t[0] / reps * 1e3, t[1] / reps * 1e3, t[2] / reps * 1e3, t[3] / reps * 1e3,
           h[0], h[1], h[2], h[3], c.splits, 2 * c.splits, 3 * c.splits, c.splits * (c.splits - 1) / 2);
  }
  return 0;
}
