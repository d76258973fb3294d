// Micro-benchmark: what does a kernel->kernel edge cost inside a replayed CUDA graph on B200, and how much of it does
// programmatic dependent launch (griddepcontrol.launch_dependents at the top of the primary, griddepcontrol.wait at
// the top of the secondary) give back?  A chain of `n` dependent streaming kernels (each reads what the previous one
// wrote), captured into one graph, three sizes of per-kernel work.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/pdl_probe tools/pdl_probe.cu && build/pdl_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

template <bool PDL>
__global__ void __launch_bounds__(256) stream_kernel(const float4* __restrict__ in, float4* __restrict__ out, long long n) {
  if (PDL) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
  }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float4 v = in[i];
    v.x += 1.f;
    out[i] = v;
  }
}

template <bool PDL>
static void launch(const float4* in, float4* out, long long n, int grid, cudaStream_t st) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(256);
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = PDL ? 1 : 0;
  CK(cudaLaunchKernelEx(&cfg, stream_kernel<PDL>, in, out, n));
}

int main() {
  const int chain = 200;
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  float4 *a, *b;
  const long long maxn = 64LL << 20;   // 1 GiB per buffer
  CK(cudaMalloc(&a, maxn * 16));
  CK(cudaMalloc(&b, maxn * 16));
  CK(cudaMemset(a, 0, maxn * 16));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  const long long sizes[] = {1LL << 12, 1LL << 20, 1LL << 23, 1LL << 25};   // 64 KB, 16 MB, 128 MB, 512 MB per kernel
  for (long long n : sizes) {
    float res[2];
    for (int pdl = 0; pdl < 2; ++pdl) {
      cudaGraph_t g;
      cudaGraphExec_t ge;
      CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
      for (int i = 0; i < chain; ++i) {
        const float4* in = (i & 1) ? b : a;
        float4* out = (i & 1) ? a : b;
        if (pdl) launch<true>(in, out, n, 296, st);
        else launch<false>(in, out, n, 296, st);
      }
      CK(cudaStreamEndCapture(st, &g));
      CK(cudaGraphInstantiate(&ge, g, 0));
      for (int w = 0; w < 3; ++w) CK(cudaGraphLaunch(ge, st));
      CK(cudaStreamSynchronize(st));
      CK(cudaEventRecord(e0, st));
      for (int r = 0; r < 5; ++r) CK(cudaGraphLaunch(ge, st));
      CK(cudaEventRecord(e1, st));
      CK(cudaEventSynchronize(e1));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      res[pdl] = ms * 1e3f / (5 * chain);
      CK(cudaGraphExecDestroy(ge));
      CK(cudaGraphDestroy(g));
    }
    // check: after the warm-up + timed replays the chain must still be a chain (value = number of kernels run)
    float h;
    CK(cudaMemcpy(&h, a, 4, cudaMemcpyDeviceToHost));
    printf("%8.2f MB per kernel: plain edge %7.2f us/kernel | programmatic edge %7.2f us/kernel | saved %5.2f us | a[0].x=%g\n",
           n * 16.0 / (1 << 20), res[0], res[1], res[0] - res[1], h);
  }
  return 0;
}
