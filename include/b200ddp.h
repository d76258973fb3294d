/* b200ddp — C ABI of the sm_100a kernel library and native runtime.
 *
 * Every entry point is `extern "C"`, takes raw device pointers plus a `cudaStream_t` (passed as void*),
 * never synchronises the device and returns 0 on success or a cudaError_t / CUresult value.
 * Python binds these with ctypes (pytorch_distributed_tutorials_b200/ops/_lib.py).
 *
 * "Reference" below is /root/reference/resnet/main.py (the whole upstream program); each block names the
 * line(s) of it, or the dependency it resolves to there, that the entry point stands in for.
 */
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------------
 * Convolution as implicit GEMM on tcgen05/TMEM (csrc/conv_igemm.cu).
 * Reference: resnet/main.py:76,121,123 (torchvision resnet -> cuDNN fprop / dgrad / wgrad).
 * x: NHWC bf16 [N,H,W,Cin]; w: OHWI bf16 [Cout,KH,KW,Cin]; y: NHWC bf16 [N,Ho,Wo,Cout]; dw: OHWI fp32.
 * Cin and Cout must be multiples of 64 (the 3-channel stem has its own kernels below).
 * ---------------------------------------------------------------------------------------------- */
int b2_conv_fprop(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                  int stride, int pad, void* stream);
int b2_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                  int stride, int pad, void* stream);
/* workspace: fp32 scratch for split-K partials, at least Cout*KH*KW*Cin*4 bytes; larger lets more SMs work.
 * accumulate != 0 adds into dw (gradient accumulation / no_sync), otherwise dw is overwritten. */
int b2_conv_wgrad(const void* x, const void* dy, float* dw, float* workspace, long long ws_bytes, int N, int H, int W,
                  int Cin, int Cout, int KH, int KW, int stride, int pad, int accumulate, void* stream);
int b2_conv_wgrad_splits(int N, int Ho, int Wo, int Cin, int Cout, int taps, long long ws_bytes);
/* Debug/tuning: force the N tile (64/128/256); 0 = automatic. */
void b2_conv_set_tile_n(int bn);

#ifdef __cplusplus
}
#endif
