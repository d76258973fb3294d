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

#define B2_MAX_WORLD 8
#define B2_XGPU_CHANNELS 256

/* Peer-memory context handed to kernels that talk to other GPUs (built by b2_symm_* below).
 * data[r]/flags[r] are this process's mappings of rank r's symmetric window and signal pad;
 * epoch is local device memory holding one counter per barrier channel.  world <= 1 or a NULL
 * pointer means "single GPU" and no peer memory is touched. */
typedef struct b2_xgpu_ctx {
  int rank, world;
  void* data[B2_MAX_WORLD];
  void* flags[B2_MAX_WORLD];
  void* epoch;
} b2_xgpu_ctx;

/* ------------------------------------------------------------------------------------------------
 * Convolution as implicit GEMM on tcgen05/TMEM (csrc/conv_igemm.cu).
 * Reference: resnet/main.py:76,121,123 (torchvision resnet -> cuDNN fprop / dgrad / wgrad).
 * x: NHWC bf16 [N,H,W,Cin]; w: OHWI bf16 [Cout,KH,KW,Cin]; y: NHWC bf16 [N,Ho,Wo,Cout]; dw: OHWI fp32.
 * Cin must be a multiple of 64 and Cout of 8 (the classifier runs as a 1x1 conv with Cout=1000; the 3-channel
 * stem has its own kernels, b2_stem_fprop / b2_stem_wgrad, or, for outputs wider than 112 pixels, goes through
 * b2_stem_im2col and then these same kernels with Cin=192).
 * ---------------------------------------------------------------------------------------------- */
int b2_conv_fprop(const void* x, const void* w, const float* bias /*nullable, fp32[Cout]*/, void* y, int N, int H,
                  int W, int Cin, int Cout, int KH, int KW, int stride, int pad, void* stream);
/* fprop that also leaves the BatchNorm statistics of y behind (fused into the persistent kernel:
 * eight extra warps read the bf16 output panels back from shared memory while the epilogue warps convert the next one).
 * stats receives [*stats_blocks][2][Cout] fp32 partial (sum, sum of squares) rows, to be handed to
 * b2_bn_fwd_train_partial; *stats_blocks == 0 means the active kernel variant did not produce them. stats must hold
 * 4 * (number of SMs) * 2 * Cout floats. */
int b2_conv_fprop_stats(const void* x, const void* w, void* y, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                        int stride, int pad, float* stats, int* stats_blocks, void* stream);
void b2_conv_set_fuse_stats(int on);
/* timing experiments only (results are wrong): 1 = skip operand loads once the smem ring is primed, 2 = skip MMAs */
void b2_conv_set_debug(int bits);
/* 1 (default): MN-major operands (dgrad weights, wgrad dy / 1x1 x) are fetched as one multi-panel TMA box */
void b2_conv_set_panels(int on);
/* 1 (default): wgrad with Cout % 256 == 0 computes one 256 x 256 tile per CTA (all 512 TMEM columns) */
void b2_conv_set_wgrad_mt2(int on);
/* split-K target of the 128-row wgrad tiles: work items per SM (1 = default: fewer, longer splits; 2) */
void b2_conv_set_wgrad_waves(int n);
/* 1 (default): the split-K sums of wgrad form in L2 - every split adds its tile into dw with 16-byte fp32 reductions,
 * no slab round trip through the workspace and no second kernel; the order of those additions is not fixed, so two
 * runs can differ in the last bits. 0: slabs + a fixed-order reduce kernel everywhere (run-to-run identical). */
void b2_conv_set_wgrad_red(int on);
/* 1 (default): plain 1x1 GEMMs with 256-wide tiles and K >= 256 (no BatchNorm statistics, no bias: the evaluation
 * forward and the dgrads that feed no BatchNorm sums) run on the CTA-pair kernel (tcgen05 cta_group::2, each CTA of a
 * 2-cluster stages half of the B tile); 0: persistent single-CTA kernel everywhere. */
void b2_conv_set_pairs_auto(int on);
int b2_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                  int stride, int pad, void* stream);
/* dgrad that also produces the partial sums of the BatchNorm backward whose output gradient it computes (statistics
 * warps of the persistent kernel): partial receives [*blocks][2][Cin] rows of (sum g, sum g*(y-mean)) with
 * g = mask ? dx : 0; hand them to b2_bn_bwd_partial. *blocks == 0: not produced, run b2_bn_bwd as usual.
 * partial must hold 2 * (number of SMs) * 2 * Cin floats. */
int b2_conv_dgrad_bnstats(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int KH,
                          int KW, int stride, int pad, const void* bn_y, const void* bn_mask, const float* bn_mean,
                          float* partial, int* blocks, void* stream);
/* dgrad with the residual fan-in fused: dx = mask ? (dgrad + add) : 0 (add: bf16 NHWC like dx; mask: 1 bit per
 * element as written by b2_bn_fwd_train, nullable), done by the statistics warps on the output panel in shared memory,
 * which also produce the BatchNorm-backward sums of dx against bn_y / bn_mean (as b2_conv_dgrad_bnstats; dx is already
 * masked, so hand it to b2_bn_bwd_partial with relu = 0). Returns -1 without launching anything when the active
 * kernel variant cannot do it. */
int b2_conv_dgrad_fused(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin, int Cout, int KH,
                        int KW, int stride, int pad, const void* add, const void* mask, const void* bn_y,
                        const float* bn_mean, float* partial, int* blocks, void* stream);
/* workspace: fp32 scratch for split-K partials, at least Cout*KH*KW*Cin*4 bytes; larger lets more SMs work.
 * accumulate bit 0: add into dw (gradient accumulation / no_sync), otherwise dw is overwritten.
 * accumulate bit 1: this call must sum its splits in a fixed order (see b2_conv_set_wgrad_red). */
int b2_conv_wgrad(const void* x, const void* dy, float* dw, float* workspace, long long ws_bytes, int N, int H, int W,
                  int Cin, int Cout, int KH, int KW, int stride, int pad, int accumulate, void* stream);
int b2_conv_wgrad_splits(int N, int Ho, int Wo, int Cin, int Cout, int taps, long long ws_bytes);
/* 3x3 / stride 1 / pad 1 with 64 or 128 reduction channels (csrc/conv3x3.cu): weight-stationary kernel - the nine taps
 * of a CTA's 64 output columns stay in shared memory, every tile's input rows (+ halo) are loaded ONCE and the taps are
 * nine tcgen05.mma passes over shifted windows of that slab. b2_conv_fprop[_stats] / b2_conv_dgrad[_bnstats] route the
 * shapes b2_conv_halo_supported(N, H, W, K = reduction channels, Ncols = output columns) accepts through it
 * (b2_conv_set_halo(0) switches that off). dgrad != 0: src = dy, dst = dx. stats: one partial row per CTA.
 * Reference: resnet/main.py:76 (conv2 of every torchvision BasicBlock / Bottleneck in layer1 and layer2 -> cuDNN). */
int b2_conv_halo_supported(int N, int H, int W, int K, int Ncols);
int b2_conv3x3_halo(int dgrad, const void* src, const void* w, void* dst, int N, int H, int W, int Cin, int Cout,
                    float* stats, int* stats_blocks, const void* bwd_y, const void* bwd_mask, const float* bwd_mean,
                    void* stream);
void b2_conv_set_halo(int on);
void b2_conv_set_halo_debug(int bits);
/* Debug/tuning: force the N tile (64/128/256); 0 = automatic. */
void b2_conv_set_tile_n(int bn);
/* Debug/tuning: staging of k>1 / strided activations: -1 auto (TMA im2col), 1 = cp.async gather. */
void b2_conv_set_act_mode(int mode);
/* Debug/tuning: 1 = one tile per CTA, 2 = persistent kernel with double-buffered TMEM accumulators (default). */
void b2_conv_set_impl(int impl);

/* ------------------------------------------------------------------------------------------------
 * BatchNorm (+ReLU, +residual add) forward/backward on NHWC bf16, fp32 statistics (csrc/bn.cu).
 * Reference: torchvision BasicBlock/Bottleneck bn/relu/add reached from resnet/main.py:76,121,123
 * (aten::native_batch_norm[_backward], relu_, add_, threshold_backward); with xg->world > 1 the statistics
 * are reduced across GPUs inside the finalize kernel (torch.nn.SyncBatchNorm's all_gather/all_reduce): every rank
 * publishes its sums at float offset xg_slot of its window, ONE device barrier, every rank adds all ranks' sums in
 * rank order. The exchange area (2*C + C/8 floats) must ALTERNATE between two offsets from one cross-GPU call to the
 * next (forward and backward calls alike), which is what makes a second, trailing barrier unnecessary.
 * y: conv output [M,C]; z: block output; partial: scratch of b2_bn_partial_floats(C) floats.
 * ---------------------------------------------------------------------------------------------- */
long long b2_bn_partial_floats(int C);
/* Tuning: load batching of the backward reduce pass. 0 = plain unrolled loop, anything else (default) = 4 row groups
 * of loads issued back to back, 2 CTAs/SM. */
void b2_bn_set_reduce_variant(int v);
/* 1 (default): backward apply pass with per-thread channel ownership and batched loads; 0: grid-stride kernels. */
void b2_bn_set_apply_batched(int on);
/* relu_mask (optional, u8 [M*C/8]): one bit per output element, set where the ReLU passed; the backward pass can
 * then be given the mask instead of z (16x fewer bytes). */
int b2_bn_fwd_train(const void* y, const void* res, void* z, void* relu_mask, long long M, int C,
                    const float* gamma, const float* beta, float* running_mean, float* running_var, long long* num_batches_tracked,
                    float momentum, float eps, int relu, float* mean, float* invstd, float* scale_shift,
                    float* partial, const b2_xgpu_ctx* xg, int xg_slot, void* stream);
/* Same, but the statistics partials come from b2_conv_fprop_stats (partial_blocks rows of [2][C]). */
int b2_bn_fwd_train_partial(const void* y, const void* res, void* z, void* relu_mask, long long M, int C,
                            const float* gamma, const float* beta, float* running_mean, float* running_var,
                            long long* num_batches_tracked, float momentum, float eps, int relu, float* mean,
                            float* invstd, float* scale_shift, float* partial, int partial_blocks,
                            const b2_xgpu_ctx* xg, int xg_slot, void* stream);
int b2_bn_fwd_eval(const void* y, const void* res, void* z, long long M, int C, const float* gamma, const float* beta,
                   const float* running_mean, const float* running_var, float eps, int relu, float* scale_shift,
                   void* stream);
/* Same operation as b2_bn_fwd_train for callers that cannot allocate device memory (a serving loop that cannot
 * afford cudaMalloc's device-wide synchronisation, a host application that owns the memory budget and forbids
 * side allocations, a GPU so full that an allocation could fail mid-step): the partial-sum scratch is a static device array inside the library instead of a
 * caller-provided buffer, so the call performs no cudaMalloc / cudaFree and no allocating library routine.
 * Inputs and outputs are still the caller's.  This entry point has no counterpart in the reference.
 *   - Limit: C <= B2_BN_NOALLOC_MAX_CHANNELS; larger C is rejected with cudaErrorInvalidValue and a message on
 *     stderr.  The array (592 * 2 * 2048 floats, about 9.25 MiB) is held by every GPU that loads the library,
 *     until the process exits.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress with that device
 *     current); make that first call before starting any stream capture.
 *   - All calls on a device share the one array, and it keeps whatever the last call wrote. Nothing in it needs
 *     to start at zero (every value that is read is written first within the same call), but two calls whose
 *     kernels overlap on the GPU (different streams) race on it: serialise them. */
#define B2_BN_NOALLOC_MAX_CHANNELS 2048
int b2_bn_fwd_train_noalloc(const void* y, const void* res, void* z, long long M, int C, const float* gamma,
                            const float* beta, float* running_mean, float* running_var,
                            long long* num_batches_tracked, float momentum, float eps, int relu, float* mean,
                            float* invstd, float* scale_shift, const b2_xgpu_ctx* xg, int xg_slot, void* stream);
/* dz' = (dz_a + dz_b) * relu_mask(z); dy = BN backward of dz'; dres (optional) receives dz'. */
int b2_bn_bwd(const void* dz_a, const void* dz_b, const void* z, const void* relu_mask, const void* y, void* dy,
              void* dres, long long M, int C, const float* gamma, const float* mean, const float* invstd, float* dgamma, float* dbeta,
              int accumulate, int relu, float* coef, float* partial, const b2_xgpu_ctx* xg, int xg_slot,
              void* stream);
/* b2_bn_bwd with the second incoming gradient given at HALF resolution: dz_b_q [N, ceil(H/2), ceil(W/2), C] stands for
 * the [N, H, W, C] tensor that equals it at the even pixels and is zero elsewhere - the input gradient of a 1x1 /
 * stride 2 downsample convolution (torchvision's `downsample[0]`), which is then a plain GEMM on the half-resolution
 * grid: no zero fill of the full tensor, a quarter of the bytes read here. relu_mask and dres are required. */
int b2_bn_bwd_q(const void* dz_a, const void* dz_b_q, int H, int W, const void* relu_mask, const void* y, void* dy,
                void* dres, long long M, int C, const float* gamma, const float* mean, const float* invstd,
                float* dgamma, float* dbeta, int accumulate, int relu, float* coef, float* partial,
                const b2_xgpu_ctx* xg, int xg_slot, void* stream);
/* Same for a single incoming gradient whose reduce pass already happened: partial holds partial_blocks rows of
 * [2][C] written by b2_conv_dgrad_bnstats. */
int b2_bn_bwd_partial(const void* dz, const void* z, const void* relu_mask, const void* y, void* dy, long long M,
                      int C, const float* gamma, const float* mean, const float* invstd, float* dgamma, float* dbeta,
                      int accumulate, int relu, float* coef, float* partial, int partial_blocks,
                      const b2_xgpu_ctx* xg, int xg_slot, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Pooling, loss, optimizer, stem helpers (csrc/misc.cu).
 * Reference: resnet/main.py:76 (maxpool / avgpool / fc of torchvision resnet), :102 (CrossEntropyLoss),
 * :103 and :124 (optim.SGD momentum 0.9, weight_decay 1e-5, step()).
 * ---------------------------------------------------------------------------------------------- */
/* 3x3 / stride 2 / pad 1. idx (optional, u8 [N,Ho,Wo,C]) records the winning tap for the backward pass. */
int b2_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, void* stream);
/* BatchNorm affine + ReLU + maxpool in one pass over the raw convolution output x: y = maxpool(relu(x*scale+shift))
 * with the intermediate rounded to bf16 exactly as the separate apply kernel stores it (same values, same winners).
 * Windows whose maximum is 0 record tap code 9, so b2_maxpool_bwd scatters nothing for them and its result is the
 * gradient of the BatchNorm output (ReLU backward included). scale / shift: fp32 [C], e.g. the scale_shift vector that
 * b2_bn_fwd_train* / b2_bn_fwd_eval leave behind when called with z == NULL (coefficients only, no apply pass). */
int b2_bn_relu_maxpool_fwd(const void* x, const float* scale, const float* shift, void* y, void* idx, int N, int H,
                           int W, int C, void* stream);
int b2_maxpool_bwd(const void* idx, const void* dy_a, const void* dy_b, void* dx, int N, int H, int W, int C,
                   void* stream);
int b2_avgpool_fwd(const void* x, void* y, int N, int HW, int C, void* stream);
int b2_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, void* stream);
int b2_ce_fwd(const void* logits, const long long* labels, float* lse, float* loss_row, float* loss, int rows,
              int ncls, int ld, void* stream);
int b2_ce_bwd(const void* logits, const long long* labels, const float* lse, const float* gscale, void* dlogits,
              float* dbias, int accumulate_bias, int rows, int ncls, int ld, void* stream);
int b2_colsum_bf16(const void* m, float* out, int rows, int cols, int ld, int accumulate, void* stream);
/* Fused classifier head (csrc/head.cu). Reference: resnet/main.py:121-122 (`ddp_model(inputs)` ends in
 * torchvision's avgpool -> flatten -> fc; `criterion(outputs, labels)` is log_softmax + nll_loss).
 * b2_head_fwd: global average pool of the NHWC bf16 feature map x [N][HW][C] and the fully connected layer in one
 * kernel (clusters of 8 CTAs share the pooled operand through distributed shared memory, tcgen05 MMA with the
 * accumulator in TMEM): logits [N][ld] fp32 = mean_hw(x) . w^T + bias, w [ncls][C] bf16; the bf16 pooled matrix
 * [N][C] is also written out (the weight-gradient GEMM of the backward pass reads it).
 * b2_head_supported: C a power-of-two multiple of 64 up to 2048, ncls <= 1024.
 * b2_ce_fwd_f32: per-row log-sum-exp (lse) and the mean NLL loss from fp32 logits, one launch; counter is one
 * zero-initialised u32 of device scratch (the kernel leaves it zero).
 * b2_ce_bwd_fused: dlogits (bf16, row stride ld_out) = (softmax - onehot) * (*gscale) / rows and the bias gradient
 * dbias[ncls] (written, or added to when accumulate_bias) in one launch. */
int b2_head_supported(int C, int ncls);
int b2_head_fwd(const void* x, const void* w_bf16, const float* bias, void* pooled, float* logits, int N, int HW,
                int C, int ncls, int ld, void* stream);
int b2_ce_fwd_f32(const float* logits, const long long* labels, float* lse, float* loss_row, float* loss,
                  unsigned int* counter, int rows, int ncls, int ld, void* stream);
int b2_ce_bwd_fused(const float* logits, const long long* labels, const float* lse, const float* gscale,
                    void* dlogits_bf16, float* dbias, int accumulate_bias, int rows, int ncls, int ld, int ld_out,
                    void* stream);
/* p, g, buf: flat fp32; shadow: optional bf16 copy of p refreshed in the same pass.
 * hyper_dev (nullable): three floats in device memory {lr, momentum, weight_decay} read by the kernel instead of the
 * by-value arguments, so a learning-rate schedule does not change the launch (and a captured CUDA graph stays valid). */
int b2_sgd_step(float* p, const float* g, float* buf, void* shadow_bf16, long long n, float lr, float momentum,
                float weight_decay, float grad_scale, int first_step, const float* hyper_dev, void* stream);
int b2_cast_f32_bf16(const float* src, void* dst, long long n, void* stream);
/* Tracing inside CUDA-graph replays (per-bucket ready / launch / done timestamps, exposed communication):
 * a one-thread marker kernel that appends the GPU's %globaltimer (ns) to a device ring. ring[0] counts the stamps,
 * ring[1 + i % capacity] holds stamp i; the ring must hold capacity + 1 u64 and start zeroed. Stamps on different
 * streams of one GPU share a clock, so "comm stream done" minus "compute stream done" is the exposed communication
 * of a step whether the step was launched from Python or replayed from a graph. No reference counterpart (the
 * reference has no timers; torch's Reducer samples similar times every 100 iterations). */
int b2_timestamp(unsigned long long* ring, int capacity, void* stream);
/* *slot = value, in stream order (a one-thread kernel, so it can sit between graph replays): lets a captured graph
 * read its input batch through a pointer that is re-aimed before every replay (b2_stem_pack_input_indirect). */
int b2_set_ptr_slot(void* slot, const void* value, void* stream);
int b2_stem_im2col(const void* x_nchw, int x_is_bf16, void* A, int N, int H, int W, void* stream);
/* Direct stem (csrc/stem.cu): the 7x7/2 convolution on the tensor cores straight from a zero-padded NHWC4 bf16 copy
 * of the batch (overlapping-window TMA boxes), no im2col matrix in memory. b2_stem_supported: output width <= 112.
 * xp holds b2_stem_padded_elems(N,H,W) bf16, written entirely (padding included) by b2_stem_pack_input;
 * wp is bf16 [64][224]; stats / stats_blocks as in b2_conv_fprop_stats (2 rows per CTA);
 * workspace holds b2_stem_wgrad_workspace_floats() floats; dw is fp32 OHWI [64][7][7][3]. */
int b2_stem_supported(int H, int W);
long long b2_stem_padded_elems(int N, int H, int W);
int b2_stem_pack_input(const void* x_nchw, int x_is_bf16, void* xp, int N, int H, int W, void* stream);
/* Same as b2_stem_pack_input, but the batch address is read from device memory (*x_slot) when the kernel runs. */
int b2_stem_pack_input_indirect(const void* const* x_slot, int x_is_bf16, void* xp, int N, int H, int W,
                                void* stream);
int b2_stem_pack_weight(const float* w_ohwi, void* wp, void* stream);
int b2_stem_fprop(const void* xp, const void* wp, void* y, int N, int H, int W, float* stats, int* stats_blocks,
                  void* stream);
long long b2_stem_wgrad_workspace_floats(void);
int b2_stem_wgrad(const void* xp, const void* dy, float* dw_ohwi, float* workspace, int N, int H, int W,
                  int accumulate, void* stream);
int b2_stem_pack_w(const float* w_ohwi, void* wp_bf16, int Cout, void* stream);
int b2_stem_unpack_dw(const float* dwp, float* dw_ohwi, int Cout, int accumulate, void* stream);
int b2_nchw_to_nhwc_bf16(const void* x, int x_is_bf16, void* y, int N, int C, int HW, void* stream);

/* ------------------------------------------------------------------------------------------------
 * DDP reducer core: bucket planner + ready/launch state machine (csrc/reducer.cpp, host only).
 * Reference: resnet/main.py:80 (DistributedDataParallel -> c10d::Reducer, compute_bucket_assignment_by_size).
 * ---------------------------------------------------------------------------------------------- */
typedef struct b2_reducer b2_reducer;
int b2_plan_buckets(const long long* sizes_bytes, int n, long long first_cap_bytes, long long cap_bytes,
                    int* bucket_of_param);
b2_reducer* b2_reducer_create(int n_params, const int* bucket_of_param, int n_buckets);
void b2_reducer_destroy(b2_reducer* r);
void b2_reducer_reset(b2_reducer* r);
int b2_reducer_mark_ready(b2_reducer* r, int idx, int* launch);
int b2_reducer_unlaunched(const b2_reducer* r);
long long b2_reducer_iterations(const b2_reducer* r);
int b2_reducer_ready_order(const b2_reducer* r, int* out);
/* Native launch path: a bucket's all-reduce launch descriptor (what b2_allreduce / b2_allreduce_bf16 need), registered
 * once per bucket (again after the momentum buffers come into existence or the buckets are rebuilt). With descriptors
 * in place b2_reducer_mark_ready_launch marks a parameter ready AND launches the fused reduction of every bucket that
 * became launchable (event on the compute stream -> wait on the communication stream -> kernel), writing the launched
 * bucket indices to `launched`; b2_reducer_join makes the compute stream wait for the reductions (finalize_backward).
 * Reference: c10d::Reducer mark_variable_ready / all_reduce_bucket / finalize_backward behind resnet/main.py:80,123. */
typedef struct b2_bucket_desc {
  b2_xgpu_ctx ctx;
  void* mc_base;
  long long offset_bytes; /* fp32 wire: the bucket inside the window */
  long long n;            /* fp32 elements */
  int algo;
  float scale;
  float* p32;             /* fused update (used when the call says fused) */
  float* mom;
  void* shadow;
  int write_grad, nblocks;
  int wire_bf16;          /* 1: b2_allreduce_bf16 with the two fields below */
  long long wire_offset_bytes;
  float* g32;
} b2_bucket_desc;
int b2_reducer_set_bucket(b2_reducer* r, int bucket, const b2_bucket_desc* desc);
int b2_reducer_mark_ready_launch(b2_reducer* r, int idx, int fused, float lr, float momentum, float weight_decay,
                                 int first_step, const float* hyper_dev, void* compute_stream, void* comm_stream,
                                 int* launched);
int b2_reducer_join(b2_reducer* r, void* compute_stream, void* comm_stream);
void b2_reducer_release_launch_state(b2_reducer* r);

/* ------------------------------------------------------------------------------------------------
 * Symmetric-memory runtime and watchdog (csrc/symm.cu).
 * Reference: resnet/main.py:74 init_process_group("nccl") -> ProcessGroupNCCL communicator, buffer
 * registration and watchdog thread.
 * ---------------------------------------------------------------------------------------------- */
int b2_ipc_alloc(long long bytes, void** dptr, unsigned char* handle64);
int b2_ipc_open(const unsigned char* handle64, void** dptr);
int b2_ipc_close(void* dptr);
int b2_ipc_free(void* dptr);
long long b2_vmm_round(long long bytes, int device, int world);
int b2_vmm_multicast_supported(int device);
int b2_vmm_alloc(long long bytes, int device, void** dptr, unsigned long long* handle, int* fd);
int b2_vmm_import(int fd, long long bytes, int device, void** dptr);
int b2_mc_create(long long bytes, int world, unsigned long long* mc_handle, int* fd);
int b2_mc_import(int fd, unsigned long long* mc_handle);
int b2_mc_add_device(unsigned long long mc_handle, int device);
int b2_mc_bind_and_map(unsigned long long mc_handle, unsigned long long mem_handle, long long bytes, int device,
                       void** mc_ptr);
/* Teardown (destroy_process_group / SymmetricWindow.close()): unbind from the multicast object, unmap every mapping
 * (own window, peers' windows, multicast address) and free its address range, release every handle. After the last
 * rank has done so the physical memory and the multicast object are gone. b2_vmm_import_h is b2_vmm_import that also
 * returns the imported handle so that it can be released. */
int b2_vmm_import_h(int fd, long long bytes, int device, void** dptr, unsigned long long* handle);
int b2_vmm_unmap(void* dptr, long long bytes);
int b2_vmm_release(unsigned long long handle);
int b2_mc_unbind(unsigned long long mc_handle, int device, long long bytes);
int b2_fd_listen(const char* name);
int b2_fd_serve(int listen_sock, const int* fds, int nfds, int count);
int b2_fd_fetch(const char* name, int which, int timeout_ms);
void b2_fd_close(int fd);
int b2_watchdog_start(int timeout_s, int device, int abort_on_fire);
void b2_watchdog_watch(void* cuda_event);
int b2_watchdog_fired(void);
void b2_watchdog_stop(void);

/* ------------------------------------------------------------------------------------------------
 * Collectives over peer memory, fused with scale + SGD (csrc/allreduce.cu). Never call NCCL.
 * Reference: resnet/main.py:80 (DDP bucket all-reduce, parameter/buffer broadcast), :124 (optimizer.step()).
 * offset_bytes locates the bucket inside every rank's window (ctx->data[r]); n = fp32 element count (mult. of 4).
 * p32 == NULL: reduce only. write_grad: leave the averaged gradient in this rank's bucket (p.grad parity).
 * ---------------------------------------------------------------------------------------------- */
enum { B2_AR_AUTO = 0, B2_AR_ONESHOT = 1, B2_AR_TWOSHOT = 2, B2_AR_MULTIMEM = 3 };
long long b2_allreduce_oneshot_max_elems(int nblocks);
int b2_allreduce(const b2_xgpu_ctx* ctx, void* mc_base, long long offset_bytes, long long n, int algo, float scale,
                 float* p32, float* mom, void* shadow_bf16, float lr, float momentum, float weight_decay,
                 int first_step, const float* hyper_dev /* nullable, see b2_sgd_step */, int write_grad, int nblocks,
                 void* stream);
int b2_broadcast(const b2_xgpu_ctx* ctx, long long offset_bytes, long long nbytes, int src, void* dst_override,
                 int nblocks, void* stream);
/* bf16 wire format (the counterpart of torch's bf16_compress_hook): the kernel casts this rank's fp32 bucket
 * (g32_in, local) into a bf16 staging area at wire_offset_bytes of the window, reduces the peers' staging areas in
 * fp32, rounds the average to bf16 once, and feeds the SGD epilogue / g32_out (optional) from that value. */
int b2_allreduce_bf16(const b2_xgpu_ctx* ctx, void* mc_base, long long wire_offset_bytes, long long n, int algo,
                      float scale, const float* g32_in, float* g32_out, float* p32, float* mom, void* shadow_bf16,
                      float lr, float momentum, float weight_decay, int first_step, const float* hyper_dev,
                      int nblocks, void* stream);
/* Broadcast through the NVSwitch multicast mapping: only the source rank stores, the switch replicates. */
int b2_broadcast_mc(const b2_xgpu_ctx* ctx, void* mc_base, long long offset_bytes, long long nbytes, int src,
                    int nblocks, void* stream);
int b2_xgpu_barrier(const b2_xgpu_ctx* ctx, int channel, void* stream);
/* Device-side waits on peers trap after this long (default 1800 s, the reference stack's process-group timeout:
 * ranks legitimately wait minutes for rank 0's evaluation + checkpoint, resnet/main.py:109-115). One setter per
 * translation unit that spins; parallel/comm.py calls both. */
int b2_set_spin_timeout_ns_allreduce(unsigned long long ns);
int b2_set_spin_timeout_ns_bn(unsigned long long ns);
/* Protocol debugging: a non-zero value makes a pseudo-random subset of the CTAs of every collective
 * kernel sleep that long before and between the phases, so that a kernel which consumes data produced by a
 * differently-numbered CTA of a peer (not ordered by the per-CTA cross-GPU barrier) fails its numerical test. */
int b2_set_debug_skew_ns_allreduce(unsigned long long ns);

#ifdef __cplusplus
}
#endif