/* b200dt — C ABI of the hand-written sm_100a kernels of this library.
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on (a cudaStream_t
 * passed as void*), launches asynchronously, and returns 0 on success, a positive cudaError_t
 * value for a CUDA failure, or a negative value for an argument the kernel cannot handle.
 * Python binds these through ctypes (pytorch_distributed_training_tutorials_b200/_lib.py);
 * the C++ gradient reducer calls them directly.
 *
 * "Replaces" cites the place in the reference tutorial (paths under /root/reference) where the
 * equivalent library op (cuBLAS / cuDNN / ATen / NCCL) is reached.
 */
#ifndef B200DT_H_
#define B200DT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------- tcgen05 GEMM / implicit GEMM
 * Replaces nn.Linear (ddp_gpus.py:81, 01.data_parallel.ipynb:171, 03.model_parallel.ipynb:444-446)
 * and the ResNet-50 convolutions (03.model_parallel.ipynb:807-833). csrc/gemm.cu */

/* C[M,N] = A[M,K] * B[N,K]^T (+ bias[N]) (ReLU); A, B bf16 row-major, C bf16 or fp32.
 * bn_sums (optional, bf16 output only): per-column sum / sum of squares of C accumulated in the
 * epilogue, in the layout b200dt_bn_apply reads (2*N*B200DT_BN_REPLICAS + 1 floats, see below). */
int b200dt_gemm_tn_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda,
                        int ldb, int ldc, const float* bias, int relu, int out_fp32, float* bn_sums,
                        void* stream);
/* C[M,N] = A[M,K] * B[K,N]; row-major bf16 inputs (Linear dgrad without a transposed weight). */
int b200dt_gemm_nn_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                        int ldc, int out_fp32, void* stream);
/* C[M,N] fp32 = A[Kr,M]^T * B[Kr,N] (weight gradients). splits = 0: library picks and clears C. */
int b200dt_gemm_nt_reduce_bf16(const void* A, const void* B, float* C, int M, int N, int Kr,
                               int lda, int ldb, int ldc, int splits, void* stream);
/* NHWC conv2d, square filter R in {1,3}, pad R/2, stride in {1,2}, Cin % 64 == 0. */
int b200dt_conv2d_fprop_bf16(const void* x, const void* w, void* y, int NB, int H, int W, int Cin,
                             int Cout, int R, int stride, int relu, float* bn_sums, void* stream);
int b200dt_conv2d_wgrad_bf16(const void* x, const void* dy, float* dw, int NB, int H, int W,
                             int Cin, int Cout, int R, int stride, void* stream);
/* dgrad reads the forward filter w [Cout,R,R,Cin] in place (MN-major B operand). */
int b200dt_conv2d_dgrad_bf16(const void* dy, const void* w, void* dx, int NB, int H, int W,
                             int Cin, int Cout, int R, int stride, void* stream);
/* dx += dgrad(dy, w): the epilogue adds to what dx holds (TMA reduce-add, bf16), e.g. the gradient
 * that reached the same tensor through a residual connection. Returns -4 when the output cannot be
 * written through a tensor map (caller then uses the plain entry point and adds). */
int b200dt_conv2d_dgrad_acc_bf16(const void* dy, const void* w, void* dx, int NB, int H, int W,
                                 int Cin, int Cout, int R, int stride, void* stream);

/* ---------------------------------------------------------------- layout / stem helpers
 * Replaces the cuDNN 7x7 stem convolution's gather (03.model_parallel.ipynb:812). csrc/layout.cu */

/* x: [NB,3,H,W] fp32 NCHW -> cols: [NB*Ho*Wo, ldk] bf16 with (r,s,c) ordering, 7x7 s2 p3. */
int b200dt_stem_im2col(const float* x, void* cols, int NB, int H, int W, int ldk, void* stream);
/* Implicit-GEMM stem (csrc/stem.cu): y[NB,H/2,W/2,64] bf16 NHWC = conv 7x7 s2 p3 of x[NB,3,H,W] fp32 NCHW with
 * w[64,7,7,3] fp32; the patch matrix only ever exists tile by tile in shared memory. bn_sums as for the other
 * convolutions (64 channels). Needs (W/2) % 16 == 0 and (H/2) % 8 == 0, else -3 (use im2col + GEMM). */
int b200dt_stem_conv_fprop(const float* x, const float* w, void* y, int NB, int H, int W, float* bn_sums,
                           void* stream);
/* dw[64,7,7,3] fp32 from dy[NB,H/2,W/2,64] bf16; scratch: b200dt_stem_wgrad_scratch_floats() floats (one partial
 * per CTA, folded in CTA order: the result is deterministic in every mode). */
long long b200dt_stem_wgrad_scratch_floats(void);
int b200dt_stem_conv_wgrad(const float* x, const void* dy, float* dw, float* scratch, int NB, int H, int W,
                           void* stream);
int b200dt_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream);
int b200dt_transpose_bf16(const void* src, void* dst, int rows, int cols, void* stream);
int b200dt_nchw_f32_to_nhwc_bf16(const float* x, void* y, int NB, int C, int H, int W, void* stream);

/* ---------------------------------------------------------------- BatchNorm / ReLU / residual
 * Replaces cuDNN BatchNorm + ATen relu/add in torchvision Bottleneck (03.model_parallel.ipynb:810).
 * csrc/bn.cu.  x, y, residual, dy, dx, dres: [rows, C] bf16 (NHWC flattened); C % 8 == 0. */

/* `sums` holds b200dt_bn_replicas_used(C) copies of {sum x [C], sum x^2 [C]} (fp32; cleared by the call; CTAs
 * spread their atomics over the copies and the last CTA folds them into copy 0, which is what the apply
 * kernels read) followed by one ticket word.  The copy count falls with C (min(16, 2048 / C), at least 1:
 * wide layers have little contention and the fold costs 2*C*copies loads), so 2*C*B200DT_BN_REPLICAS + 1
 * floats is always enough; 2*C*b200dt_bn_replicas_used(C) + 1 is exact.
 * b200dt_set_sums_prezeroed(1): the caller promises that every `sums` / `bn_sums` buffer it passes is already
 * zero (all copies and the ticket word), and the entry points stop issuing their own memset -- a process-wide
 * switch for callers that carve the buffers from zero-filled memory (ops/native.py does: ~160 memset nodes
 * less per ResNet-50 step).  b200dt_capture_id: id of the stream capture `stream` is in, 0 if none. */
#define B200DT_BN_REPLICAS 16
int b200dt_bn_replicas_used(int C);
void b200dt_set_sums_prezeroed(int on);
int b200dt_get_sums_prezeroed(void);
unsigned long long b200dt_capture_id(void* stream);
/* Deterministic mode (env B200DT_DETERMINISTIC=1 or b200dt_set_deterministic(1); no counterpart call in
 * the reference, it is a regression-testing aid): every fp32 sum that is otherwise combined with atomics
 * (BatchNorm statistics, BatchNorm-backward partials, split-K weight gradients, bias gradients, the SIMT
 * Linear's slices) is combined in a fixed order, so identical launches give identical bits. `sums` buffers
 * then hold B200DT_BN_REPLICAS_DET copies (one per CTA slot): size them with b200dt_bn_replicas(). Split-K
 * weight gradients take a stream-ordered scratch (cudaMallocAsync / cudaFreeAsync on the launch stream). */
#define B200DT_BN_REPLICAS_DET 160
void b200dt_set_deterministic(int on);
int b200dt_get_deterministic(void);
int b200dt_bn_replicas(void);
int b200dt_bn_stats(const void* x, float* sums, long long rows, int C, void* stream);
/* fp32 SIMT Linear for feature counts the tensor-core path cannot take (K or N not a multiple of 8;
 * the tutorial's Linear(20,1), (32,2), (10000,10), (10,5)). y[B,N] = x[B,K] * w[N,K]^T + bias (bias may be
 * NULL); bwd writes any of dx[B,K], dw[N,K], db[N] that is not NULL. The *_host variants run the same
 * element code serially on the CPU (host pointers): they exist so the indexing is testable without a GPU. */
int b200dt_skinny_linear_fwd(const float* x, const float* w, const float* bias, float* y, int B, int K,
                             int N, void* stream);
int b200dt_skinny_linear_bwd(const float* dy, const float* x, const float* w, float* dx, float* dw,
                             float* db, int B, int K, int N, void* stream);
int b200dt_skinny_linear_fwd_host(const float* x, const float* w, const float* bias, float* y, int B,
                                  int K, int N);
int b200dt_skinny_linear_bwd_host(const float* dy, const float* x, const float* w, float* dx, float* dw,
                                  float* db, int B, int K, int N);
/* Backward of y = relu(linear): g = dy * (y > 0) (bf16, y == g == NULL skips the mask) and
 * db[C] = column sums of g in fp32 (db == NULL skips them); one pass over [rows, C], C % 8 == 0. */
int b200dt_relu_bwd_bias(const void* dy, const void* y, void* g, float* db, long long rows, int C,
                         void* stream);
/* y = act(gamma * (x - mean) * invstd + beta (+ residual)); writes mean/invstd (fp32 [C] each)
 * and updates running stats with `momentum` (unbiased variance), all from `sums`.
 * batches_tracked (optional, device int64): incremented by one, like nn.BatchNorm's counter.
 * relu_mask (optional, rows*C/8 bytes, relu only): bit j of byte [row][c/8] = output element > 0. */
int b200dt_bn_apply(const void* x, const void* residual, void* y, const float* sums,
                    const float* gamma, const float* beta, float* running_mean, float* running_var,
                    float* save_mean, float* save_invstd, long long rows, int C, float eps,
                    float momentum, int relu, long long* batches_tracked, unsigned char* relu_mask,
                    void* stream);
/* Inference-style apply with given statistics. */
int b200dt_bn_apply_eval(const void* x, const void* residual, void* y, const float* gamma,
                         const float* beta, const float* mean, const float* var, long long rows,
                         int C, float eps, int relu, void* stream);
/* Backward phase 1: g = dy * (y > 0 if relu); sums[0..C) = sum g, sums[C..2C) = sum g * xhat;
 * optionally writes g to dres (gradient of the residual branch). */
/* With relu != 0 and y == NULL the mask is recomputed from x, gamma, beta (no residual case).
 * relu == 2: y points at the bit mask written by b200dt_bn_apply (dres is required). */
int b200dt_bn_bwd_reduce(const void* dy, const void* y, const void* x, const float* save_mean,
                         const float* save_invstd, const float* gamma, const float* beta, float* sums,
                         void* dres, long long rows, int C, int relu, void* stream);
/* Backward phase 2: dx = gamma*invstd*(g - sum_g/rows - xhat*sum_gx/rows); dgamma, dbeta fp32. */
int b200dt_bn_bwd_apply(const void* dy, const void* y, const void* x, const float* save_mean,
                        const float* save_invstd, const float* gamma, const float* beta,
                        const float* sums, void* dx, float* dgamma, float* dbeta, long long rows, int C,
                        int relu, void* stream);

/* ---------------------------------------------------------------- pooling (csrc/pool.cu)
 * Replaces nn.MaxPool2d(3,2,1) / AdaptiveAvgPool2d(1) (03.model_parallel.ipynb:815,824). */
int b200dt_maxpool3x3s2_fwd(const void* x, void* y, uint8_t* argmax, int NB, int H, int W, int C,
                            void* stream);
int b200dt_maxpool3x3s2_bwd(const void* dy, const uint8_t* argmax, void* dx, int NB, int H, int W,
                            int C, void* stream);
int b200dt_avgpool_fwd(const void* x, void* y, int NB, int HW, int C, void* stream);
int b200dt_avgpool_bwd(const void* dy, void* dx, int NB, int HW, int C, void* stream);

/* ---------------------------------------------------------------- losses (csrc/loss.cu)
 * Replaces F.cross_entropy (ddp_gpus.py:37), nn.MSELoss (03.model_parallel.ipynb:533,976) and
 * output.sum() (01.data_parallel.ipynb:485). Each call computes loss and d(loss)/d(pred). */

/* Scratch for the two-pass reductions below: at least b200dt_loss_scratch_bytes() bytes. */
long long b200dt_loss_scratch_bytes(void);
/* loss = mean((pred - target)^2); grad = 2*(pred - target)/n * grad_scale. pred/target fp32 or
 * bf16 (is_bf16), grad has the dtype of pred, loss fp32[1]. grad may be NULL. */
int b200dt_mse_loss(const void* pred, const void* target, void* grad, float* loss, long long n,
                    float grad_scale, int is_bf16, void* scratch, void* stream);
/* loss = sum(x); grad = grad_scale everywhere (may be NULL). */
int b200dt_sum_loss(const void* x, void* grad, float* loss, long long n, float grad_scale,
                    int is_bf16, void* scratch, void* stream);
/* Soft-target cross entropy, mean over rows: logits [B,C], probs [B,C] fp32 targets. */
int b200dt_ce_prob_loss(const void* logits, const float* probs, void* grad, float* loss, int B,
                        int C, float grad_scale, int is_bf16, void* scratch, void* stream);
/* Class-index cross entropy, mean over rows: labels int64[B]. */
int b200dt_ce_index_loss(const void* logits, const long long* labels, void* grad, float* loss,
                         int B, int C, float grad_scale, int is_bf16, void* scratch, void* stream);

/* MSE loss for callers that cannot allocate device memory (no counterpart in the reference, which
 * never manages memory itself). Same result as b200dt_mse_loss, but the reduction scratch is a
 * fixed B200DT_MSE_STATIC_SCRATCH_BYTES array inside the library's device image, so the call does
 * no cudaMalloc / cudaFree and needs no scratch argument; inputs and outputs are still the
 * caller's. Meant for hosts with a pre-carved memory arena, plugins inside a runtime that owns
 * the heap, or phases where allocating is forbidden.
 *  - n must be <= B200DT_MSE_STATIC_MAX_ELEMS (one partial-sum slot per 4096 elements, 256
 *    slots); larger inputs return -9 and launch nothing.
 *  - The first call on each device looks up the device address of the array
 *    (cudaGetSymbolAddress, with that device current); make that first call before starting any
 *    stream capture.
 *  - All calls on a device share the one array: two calls whose kernels may overlap on the GPU
 *    (different streams without ordering) race on it. Order them, or use b200dt_mse_loss with
 *    separate scratch buffers. The array lives on every GPU that loads the library until exit.
 *  - `stream` must belong to the device that is current when the call is made: the array used is the
 *    current device's copy. The cached address does not survive cudaDeviceReset(). */
#define B200DT_MSE_STATIC_SCRATCH_BYTES 1040
#define B200DT_MSE_STATIC_MAX_ELEMS 1048576LL
int b200dt_mse_loss_static(const void* pred, const void* target, void* grad, float* loss,
                           long long n, float grad_scale, int is_bf16, void* stream);

/* ---------------------------------------------------------------- optimizers (csrc/optim.cu)
 * Replaces torch.optim.SGD (ddp_gpus.py:82) / Adam (01.data_parallel.ipynb:287) foreach kernels.
 * meta: device int64 [n_tensors, 8] = {param, grad, shadow_bf16|0, numel, exp_avg|0, exp_avg_sq|0,
 * 0, 0}; chunks: device int32 [n_chunks, 2] = {tensor index, first element}; every chunk covers
 * at most B200DT_OPT_CHUNK elements. */
#define B200DT_OPT_CHUNK 8192
int b200dt_multi_sgd(const long long* meta, const int* chunks, int n_chunks, float lr,
                     float weight_decay, float grad_scale, int zero_grad, void* stream);
int b200dt_multi_adam(const long long* meta, const int* chunks, int n_chunks, float lr, float beta1,
                      float beta2, float eps, float weight_decay, int step, float grad_scale,
                      int zero_grad, void* stream);
/* Same update with the step count on the device (torch's Adam(capturable=True)): step_state = 3 floats
 * {step, 1 - beta1^step, 1/sqrt(1 - beta2^step)}; the call increments step first. Capturable in a CUDA graph,
 * which the reference's DataParallel loop (01.data_parallel.ipynb:287,478-487) needs to be replayed from one. */
int b200dt_multi_adam_capturable(const long long* meta, const int* chunks, int n_chunks, float lr, float beta1,
                                 float beta2, float eps, float weight_decay, float* step_state,
                                 float grad_scale, int zero_grad, void* stream);

/* ---------------------------------------------------------------- multi-process collectives
 * Replaces the NCCL all-reduce / broadcast issued by DistributedDataParallel (ddp_gpus.py:32,38)
 * together with the Reducer's copy-in / copy-out. csrc/allreduce.cu.
 *
 * One launch = gather the listed tensors (scaled, cast to the wire dtype) into this rank's
 * symmetric staging buffer -> cross-GPU barrier -> reduce over peer memory (mode 0 one-shot,
 * 1 two-shot, 2 NVLS multimem) or broadcast from `root` (mode 3) -> scatter the result back into
 * the tensors. `bufs` / `sigs` are HOST arrays of `world` peer-mapped device addresses: every
 * rank's symmetric buffer and its signal pad (uint32 words, zero-initialised, at least
 * B200DT_SIGNAL_WORDS of them). `mc` is the multicast address of the symmetric buffer (0 if none;
 * required for mode 2). `tensors`: device array of {void* ptr; int64 numel}. `chunks`: device array
 * of {int32 tensor, int32 start, int32 len, int32 pad, int64 stage_off}; stage_off is in elements,
 * a multiple of 8. All ranks must pass identical chunk tables, `grid` and `sig_base`; launches that may
 * run concurrently (different streams) need disjoint [sig_base, sig_base + grid*world) ranges, and the
 * last signal word is reserved (abort flag). `status` (3 x int32, ideally host-mapped memory from
 * b200dt_status_alloc) receives {code, cta << 8 | peer, waited ms} when a flag wait exceeds the time limit
 * (B200DT_COLL_TIMEOUT_S, default 30 s; b200dt_set_coll_timeout_ms); the kernel then raises the abort
 * word in every peer's pad and traps, so no rank ever writes gradients reduced from partial data.
 * in_place != 0 (modes 1 and 2, wire dtype == tensor dtype): the staging region is the home of the
 * tensors ("gradients as bucket views"): a tensor whose address already is its staging slot is not
 * copied, others are copied in unscaled, `scale` is applied to the SUM, and nothing is copied back
 * (the result is read from the staging region). */
#define B200DT_SIGNAL_WORDS 8192
#define B200DT_MAX_WORLD 8
int b200dt_coll_chunk_bytes(void);
void b200dt_set_coll_timeout_ms(long long ms);
int b200dt_coll_launch(int mode, const uint64_t* bufs, const uint64_t* sigs, uint64_t mc, int rank,
                       int world, int root, const void* tensors, const void* chunks, int n_chunks,
                       long long stage_base_bytes, int tensor_bf16, int wire_bf16, float scale,
                       int sig_base, int grid, int* status, int in_place, void* stream);
/* fp32 gradient -> (scaled) bucket slot, optionally as bf16 (NCCL / gloo fallback copy-in). */
int b200dt_scale_cast(const float* src, void* dst, long long n, float scale, int dst_bf16,
                      void* stream);
/* bucket slot (fp32 or bf16) -> fp32 gradient (fallback copy-out). */
int b200dt_cast_to_f32(const void* src, float* dst, long long n, int src_bf16, void* stream);
/* Single-GPU emulation of the reduce arithmetic of `world` ranks: out = sum_r scale * bufs[r];
 * dtype 0 fp32, 1 bf16 (for tests on one GPU). bufs: HOST array of device addresses. */
int b200dt_allreduce_local_emulate(const uint64_t* bufs, int world, void* out, long long n,
                                   int dtype, float scale, void* stream);

/* ---------------------------------------------------------------- symmetric arena (csrc/arena.cu)
 * Replaces the buffer registration / peer mapping ProcessGroupNCCL does behind init_process_group
 * (ddp_gpus.py:16): one cuMemCreate allocation per rank, exported as a POSIX fd, mapped into every peer
 * process, and bound to ONE NVLS multicast object (cuMulticastCreate / AddDevice / BindMem) whose address
 * the multimem kernels use. fds travel between processes out of band (ops/arena.py uses AF_UNIX sockets).
 * Call order: create (all) -> export_fd / map_peer (all pairs) -> mc_create (rank 0) / mc_import (others)
 * -> mc_add_device (all) -> barrier -> mc_bind (all) -> barrier. Returns -7 when a driver call fails
 * (b200dt_arena_last_error gives the CUresult). */
int b200dt_multicast_supported(int dev);
int b200dt_arena_create(int dev, int rank, int world, long long nbytes, void** arena, long long* size_out);
int b200dt_arena_export_fd(void* arena, int* fd);
int b200dt_arena_map_peer(void* arena, int peer, int fd);
int b200dt_arena_mc_create(void* arena, int* fd);
int b200dt_arena_mc_import(void* arena, int fd);
int b200dt_arena_mc_add_device(void* arena);
int b200dt_arena_mc_bind(void* arena);
unsigned long long b200dt_arena_ptr(void* arena, int peer);
unsigned long long b200dt_arena_mc_ptr(void* arena);
int b200dt_arena_last_error(void* arena);
int b200dt_arena_destroy(void* arena);
/* 64 bytes of pinned host memory mapped into the device: the collectives' status words. */
int b200dt_status_alloc(int** host_ptr, unsigned long long* dev_ptr);
int b200dt_status_free(int* host_ptr);

/* ---------------------------------------------------------------- single-process peer kernels
 * Replaces torch._C._broadcast_coalesced / _scatter / _gather / nccl.reduce behind nn.DataParallel
 * (01.data_parallel.ipynb:275,482,486). csrc/p2p.cu. Pointers are peer-mapped device pointers of
 * ONE process; ordering between devices is by CUDA events, not in-kernel flags. */
int b200dt_enable_peer_access(int dev, int peer);
/* src on the current device is stored to up to 8 destinations (peer pointers) in one pass. */
int b200dt_p2p_broadcast(const void* src, const uint64_t* dsts_host, int ndst, long long nbytes,
                         void* stream);
/* dst (current device) = sum of nsrc peer buffers (fp32), scaled. */
int b200dt_p2p_reduce_f32(float* dst, const uint64_t* srcs_host, int nsrc, long long n,
                          float scale, int accumulate, void* stream);
/* n <= 8 independent copies dst[k][0..nbytes[k]) = src[k][...] in ONE launch on the current device; any of
 * the pointers may be peer memory. DataParallel's scatter (root slices -> peers), gather (peers -> slices
 * of the root tensor) and the slice hand-out of its two-level broadcast. Host arrays. */
int b200dt_p2p_multi_copy(const uint64_t* srcs_host, const uint64_t* dsts_host, const long long* nbytes_host,
                          int n, void* stream);
/* Copy rows [row0, row0+rows) x row_bytes from the current device into dst (peer) and, if
 * partial != NULL, also add sum(src) (fp32 data) into *partial: gather fused with the loss sum. */
int b200dt_p2p_gather_rows(const void* src, void* dst_peer, long long nbytes, float* partial_peer,
                           int is_f32, void* stream);

/* ---------------------------------------------------------------- misc */
/* cudaMemcpyAsync(H2D) from caller-owned pinned memory (table uploads; capturable). */
int b200dt_memcpy_h2d_async(void* dst, const void* src_pinned, long long nbytes, void* stream);
int b200dt_device_sm_count(void);
const char* b200dt_error_string(int cuda_error_code);
const char* b200dt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200DT_H_ */
