/* dtf_b200 -- C ABI of the B200-native kernel library (libdtf_b200_cuda.so).
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on (a cudaStream_t passed as void*),
 * returns 0 on success or a DTF_ERR_* code, and never synchronises the device. Activations are NHWC bf16 with the
 * channel count a multiple of 8; parameters are fp32 masters with bf16 "shadow" copies for the tensor cores.
 * The Python bindings (distributed_tensorflow_framework_b200/ops/_lib.py) parse THIS file to build their ctypes
 * signatures, so a declaration here is the single source of truth: one function per declaration, scalar types
 * limited to int / long long / unsigned int / unsigned long long / size_t / float, everything else is a pointer.
 *
 * Reference parity citations are relative to /root/reference/.
 */
#ifndef DTF_B200_H_
#define DTF_B200_H_
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DTF_OK 0
#define DTF_ERR_ARG 1
#define DTF_ERR_ALIGN 2
#define DTF_ERR_CUDA 3
#define DTF_ERR_DRIVER 4
#define DTF_ERR_TENSORMAP 5
#define DTF_ERR_TOO_LARGE 6
#define DTF_ERR_UNSUPPORTED 7

/* ---- tcgen05 / TMEM / TMA GEMM family (csrc/igemm.cu) --------------------------------------------------------- */

/* C[M,N] = act(A[M,K] * B[K,N] + bias). B row-major [K,N] (ldb). A row-major [M,K] (lda), or [K,M] when flag bit3.
 * flags: bit0 relu, bit1 fp32 output, bit2 accumulate into fp32 C with red.add (split-K), bit3 A stored [K,M].
 * Replaces tf.matmul + bias_add + relu of FC_layer (distribute_tools.py:190-208) and its gradients. */
int dtf_gemm_bf16(const void* A, const void* B, void* C, const float* bias, int M, int N, int K, int lda, int ldb,
                  int ldc, int flags, void* stream);
int dtf_gemm_bf16_ex(const void* A, const void* B, void* C, const float* bias, const void* addend, float* stats,
                     int M, int N, int K, int lda, int ldb, int ldc, int flags, void* stream);
int dtf_gemm_stats_rows(const void* A, const void* B, const void* C, int M, int N, int K, int lda, int ldb, int ldc);

/* Dense GEMMs run a persistent kernel (tile loop per CTA, smem ring kept full across tiles, accumulator double
 * buffered in TMEM so the epilogue overlaps the next tile). bit0 of `on`: persistent kernel (0 = one tile per CTA);
 * bit1: allow 128x256 tiles for N >= 256. Default 3. Returns the previous setting. */
int dtf_set_persistent_gemm(int on);
int dtf_set_persistent_halo(int on);

/* K-AG-GEMM: the all-gather of parameter-server weights fused into the consuming GEMM. B is sharded over the heaps
 * of `world` ranks (peer_b[r] = rank r's mapping of the matrix, `bounds` = owned arena ranges, b_arena_off = arena
 * offset of the matrix); the TMA producer reads every tile from its OWNER over NVLink, no gather pass and no local
 * copy. shard_by_n 0: owners are row (K) ranges, multiples of 64 rows; 1: owners are column (N) ranges of a
 * transposed shadow, multiples of the N tile. Replaces the ps->worker variable pull (distribute_train.py:109-110). */
int dtf_gemm_bf16_ag(const void* A, const void* const* peer_b, void* C, const float* bias, int M, int N, int K, int lda,
                     int ldb, int ldc, int flags, int world, const long long* bounds, long long b_arena_off,
                     int shard_by_n, void* stream);

/* Implicit-GEMM convolution on NHWC bf16. geo 0: forward conv (gather from x[n_img,H,W,C], output pixel grid PHxPW);
 * geo 1: data gradient / transposed conv (gather from dy[n_img,H,W,C], output grid PHxPW is the input image grid).
 * W2d is the row-major [KH*KW*C, n_out] filter matrix. flags: bit0 relu, bit1 fp32 output.
 * Replaces tf.nn.conv2d+bias_add+relu (distribute_tools.py:69-121) and tf.nn.conv2d_transpose (:131-157). */
int dtf_conv_igemm_bf16(const void* src, const void* W2d, void* out, const float* bias, int n_img, int H, int W,
                        int C, int PH, int PW, int KH, int KW, int n_out, int stride, int pad_h, int pad_w, int geo,
                        int flags, void* stream);

/* Halo-tile convolution (csrc/conv_halo.cu): stride-1 KxK conv / data gradient with C % 64 == 0. One 4-D TMA load per
 * (tile, 64-channel block) brings a zero-padded halo into shared memory and all KH*KW taps are shifted tcgen05
 * descriptor views of it (KH*KW times less L2 traffic than the gather kernel). mirror = 1: data gradient with
 * per-tap transposed filters. flags: bit0 relu, bit8 descriptor base-offset mode. */
int dtf_conv_halo_supported(int H, int W, int C, int KH, int KW, int n_out);
int dtf_conv_halo_bf16(const void* src, const void* W2d, void* out, const float* bias, int n_img, int H, int W, int C,
                       int KH, int KW, int n_out, int pad_h, int pad_w, int mirror, int flags, void* stream);
int dtf_conv_halo_bf16_ex(const void* src, const void* W2d, void* out, const float* bias, float* stats, int n_img,
                          int H, int W, int C, int KH, int KW, int n_out, int pad_h, int pad_w, int mirror, int flags,
                          int n_phase, const unsigned long long* phase_mask, void* stream);
int dtf_conv_halo_stats_rows(int n_img, int H, int W, int C, int KH, int KW, int n_out);

/* Halo-tile filter gradient (stride 1, C % 64 == 0, n_out % 64 == 0, <= 16 taps): x halo + dy tile per image tile,
 * all taps accumulate in TMEM (tap pairs share one M=128 MMA through the descriptor's leading-dimension offset),
 * one red.add pass into dW at the end. dW is fp32 and accumulated: zero it first. */
int dtf_conv_wgrad_halo_supported(int H, int W, int C, int KH, int KW, int n_out);
int dtf_conv_wgrad_halo_bf16(const void* x, const void* dy, float* dw, int n_img, int H, int W, int C, int KH, int KW,
                             int n_out, int pad_h, int pad_w, void* stream);

/* Filter gradient dW[KH*KW*C, n_out] (fp32, red.add accumulated; zero it first) from x[n_img,H,W,C] and
 * dy[n_img,PH,PW,n_out]. Backward of the conv layers above (TF autodiff, distribute_tower.py:26-27). */
int dtf_conv_wgrad_bf16(const void* x, const void* dy, float* dw, int n_img, int H, int W, int C, int PH, int PW,
                        int KH, int KW, int n_out, int stride, int pad_h, int pad_w, void* stream);

/* ---- memory-bound ops (csrc/ops.cu) --------------------------------------------------------------------------- */

/* [pixels,C] fp32 or uint8 -> [pixels,Cpad] bf16, y = x*scale + shift, pad channels zero (input pipeline tail,
 * distribute_input.py:54-106). */
int dtf_convert_input(const void* src, void* dst, long long pixels, int C, int Cpad, float scale, float shift,
                      int src_is_u8, void* stream);
/* y[n,oh,ow,:] = x[n,oh*s,ow*s,:] and its adjoint (zero-filled scatter): a 1x1 stride-s convolution becomes
 * "select pixels, then dense GEMM" instead of an im2col gather. */
int dtf_set_tuning(int key, int value);
int dtf_pack_window(const void* x, void* P, int n_img, int H, int W, int Cx, int C, int Hp, int Wq, int Cp,
                    int rp, int kwn, int sx, int cs, int oy, int ox, void* stream);
int dtf_unpack_s2d(const void* dP, void* dx, int n_img, int H, int W, int C, int Hp, int Wq, int oy, int ox,
                   void* stream);
int dtf_block_copy(const void* src, int ld_src, void* dst, int ld_dst, const void* descs, int n_blocks,
                   int max_elems, int dtype, int accumulate, void* stream);
int dtf_subsample(const void* x, void* y, int n_img, int H, int W, int C, int OH, int OW, int s, void* stream);
int dtf_upsample_zero(const void* dy, void* dx, int n_img, int H, int W, int C, int OH, int OW, int s, void* stream);
int dtf_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream);
int dtf_cast_bf16_to_f32(const void* src, float* dst, long long n, void* stream);

/* acti_layer (distribute_tools.py:123-128) and its gradient; z = a + b (+relu) for residual joins. */
int dtf_relu_fwd(const void* x, void* y, long long n, void* stream);
int dtf_relu_bwd(const void* dy, const void* y, void* dx, long long n, void* stream);
int dtf_add(const void* a, const void* b, void* z, long long n, int relu, void* stream);

/* db[N] += column sums of dy[R,N]; fused variant also writes dz = dy*(y>0). (bias_add gradient) */
int dtf_bias_grad(const void* dy, float* db, long long R, int N, void* stream);
int dtf_relu_bwd_bias_grad(const void* dy, const void* y, void* dz, float* db, long long R, int N, void* stream);

/* Batch norm over the rows of x[R,N]: per-channel BN uses R = batch*H*W, N = C; the reference's literal
 * batch_norm (moments over axis 0, no affine, eps 1e-3; distribute_tools.py:168-180) uses R = batch, N = H*W*C. */
/* The column reductions write per-row-block partial sums part[dtf_colreduce_blocks(R,N)][2][N] (no atomics); the
 * finalize / coefficient kernels add them up in a fixed order, so batch-norm results are bitwise reproducible. */
int dtf_colreduce_blocks(long long R, int N);
int dtf_bn_stats(const void* x, float* part, long long R, int N, void* stream);
int dtf_bn_finalize(const float* part, const float* gamma, const float* beta, float* mean, float* rstd, float* scale,
                    float* shift, float* run_mean, float* run_var, int N, long long R, float eps, float momentum,
                    void* stream);
int dtf_bn_finalize_rows(const float* part, int n_rows, const float* gamma, const float* beta, float* mean,
                         float* rstd, float* scale, float* shift, float* run_mean, float* run_var, int N, long long R,
                         float eps, float momentum, void* stream);
int dtf_bn_fold(const float* run_mean, const float* run_var, const float* gamma, const float* beta, float* scale,
                float* shift, int N, float eps, void* stream);
int dtf_bn_apply(const void* x, const void* res, void* y, const float* scale, const float* shift, long long R, int N,
                 int relu, void* stream);
/* Same, and with relu also writes the ReLU decisions as a bitmask relu_mask[R][N/8] (bit j of a byte = channel 8k+j
 * is positive): the backward of a batch norm with a residual input reads 1 bit per element instead of y. */
int dtf_bn_apply_mask(const void* x, const void* res, void* y, void* relu_mask, const float* scale,
                      const float* shift, long long R, int N, int relu, void* stream);
/* Same, with the residual itself passing through a batch norm: y = act(x*scale + shift + res*res_scale + res_shift).
 * A projection shortcut's normalised tensor is then never materialised (res = the shortcut convolution's raw output). */
int dtf_bn_apply_dual(const void* x, const void* res, void* y, void* relu_mask, const float* scale, const float* shift,
                      const float* res_scale, const float* res_shift, long long R, int N, int relu, void* stream);
/* relu: 0 none, 1 mask from y, 2 mask recomputed from x with the forward's folded scale/shift (no y read, no dz),
 * 3 `y` is the bitmask of dtf_bn_apply_mask. */
int dtf_bn_bwd_reduce(const void* dy, const void* y, const void* x, void* dz, float* part, const float* mean,
                      const float* rstd, const float* mask_scale, const float* mask_shift, long long R, int N,
                      int relu, void* stream);
/* coef: [3][N] float workspace. dgamma_acc / dbeta_acc (may be NULL) receive += sum dz*xhat / sum dz. */
int dtf_bn_bwd_apply(const void* dz, const void* x, void* dx, const float* part, const float* mean, const float* rstd,
                     const float* gamma, float* coef, float* dgamma_acc, float* dbeta_acc, const float* mask_scale,
                     const float* mask_shift, long long R, int N, void* stream);

/* pool (distribute_tools.py:160-165): max / avg, TF SAME semantics via explicit top/left padding. Max pooling
 * records the in-window position of the first maximum (one byte per output element, `idx`) for the backward. */
int dtf_pool_fwd(const void* x, void* y, void* idx, int n_img, int H, int W, int C, int OH, int OW, int kh, int kw,
                 int sh, int sw, int pt, int pl, int is_max, void* stream);
/* Pooling over relu(x * pre_scale[c] + pre_shift[c]) (3x3 / 2x2 windows): a batch norm + ReLU in front of a pooling
 * layer is applied while pooling and its output is never stored (ResNet stem). */
int dtf_pool_fwd_affine(const void* x, void* y, void* idx, const float* pre_scale, const float* pre_shift, int n_img,
                        int H, int W, int C, int OH, int OW, int kh, int kw, int sh, int sw, int pt, int pl,
                        int is_max, void* stream);
int dtf_pool_bwd(const void* idx, const void* dy, void* dx, int n_img, int H, int W, int C, int OH, int OW, int kh,
                 int kw, int sh, int sw, int pt, int pl, int is_max, void* stream);

/* norm / tf.nn.lrn (distribute_tools.py:183-187): y = x / (bias + alpha * sum_{|j-c|<=radius} x_j^2)^beta. */
int dtf_lrn_fwd(const void* x, void* y, long long pixels, int C, int radius, float bias, float alpha, float beta,
                void* stream);
int dtf_lrn_bwd(const void* x, const void* dy, void* dx, long long pixels, int C, int radius, float bias, float alpha,
                float beta, void* stream);

/* Fused loss forward+backward (Loss.loss implementations for distribute_tower.py:60-76). `scratch` is global
 * memory provided by the caller: 1 + B floats (softmax_xent) or 1 + dtf_loss_scratch_blocks() floats (mse, l2).
 * Word 0 must be zero before the first call; every call leaves it zero. dlogits / dpred may be NULL. */
int dtf_loss_scratch_blocks(void);
int dtf_softmax_xent(const void* logits, const long long* labels, void* dlogits, float* loss, float* scratch, int B,
                     int C, float gscale, int is_bf16, void* stream);
int dtf_mse(const void* pred, const void* target, void* dpred, float* loss, float* scratch, long long n, float gscale,
            int is_bf16, void* stream);
/* out = 0.5 * scale * sum(w^2): tf.nn.l2_loss weight-decay term (distribute_tools.py:63-65). */
int dtf_l2_loss(const float* w, float* out, float* scratch, long long n, float scale, void* stream);
/* Allocation-free variant of dtf_softmax_xent (no counterpart in the reference, which has no native code or
 * allocation contract at all). For callers that cannot allocate device memory -- a host application that already
 * owns all GPU memory through its own pool, a latency-bound loop where cudaMalloc would stall the device, a plug-in
 * contract that forbids allocation after start-up: the scratch area comes from a fixed array inside the library
 * (256 KiB + 4 B per GPU, resident from library load to process exit) instead of from the caller. At most
 * dtf_softmax_xent_noalloc_max_rows() = 65536 rows per call; a larger B returns DTF_ERR_TOO_LARGE and launches
 * nothing. The first call on each GPU resolves the array's device address: make it at start-up with that GPU
 * current, before any stream capture. All calls on a GPU share the array (it keeps the previous call's partial
 * losses; the arrival counter is cleared on the caller's stream before each launch), so calls that can overlap on
 * the GPU race: serialise them. Neither cudaMalloc, cudaFree nor any allocating routine is called. */
int dtf_softmax_xent_noalloc_max_rows(void);
int dtf_softmax_xent_noalloc(const void* logits, const long long* labels, void* dlogits, float* loss, int B, int C,
                             float gscale, int is_bf16, void* stream);
int dtf_top1_hits(const void* logits, const long long* labels, unsigned int* hits, int B, int C, int is_bf16,
                  void* stream);

/* Optimizer apply on a flat arena (distribute.py:26,112; distribute_train.py:151-158). kind 0 SGD, 1 momentum,
 * 2 Adam. g is multiplied by gscale; wd*p is added to it; the bf16 shadow (may be NULL) receives the new weights. */
int dtf_optim_step(int kind, float* p, const float* g, float* m, float* v, void* shadow, long long n, float lr,
                   float b1, float b2, float eps, float wd, float gscale, long long step, const float* hyper,
                   void* stream);
/* Device-resident hyper-parameters. `hyper` (4 floats of device memory, optional in every optimizer entry point):
 * [0] learning rate, written by the host when the schedule moves it (LearningRate handler,
 * distribute_learningrate.py:12-50); [1] Adam's bias-corrected rate, recomputed by dtf_optim_hyper_step; word 2 the
 * int32 step counter. When `hyper` is non-NULL the kernels read lr / lr_t from it at run time and ignore the by-value
 * lr / step, so a captured CUDA graph follows the schedule. dtf_optim_hyper_step: step += 1, lr_t refreshed. */
int dtf_optim_hyper_step(float* hyper, int kind, float b1, float b2, void* stream);
/* Batched per-tap transpose [T,Ci,Co] -> [T,Co,Ci] of bf16 filters (second shadow used by dgrad / deconv). */
int dtf_tap_transpose(const void* src, void* dst, const void* descs_dev, int ndesc, int total_tiles, void* stream);
/* Philox-4x32-10 initialisers: kind 0 constant a, 1 uniform [a,b), 2 normal(a,b), 3 truncated normal(a,b) |z|<=2
 * (tf.truncated_normal_initializer / xavier / constant, distribute_tools.py:62,100,103,135,200,203). */
int dtf_init(float* out, long long n, int kind, float a, float b, unsigned long long seed,
             unsigned long long stream_id, void* stream);
int dtf_fill_u32(void* p, long long n, unsigned int v, void* stream);
/* 0 = not capturing, 1 = capturing, 2 = capture invalidated, -1 = query failed. */
int dtf_stream_capture_status(void* stream);

/* ---- peer-memory collectives (csrc/comm.cu) ------------------------------------------------------------------- */

int dtf_comm_flag_page_bytes(void);
int dtf_comm_max_blocks(void);
int dtf_comm_heap_alloc(size_t bytes, void** ptr_out, void* handle_out64);
int dtf_comm_heap_free(void* ptr);
int dtf_comm_create(int rank, int world, void* local_heap, size_t heap_bytes, const void* handles, void** comm_out);
int dtf_comm_create_local(int rank, int world, void* const* heaps, size_t heap_bytes, void** comm_out);
int dtf_comm_destroy(void* comm);
void* dtf_comm_peer_ptr(void* comm, int r);
/* NVLS: attach the multicast mapping of the heap; K-RS-OPT then reduces with multimem.ld_reduce (in-switch add) and
 * publishes weights with multimem.st (in-switch replication). */
int dtf_comm_set_multicast(void* comm, void* mc_base);

/* K-AR: one-shot allreduce of grad[start, start+count) (heap offset grad_off) fused with x gscale and the optimizer
 * update of the local replica. Replaces Tower.average_gradients + apply (distribute_tower.py:78-114). With
 * master == NULL the averaged gradient is written to avg_out instead. */
int dtf_comm_allreduce_optim(void* comm, size_t grad_off, long long start, long long count, int kind, float lr,
                             float b1, float b2, float eps, float wd, float gscale, long long step,
                             const float* hyper, float* master, float* m, float* v, void* shadow, float* avg_out,
                             int blocks, void* stream);
/* K-RS-OPT: reduce-scatter into the owner shard fused with the owner-side optimizer, then all-gather of the new
 * bf16 weights by peer stores (except the `nopush_ranges` [lo,hi) element ranges, which consumers pull with
 * dtf_gemm_bf16_ag). Inside [f32_lo, f32_hi) the owner also broadcasts the fp32 master (heap offset master_off;
 * (size_t)-1 = never): that region holds the variables which kernels read in fp32 (biases, batch-norm gamma/beta).
 * Replaces the worker->ps push, ps-side apply and ps->worker pull of SyncReplicasOptimizer /
 * replica_device_setter (distribute_train.py:109-110,146-158).
 * backup_r in (0, world): backup workers (replicas_to_aggregate = backup_r, distribute_train.py:146-155,
 * distribute_flags.py:26-29): each owner aggregates the first backup_r gradients to ARRIVE (ticket counter in its
 * flag page), drops the late ones; pass gscale = 1/backup_r. 0 or world: every rank is aggregated. */
int dtf_comm_rs_optim_ag(void* comm, size_t grad_off, const long long* bounds, int kind, float lr, float b1, float b2,
                         float eps, float wd, float gscale, long long step, const float* hyper, float* master,
                         float* m, float* v, size_t shadow_off, size_t master_off, long long f32_lo,
                         long long f32_hi, int n_nopush, const long long* nopush_ranges, int backup_r, int blocks,
                         void* stream);
int dtf_comm_allgather_slices(void* comm, size_t buf_off, const long long* bounds16, int blocks, void* stream);
/* K-ASYNC-PUSH / owner apply: asynchronous parameter-server mode (extension; the reference is sync-only,
 * distribute_train.py:151). */
/* Version stamps and staleness clocks live in the flag page and are advanced by the kernels themselves, so a replayed
 * CUDA graph keeps publishing new versions. async_begin (start of a step): remember which update of every owner the
 * weights of this step correspond to. async_push: gradient slices into the owners' mailboxes + next version + that
 * stamp. async_apply: the owner applies every newer mailbox (one optimizer update per push, never waits), drops a
 * push computed more than max_staleness owner updates ago (< 0: no bound), broadcasts bf16 weights (+ fp32 master in
 * [f32_lo, f32_hi)). async_stats: {dropped, updates applied by this owner, versions pushed by this rank}. */
int dtf_comm_async_begin(void* comm, void* stream);
int dtf_comm_async_push(void* comm, size_t grad_off, size_t mailbox_off, long long slice_cap, const long long* bounds,
                        int blocks, void* stream);
int dtf_comm_async_apply(void* comm, size_t mailbox_off, long long slice_cap, const long long* bounds, int kind,
                         float lr, float b1, float b2, float eps, float wd, float gscale, long long step,
                         const float* hyper, float* master, float* m, float* v, size_t shadow_off, size_t master_off,
                         long long f32_lo, long long f32_hi, int max_staleness, unsigned int* applied_out_dev,
                         int blocks, void* stream);
int dtf_comm_async_stats(void* comm, unsigned int* out3, void* stream);
/* Device-side barrier / liveness (token queue, done queue, init handshake: distribute_train.py:43-46,86-90,160-167). */
int dtf_comm_barrier(void* comm, void* stream);
int dtf_comm_heartbeat(void* comm, unsigned int beat, unsigned int* seen_out_dev, void* stream);
/* Failure detection (the reference hangs forever when a worker dies, distribute_train.py:86-90): every
 * device-side wait is bounded in TIME (default 30 s, DTF_COMM_TIMEOUT_MS, 0 = unbounded). A wait that expires makes
 * its kernel return and records {code 1 = barrier / 2 = arrivals, block, peer, epoch} in mapped host memory;
 * dtf_comm_error copies that record without any CUDA call (a watchdog thread may poll it), dtf_comm_clear_error
 * re-arms the communicator after the group has recovered. */
int dtf_comm_set_timeout_ms(void* comm, long long ms);
int dtf_comm_error(void* comm, unsigned int* out4);
int dtf_comm_clear_error(void* comm, void* stream);
/* Error code 3 (collective mismatch): every barrier carries the kind of collective that entered it; a rank that meets a
 * peer at the same epoch inside a different collective records it instead of silently exchanging the wrong data. */
/* Fault injection / tests: occupy the stream for `ns` nanoseconds (an artificial straggler). */
int dtf_debug_sleep(long long ns, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DTF_B200_H_ */
