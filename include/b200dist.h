/* b200dist native library — C ABI of b200dist/_native/libb200dist.so (sm_100a only).
 *
 * Every entry point returns 0 on success or one of the B200_ERR_* codes; b200_last_error() returns the
 * message of the last failure on the calling thread.  All pointers are DEVICE pointers unless stated
 * otherwise, `stream` is the CUDA stream the work is enqueued on, dtype codes are 0 = bf16, 1 = fp32.
 * Entry points never synchronise the device and (unless their comment says so) never allocate: scratch
 * is passed in by the caller.  Citations are paths under the reference repository
 * (negin513/distributed-pytorch-hpc) naming the library call each kernel replaces.
 */
#ifndef B200DIST_H_
#define B200DIST_H_

#include <cuda_runtime_api.h>
#include <stddef.h>

#ifdef __cplusplus
#define B200DIST_API extern "C"
#else
#define B200DIST_API
#endif

enum { B200_OK = 0, B200_ERR_CUDA = 1, B200_ERR_SHAPE = 2, B200_ERR_UNSUPPORTED = 3, B200_ERR_LIMIT = 4 };

/* ---- Runtime ----
 * no counterpart in the reference (it has no native code) */
B200DIST_API int b200_version(void);
B200DIST_API const char* b200_last_error(void);
B200DIST_API int b200_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin);

/* ---- GEMM (tcgen05 / TMEM / TMA) ----
 * replaces cuBLAS behind nn.Linear / torch.matmul: fsdp_tp/llama2_model.py:207,228,267,447 and their autograd */
B200DIST_API int b200_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K,
                            long long lda, long long ldb, long long ldd, int a_mn_major,
                            int b_mn_major, int out_fp32, const void* bias, const void* residual,
                            long long ldr, float alpha, int act, int accumulate, int config,
                            int max_ctas, cudaStream_t stream);

/* ---- Fused GEMM + collective over peer memory (TP / SP hot ops) ----
 * ag_gemm: D[M,N] = allgather_seq(A) . B^T.  a_ptrs[r] = rank r's [M/world, K] bf16 shard in symmetric memory; comm CTAs
 *          inside the GEMM grid pull peer shards with TMA bulk copies and flag each `rows_per_flag` rows ready, the MMA
 *          CTAs consume them tile by tile (replaces all_gather_into_tensor + cuBLAS behind ColwiseParallel,
 *          scripts/06_hybrid_parallelism/01_fsdp_tp_hybrid.py:129-141).
 * gemm_rs: partial = A . B^T is stored row-wise straight into the owning rank's staging slab (stage_ptrs[o] =
 *          [world, M/world, N] bf16); b200_comm_slab_reduce then sums the slabs (+residual) on the owner
 *          (replaces cuBLAS + reduce_scatter_tensor behind RowwiseParallel(output_layouts=Shard(1)), same file :136,143). */
B200DIST_API int b200_ag_gemm_bf16(const void* const* a_ptrs, const void* B, void* D, int M, int N, int K, long long ldb,
                               long long ldd, int b_mn_major, const void* bias, const void* residual, long long ldr,
                               int act, void* flags, int rows_per_flag, unsigned epoch, int n_comm_clusters,
                               const void* const* sig_ptrs, void* counter, int channel, unsigned seq, int world,
                               int rank, cudaStream_t stream);
B200DIST_API int b200_gemm_rs_bf16(const void* A, const void* B, const void* const* stage_ptrs, int M, int N, int K,
                               long long lda, long long ldb, int b_mn_major, const void* residual, long long ldr,
                               int world, int rank, cudaStream_t stream);

/* implicit-GEMM convolution (stride 1, NHWC bf16): the A operand is read straight from the activation through a 4-D
 * tensor map, one shifted TMA box per (filter tap, 64-channel slice); also serves dgrad (dY, flipped filter, pad' = K-1-pad).
 * replaces cuDNN behind torchvision ResNet / SimpleUNet convolutions: scripts/main.py:249, multinode_ddp_unet.py:190-198 */
B200DIST_API int b200_conv2d_igemm_bf16(const void* x, const void* wmat, void* out, int N, int H, int W, int C, int Cout,
                                    int KH, int KW, int pad, const void* bias, const void* residual, int act,
                                    cudaStream_t stream);
B200DIST_API int b200_conv2d_igemm_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int Cout,
                                     int KH, int KW, int pad, int accumulate, cudaStream_t stream);
/* stride 1 or 2 (torchvision ResNet stage transitions, /root/reference/scripts/main.py:249): the activation tensor map
 * steps over every `stride`-th pixel, so strided forward and weight-gradient convolutions need no patch matrix either */
B200DIST_API int b200_conv2d_igemm_strided_bf16(const void* x, const void* wmat, void* out, int N, int H, int W, int C, int Cout,
                                                int KH, int KW, int pad, int stride, const void* bias, const void* residual,
                                                int act, cudaStream_t stream);
B200DIST_API int b200_conv2d_igemm_wgrad_strided(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int Cout,
                                                 int KH, int KW, int pad, int stride, int accumulate, cudaStream_t stream);
/* small-channel stride-2 stem (ResNet conv1, 7x7 / 2 on RGB; /root/reference/scripts/main.py:249 via torchvision): the input
 * comes zero-padded to 8 channels and spatially, a sliding-window tensor map turns (8 taps x 8 channels) into one 128-byte operand row */
B200DIST_API int b200_conv2d_stem_bf16(const void* xp, const void* wmat, void* out, int N, int Hp, int Wp, int Ho, int Wo,
                                       int Cout, int KH, cudaStream_t stream);
B200DIST_API int b200_conv2d_stem_wgrad(const void* xp, const void* dy, float* dw, int N, int Hp, int Wp, int Ho, int Wo,
                                        int Cout, int KH, cudaStream_t stream);

/* ---- Attention (tcgen05 flash attention) ----
 * replaces F.scaled_dot_product_attention: fsdp_tp/llama2_model.py:218-223; explicit softmax attention: scripts/03_tensor_parallel_tp/tensor_parallel_vit.py:120-122 */
B200DIST_API int b200_flash_attn_fwd(const void* Q, const void* K, const void* V, void* O, float* LSE,
                                 int B, int H, int Hkv, int Sq, int Sk, int hd,
                                 long long q_sb, long long q_ss, long long q_sh,
                                 long long k_sb, long long k_ss, long long k_sh,
                                 long long v_sb, long long v_ss, long long v_sh,
                                 long long o_sb, long long o_ss, long long o_sh,
                                 float scale, int causal, int q_offset, int k_offset,
                                 cudaStream_t stream);
/* ring (context-parallel) attention forward as ONE kernel: communication CTAs stage the peers' K/V chunks into the local
 * gathered buffer with plain TMA bulk copies while the attention CTAs consume arrived 128-key blocks; (O, m, l) stay on chip
 * (the reference only describes the algorithm: docs/guide/08_sequence_parallel.md:82-143) */
B200DIST_API int b200_flash_attn_fwd_ring(const void* Q, const void* const* kv_ptrs, void* O, float* LSE, int B, int H, int Hkv,
                                          int Sq, int chunk_rows, int hd, long long q_sb, long long q_ss, long long q_sh,
                                          long long o_sb, long long o_ss, long long o_sh, float scale, int causal, void* flags,
                                          unsigned epoch, int n_comm, const void* const* sig_ptrs, void* counter, int channel,
                                          unsigned seq, int world, int rank, cudaStream_t stream);
B200DIST_API int b200_flash_attn_bwd(const void* Q, const void* K, const void* V, const void* O,
                                 const void* dO, const float* LSE, float* delta, void* dQ, void* dK,
                                 void* dV, int B, int H, int Hkv, int Sq, int Sk, int hd,
                                 const long long* strides /* 24: q,k,v,o,do,dq,dk,dv x (b,s,h) */,
                                 float scale, int causal, int q_offset, int k_offset,
                                 cudaStream_t stream);

/* ---- Normalisation ----
 * RMSNorm fsdp_tp/llama2_model.py:115-142; nn.LayerNorm tensor_parallel_vit.py:141-143 */
B200DIST_API int b200_norm_bwd_partials(int rows);
B200DIST_API int b200_rmsnorm_fwd(const void* x, const void* w, void* y, float* rstd, int rows, int D,
                              float eps, int dtype, cudaStream_t stream);
B200DIST_API int b200_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd,
                              void* dx, void* dw, float* scratch, int rows, int D, int add_to_dx,
                              int accumulate_dw, int dtype, cudaStream_t stream);
/* dx = dres + d/dx: the residual branch of a pre-norm transformer block folded into the norm backward (dres may alias dx) */
B200DIST_API int b200_rmsnorm_bwd_res(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                                      void* dx, void* dw, float* scratch, int rows, int D, int accumulate_dw, int dtype,
                                      cudaStream_t stream);
/* No counterpart in the reference.  Same result as b200_rmsnorm_bwd, but the partial-sum scratch is a fixed
 * 4.6 MiB array inside the library instead of a caller buffer, for hosts that cannot allocate device memory
 * (no allocator of their own, HBM fully pre-partitioned, or cudaMalloc banned from the hot loop).  Calls
 * nothing that allocates.  D <= 8192 (else B200_ERR_LIMIT).  The first call on a device looks up the array's
 * device address: issue it before any stream capture.  All calls on a GPU share the array, so calls that can
 * overlap on that GPU race; serialise them. */
B200DIST_API int b200_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* w, const float* rstd,
                                      void* dx, void* dw, int rows, int D, int add_to_dx,
                                      int accumulate_dw, int dtype, cudaStream_t stream);
B200DIST_API int b200_layernorm_fwd(const void* x, const void* w, const void* b, void* y, float* mean,
                                float* rstd, int rows, int D, float eps, int dtype,
                                cudaStream_t stream);
B200DIST_API int b200_layernorm_bwd(const void* dy, const void* x, const void* w, const float* mean,
                                const float* rstd, void* dx, void* dw, void* db, float* scratch,
                                int rows, int D, int accumulate, int dtype, cudaStream_t stream);

/* ---- Elementwise / embedding / loss ----
 * RoPE llama2_model.py:73-100; SwiGLU :266-267; nn.Embedding :374,440; CrossEntropyLoss scripts/main.py:308 */
B200DIST_API int b200_rope(const void* x, void* y, const float* freqs, long long rows, int S, int H,
                       int hd, int pos_offset, int backward, int dtype, cudaStream_t stream);
/* the same rotation inside a fused QKV projection output: rotates the first n_heads heads (q then k) of every
 * [tokens, ld] row in one launch, in place if y == x (apply_rotary_emb on xq and xk, llama2_model.py:73-100) */
B200DIST_API int b200_rope_strided(const void* x, void* y, const float* freqs, long long tokens, int S, int n_heads, int hd,
                                   long long ldx, long long ldy, int pos_offset, int backward, int dtype,
                                   cudaStream_t stream);
B200DIST_API int b200_swiglu_fwd(const void* a, const void* b, void* y, long long rows, int F,
                             long long lda, long long ldb, long long ldy, int dtype,
                             cudaStream_t stream);
B200DIST_API int b200_swiglu_bwd(const void* dy, const void* a, const void* b, void* da, void* db,
                             long long rows, int F, long long ldy, long long lda, long long ldb,
                             long long ldda, long long lddb, int dtype, cudaStream_t stream);
/* softmax over attention scores with Philox dropout on the probabilities (mask regenerated from seed / offset, never stored):
   probs = softmax(scale * scores [+ causal]), probs_dropped = probs * mask / (1 - p); rows = B * H * sq, cols % 4 == 0 */
B200DIST_API int b200_softmax_dropout_fwd(const void* scores, void* probs, void* probs_dropped, long long rows, int cols, int sq,
                                      float scale, int causal, float p_drop, unsigned long long seed,
                                      unsigned long long offset, int dtype, cudaStream_t stream);
B200DIST_API int b200_softmax_dropout_bwd(const void* probs, const void* d_probs_dropped, void* d_scores, long long rows, int cols,
                                      float scale, float p_drop, unsigned long long seed, unsigned long long offset,
                                      cudaStream_t stream);
B200DIST_API int b200_elementwise(const void* x, const void* y2, void* out, long long n, int op,
                              float alpha, int dtype, cudaStream_t stream);
B200DIST_API int b200_cast_scale(const void* src, int src_dtype, void* dst, int dst_dtype, long long n,
                             float alpha, cudaStream_t stream);
B200DIST_API int b200_embedding_fwd(const long long* idx, const void* w, void* out, long long ntok, int D,
                                long long vocab_start, long long vocab_local, int dtype,
                                cudaStream_t stream);
B200DIST_API int b200_embedding_bwd(const long long* idx, const void* dy, float* dw, long long ntok, int D,
                                long long vocab_start, long long vocab_local, int dtype,
                                cudaStream_t stream);
B200DIST_API int b200_xent_fwd(const void* logits, const long long* target, float* loss, float* lse,
                           int rows, int V, long long ld, long long ignore_index, int dtype,
                           cudaStream_t stream);
B200DIST_API int b200_xent_bwd(const void* logits, const long long* target, const float* lse,
                           const float* gscale, void* dlogits, int rows, int V, long long ld,
                           long long ldd, long long ignore_index, int dtype, cudaStream_t stream);
/* vocabulary-parallel pieces: local columns [vocab_start, vocab_start + V) of a [rows, v_total] logits matrix sharded
 * over tensor-parallel ranks (no counterpart in the reference, whose hybrid recipe gathers the full logits) */
B200DIST_API int b200_xent_fwd_vp(const void* logits, const long long* target, float* loss, float* lse, int rows, int V,
                              long long ld, long long ignore_index, long long vocab_start, int dtype,
                              cudaStream_t stream);
B200DIST_API int b200_xent_bwd_vp(const void* logits, const long long* target, const float* lse, const float* gscale,
                              void* dlogits, int rows, int V, long long ld, long long ldd, long long ignore_index,
                              long long vocab_start, long long v_total, int dtype, cudaStream_t stream);
B200DIST_API int b200_sum_scale(const float* x, float* out, int n, float scale, cudaStream_t stream);

/* ---- Optimizers ----
 * torch.optim.AdamW 01_fsdp_tp_hybrid.py:162; torch.optim.SGD scripts/main.py:309-311 */
B200DIST_API int b200_adamw(void* p, int p_dtype, void* p_lowp, const void* g, int g_dtype, float* m,
                        float* v, long long n, float lr, float beta1, float beta2, float eps,
                        float wd, float bc1, float bc2, float grad_scale, cudaStream_t stream);
/* AdamW with the step-dependent scalars {lr, 1 - beta1^t, sqrt(1 - beta2^t)} read from device memory: usable inside a CUDA
 * graph that is replayed every step (no counterpart in the reference, which runs eager torch.optim.AdamW) */
B200DIST_API int b200_adamw_dev(void* p, int p_dtype, void* p_lowp, const void* g, int g_dtype, float* m, float* v,
                            long long n, float beta1, float beta2, float eps, float wd, const float* scalars,
                            float grad_scale, cudaStream_t stream);
B200DIST_API int b200_sgd(void* p, int p_dtype, void* p_lowp, const void* g, int g_dtype, float* buf,
                      long long n, float lr, float momentum, float wd, float grad_scale,
                      int first_step, cudaStream_t stream);
B200DIST_API int b200_sumsq(const void* x, float* out, long long n, int dtype, cudaStream_t stream);

/* ---- Convolution-network ops (NHWC) ----
 * cuDNN conv/BN/pool via torchvision scripts/main.py:249; UNet blocks multinode_ddp_unet.py:171-229 */
B200DIST_API int b200_im2col(const void* x, void* col, int N, int H, int W, int C, int Cp, int KH, int KW, int stride, int pad, int Ho, int Wo, int dtype, cudaStream_t stream);
B200DIST_API int b200_col2im(const void* dcol, void* dx, int N, int H, int W, int C, int Cp, int KH, int KW, int stride, int pad, int Ho, int Wo, int dtype, cudaStream_t stream);
B200DIST_API int b200_col_reduce(const void* a, const void* xin, const void* y, const float* mean, const float* rstd,
                             float* sums, long long rows, int C, int mode, int relu_mask, int dtype,
                             cudaStream_t stream);
B200DIST_API int b200_bn_finalize(const float* sums, float* mean, float* rstd, float* running_mean, float* running_var,
                              long long rows, int C, float eps, float momentum, cudaStream_t stream);
B200DIST_API int b200_bn_apply(const void* x, const float* mean, const float* rstd, const float* gamma, const float* beta,
                           const void* residual, void* y, long long rows, int C, int relu, int dtype,
                           cudaStream_t stream);
B200DIST_API int b200_bn_bwd_apply(const void* dy, const void* x, const void* y, const float* mean, const float* rstd,
                               const float* gamma, const float* sums, void* dx, void* dres, long long rows, int C,
                               int relu_mask, int dtype, cudaStream_t stream);
B200DIST_API int b200_maxpool(const void* a, const void* x, void* idx, void* out, int N, int H, int W, int C, int K,
                          int stride, int pad, int Ho, int Wo, int backward, int dtype, cudaStream_t stream);
B200DIST_API int b200_avgpool(const void* in, void* out, int N, int HW, int C, int backward, int dtype, cudaStream_t stream);
B200DIST_API int b200_pixel_shuffle2(const void* src, void* dst, int N, int H, int W, int K, int dir, int dtype,
                                 cudaStream_t stream);
B200DIST_API int b200_bilinear(const void* src, void* dst, int N, int Hi, int Wi, int Ho, int Wo, int C, int backward,
                           int dtype, cudaStream_t stream);
B200DIST_API int b200_lat_mse(const void* pred, const void* target, const float* wlat, float* loss_sum, void* grad,
                          const float* gscale, long long total, int H, long long h_div, int backward, int dtype,
                          cudaStream_t stream);

/* ---- Peer memory (CUDA IPC symmetric buffers) ----
 * plays the role of NCCL's buffer registration, SURVEY.md 5.8 */
B200DIST_API int b200_peer_alloc(size_t bytes, void** out);
B200DIST_API int b200_peer_free(void* p);
B200DIST_API int b200_peer_export(void* p, unsigned char* handle64);
B200DIST_API int b200_peer_import(const unsigned char* handle64, void** out);
B200DIST_API int b200_peer_close(void* p);
/* copy-engine (DMA) transfer between local and imported peer memory: the FSDP parameter all-gather path, no SMs used */
B200DIST_API int b200_peer_memcpy(void* dst, const void* src, size_t bytes, cudaStream_t stream);

/* ---- Limit on cross-rank waits (arrive / depart / NVLS / fused-GEMM ready flags / p2p mailboxes).  NCCL, which the
 * reference relies on, tolerates arbitrary rank skew; so does this library by default: 10 minutes, then report and
 * trap.  seconds <= 0 = wait for ever; trap = 0 = report (device printf + counter) and keep waiting.  Applies to the
 * current device; b200_peer_timeout_count() can be polled from the host while a kernel is still spinning. */
B200DIST_API int b200_set_peer_timeout(double seconds, int trap);
B200DIST_API int b200_peer_timeout_count(void);

/* ---- Point-to-point signalling for pipeline stage boundaries (replaces dist.isend/irecv,
 * scripts/04_pipeline_parallel_pp/01_manual_model_split.py:108-125): copy with b200_peer_memcpy, then signal; the
 * receiver's stream waits on its own flag word.  Flags are 32-bit sequence numbers that only grow. */
B200DIST_API int b200_p2p_signal(void* remote_flag, unsigned value, cudaStream_t stream);
B200DIST_API int b200_p2p_wait(const void* local_flag, unsigned value, cudaStream_t stream);

/* ---- Collectives over NVLink peer memory ----
 * replace c10d/NCCL: tests/torch_comm_bench.py:49,79; DDP / FSDP / DTensor implicit collectives */
B200DIST_API int b200_comm_barrier(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, cudaStream_t stream);
B200DIST_API int b200_comm_allreduce(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, void* out, long long in_off, long long n, float scale,
                                 int in_dtype, int out_dtype, int algo, void* local_gate, int max_ctas,
                                 cudaStream_t stream);
B200DIST_API int b200_comm_allgather(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, void* out, long long in_off_bytes, long long bytes,
                                 int max_ctas, cudaStream_t stream);
B200DIST_API int b200_comm_reducescatter(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, void* out, long long in_off, long long n, float scale,
                                     int in_dtype, int out_dtype, int max_ctas, cudaStream_t stream);
B200DIST_API int b200_comm_rs_adamw(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, long long in_off, long long n, float grad_scale, int in_dtype,
                                float* master, void* lowp, float* m, float* v, float lr, float beta1,
                                float beta2, float eps, float wd, float bc1, float bc2, int max_ctas,
                                cudaStream_t stream);
/* local variant: slab_ptrs[r] are `world` LOCAL n-element gradient slabs (peers' slices staged by b200_peer_memcpy) */
B200DIST_API int b200_comm_rs_adamw_local(const void* const* slab_ptrs, int world, long long n, float grad_scale, int in_dtype,
                                      float* master, void* lowp, float* m, float* v, float lr, float beta1,
                                      float beta2, float eps, float wd, float bc1, float bc2, int max_ctas,
                                      cudaStream_t stream);
/* latency path: input pushed straight from the user tensor into every peer's slot buffer, no staging copy, no depart */
B200DIST_API int b200_comm_allreduce_push(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, const void* in, void* out, long long n, float scale, int in_dtype,
                                      int out_dtype, long long slot_elems, int parity, cudaStream_t stream);
/* the same latency path for reduce-scatter (in = world slices of n elements) ... */
B200DIST_API int b200_comm_reducescatter_push(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, const void* in, void* out, long long n, float scale, int in_dtype,
                                      int out_dtype, long long slot_elems, int parity, cudaStream_t stream);
/* ... and for all-gather (root < 0) / broadcast (root >= 0) of `bytes` <= slot_bytes */
B200DIST_API int b200_comm_gather_push(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, const void* in, void* out, long long bytes, long long slot_bytes,
                                      int root, int parity, cudaStream_t stream);
/* large-message broadcast: scatter from the root, then all-gather among the ranks */
B200DIST_API int b200_comm_broadcast2(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, void* out, long long in_off_bytes, long long bytes, int root,
                                  void* local_gate, int max_ctas, cudaStream_t stream);
B200DIST_API int b200_comm_broadcast(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, void* out, long long in_off_bytes, long long bytes, int root,
                                 int max_ctas, cudaStream_t stream);
B200DIST_API int b200_comm_alltoall(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, void* out, long long in_off_bytes, long long bytes_per_peer,
                                int max_ctas, cudaStream_t stream);
/* neighbour (halo) exchange along a chain of ranks: every rank reads two strips - from rank-1 and rank+1 - whatever the
 * world size (domain-parallel convolution; the reference only describes it, docs/guide/10_domain_parallel.md:69-103) */
B200DIST_API int b200_comm_neighbor_exchange(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                                             int world, int channel, unsigned seq, void* from_prev, void* from_next,
                                             long long in_off_bytes, long long bytes, int max_ctas, cudaStream_t stream);
B200DIST_API int b200_comm_slab_reduce(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank,
                  int world, int channel, unsigned seq, void* out, const void* residual, long long n, int max_ctas,
                                   cudaStream_t stream);

/* ---- NVLS: NVSwitch multicast objects + in-switch reduction (csrc/multicast.cu) ----
 * rank 0: b200_mc_create (handle + POSIX fd to pass to the peers); others: b200_mc_import(fd); all: b200_mc_add_device,
 * barrier, b200_mc_bind_map -> local unicast pointer + multicast pointer.  b200_comm_allreduce_nvls reduces the first n
 * elements of the multicast buffer in place (sum * scale); n % (world * 16 / element size) == 0.  Plays the role of
 * NCCL's NVLS algorithm behind dist.all_reduce (tests/torch_comm_bench.py:79, DDP buckets). */
B200DIST_API int b200_mc_supported(int* supported);
B200DIST_API int b200_mc_granularity(size_t bytes, int world, size_t* rounded);
B200DIST_API int b200_mc_create(size_t bytes, int world, unsigned long long* handle, int* fd);
B200DIST_API int b200_mc_import(int fd, unsigned long long* handle);
B200DIST_API int b200_mc_add_device(unsigned long long handle);
B200DIST_API int b200_mc_bind_map(unsigned long long handle, size_t bytes, void** local_ptr, void** mc_ptr, void** state);
B200DIST_API int b200_mc_release(void* state);
B200DIST_API int b200_comm_allreduce_nvls(const void* const* flag_ptrs, void* counter, int rank, int world, int channel,
                                      unsigned seq, void* mc_ptr, long long n, float scale, int dtype, int max_ctas, int split,
                                      cudaStream_t stream);
B200DIST_API int b200_comm_broadcast_nvls(const void* const* flag_ptrs, void* counter, int rank, int world, int channel,
                                      unsigned seq, const void* src, void* mc_ptr, const void* replica_ptr, void* out,
                                      long long bytes, int root, int max_ctas, cudaStream_t stream);
/* reduce-scatter / fused reduce-scatter + AdamW / all-gather through the NVSwitch (multimem.ld_reduce, multimem.st):
 * the region at mc_ptr is [world * n] elements (resp. [world * bytes]) inside a buffer from b200_mc_bind_map; replace
 * dist.reduce_scatter_tensor / dist.all_gather_into_tensor and FSDP's gradient + optimizer step */
B200DIST_API int b200_comm_reducescatter_nvls(const void* const* flag_ptrs, void* counter, int rank, int world, int channel,
                                              unsigned seq, const void* mc_ptr, void* out, long long n, float scale,
                                              int in_dtype, int out_dtype, int max_ctas, int split, cudaStream_t stream);
B200DIST_API int b200_comm_rs_adamw_nvls(const void* const* flag_ptrs, void* counter, int rank, int world, int channel,
                                         unsigned seq, const void* mc_ptr, long long n, float grad_scale, int in_dtype,
                                         float* master, void* lowp, float* m, float* v, float lr, float beta1, float beta2,
                                         float eps, float wd, float bc1, float bc2, int max_ctas, int split, cudaStream_t stream);
B200DIST_API int b200_comm_allgather_nvls(const void* const* flag_ptrs, void* counter, int rank, int world, int channel,
                                          unsigned seq, const void* src, void* mc_ptr, const void* replica_ptr, void* out,
                                          long long bytes, int max_ctas, cudaStream_t stream);

#endif  /* B200DIST_H_ */
