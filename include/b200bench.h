/* b200bench native library: C ABI.
 *
 * Every entry point takes raw device pointers and a cudaStream_t (as void*), launches
 * asynchronously on that stream and returns 0 on success, a cudaError_t value (> 0) for a CUDA
 * failure, or a negative number for an argument error.  Unless stated otherwise no entry point
 * allocates memory.  bf16 = __nv_bfloat16, row-major everywhere.
 *
 * Citations are path:line under the reference repository
 * (deepaksatna/Distributed-LLM-Training-Benchmark-Framework); TH = benchmarking/train_harness.py.
 */
#ifndef B200BENCH_H_
#define B200BENCH_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* GEMM epilogues */
#define B200_EPI_NONE 0       /* D = acc                                         */
#define B200_EPI_BIAS 1       /* D = acc + bias[n]                               */
#define B200_EPI_BIAS_GELU 2  /* D = acc + bias ; D2 = gelu_erf(D)   (TH:119-120) */
#define B200_EPI_BIAS_RESID 3 /* D = aux + dropout(acc + bias)       (TH:121-122,128,130) */
#define B200_EPI_DGELU 4      /* D = acc * gelu_erf'(aux)            (backward of TH:120) */
#define B200_EPI_ACCUM 5      /* D = D + acc                         (wgrad accumulation) */
#define B200_EPI_BIAS_GELU_GRAD 7 /* D = gelu_erf'(acc + bias) ; D2 = gelu_erf(acc + bias): the forward saves the
                                     derivative instead of the pre-activation, so that the backward epilogue
                                     (B200_EPI_MUL) is one multiply instead of an erf and an exp per element */
#define B200_EPI_MUL 8        /* D = acc * aux                       (backward of TH:120 with the saved derivative) */
#define B200_EPI_ROWDOT 6     /* D = acc ; rowdot[(b*H+h)*T+t] = sum over head h of acc*aux   (attention-backward
                                 D = rowsum(dO o O), produced by the out-proj dgrad that writes dO) */

/* D[M,N] = epi(A * B^T), tcgen05/TMEM/TMA.  a_mn / b_mn: 0 = operand stored [rows, K]
 * (K-major), 1 = stored [K, rows] (MN-major).  lda/ldb/ldd in elements, multiples of 8.
 * bn_hint: 0 (auto), 128 / 256 (single-CTA tile width) or 2128 / 2256 (CTA-pair cta_group::2 mode,
 * 256 x 128 / 256 x 256 tiles per pair).  max_ctas: 0 = one CTA per SM, else caps the persistent grid
 * (used to leave SMs free for an overlapped communication kernel).
 * Replaces cuBLAS under nn.Linear / MultiheadAttention projections (TH:58,113-123,127). */
int b200_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int lda, int ldb,
                   int ldd, int a_mn, int b_mn, int epi, const void* bias, const void* aux,
                   void* D2, float drop_p, unsigned int drop_seed, int bn_hint, int max_ctas,
                   void* stream);

/* ---- attention (TH:127) ------------------------------------------------------
 * qkv [B*T, 3*H*hd] bf16 (columns q|k|v, head h at h*hd in each), out [B*T, H*hd] bf16,
 * lse [B*H, T] fp32 (natural log).  head_dim must be 64.  causal=0 reproduces the reference
 * (no mask).  drop_p > 0 drops attention probabilities after the softmax (counter-based RNG,
 * the backward regenerates the mask from the same seed). */
int b200_attention_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int head_dim,
                       int causal, float drop_p, unsigned int seed, void* stream);
/* Forward variant: 1 (default) = O accumulates in tensor memory and is rescaled only when a block raises the row
 * maximum by more than 2^8; 0 = O in registers with the exact running-max rescale after every block.  Returns the
 * previous setting; a negative argument only queries.  Environment: B200BENCH_ATTN_FWD=exact starts with 0. */
int b200_set_attention_fwd_variant(int lazy);
/* dqkv [B*T, 3*H*hd] bf16 receives dQ|dK|dV.  D [B*H, T] fp32 and dq_acc [B*T, H*hd] fp32 are
 * caller-provided scratch (dq_acc is zeroed inside). */
int b200_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse,
                       void* dqkv, float* D, float* dq_acc, int B, int T, int H, int head_dim,
                       int causal, float drop_p, unsigned int seed, void* stream);

/* Same, with the two helper passes optional.  flags:
 *   B200_ATTN_BWD_HAVE_D       D already holds rowsum(dO o O) per (batch, head, query) - e.g. written by the out-proj
 *                              dgrad GEMM with B200_EPI_ROWDOT - so the preparation kernel is skipped;
 *   B200_ATTN_BWD_ACC_IS_ZERO  dq_acc is all zeros on entry (every call leaves it zeroed again), so it is not cleared.
 * With both flags the backward is two launches: the main kernel and the fp32 -> bf16 dQ conversion. */
#define B200_ATTN_BWD_HAVE_D 1
#define B200_ATTN_BWD_ACC_IS_ZERO 2
int b200_attention_bwd_ex(const void* qkv, const void* out, const void* dout, const float* lse,
                          void* dqkv, float* D, float* dq_acc, int B, int T, int H, int head_dim,
                          int causal, float drop_p, unsigned int seed, void* stream, int flags);

/* ---- LayerNorm (TH:113,117,56) ------------------------------------------------ */
int b200_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean,
                       float* rstd, int M, int d, float eps, void* stream);
/* dx = LayerNorm'(dy) + dres (dres may be NULL) */
int b200_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                       const float* rstd, const void* dres, void* dx, int M, int d, void* stream);
/* LayerNorm backward in one pass over dy and x: dx (+ dres), optionally dx_drop = dropout_mask(drop_seed) * dx /
 * (1 - drop_p) with the element indexing of the GEMM dropout epilogue (row * d + col), and optionally (partials !=
 * NULL) per-CTA fp32 partial column sums of dy*xhat and dy, b200_layernorm_bwd_fused_scratch_bytes(M, d) bytes.
 * b200_layernorm_param_fold turns those partials into dgamma / dbeta (accumulate != 0 adds onto them); it only
 * feeds the optimizer, so it can run on another stream (after the producer, before the partials are reused).
 * Both steps sum in a fixed order (no atomics). */
size_t b200_layernorm_bwd_fused_scratch_bytes(int M, int d);
int b200_layernorm_bwd_fused(const void* dy, const void* x, const void* gamma, const float* mean,
                             const float* rstd, const void* dres, void* dx, void* dx_drop, float drop_p,
                             unsigned int drop_seed, void* partials, size_t partials_bytes, int M, int d,
                             void* stream);
int b200_layernorm_param_fold(const void* partials, int M, int d, void* dgamma, void* dbeta, int accumulate,
                              void* stream);
/* Scratch needed by the two column reductions below: 256 per-column-block completion
 * counters followed by fp32 partial sums.  The buffer must be zero-filled once before its first use; every call
 * leaves the counters zeroed again.  Calls sharing one scratch buffer must be stream-ordered. */
size_t b200_colreduce_scratch_bytes(int M, int N, int outputs);
/* dgamma[n] (+)= sum_m dy*xhat, dbeta[n] (+)= sum_m dy.  scratch >= colreduce_scratch_bytes(M,d,2) */
int b200_layernorm_bwd_params(const void* dy, const void* x, const float* mean, const float* rstd,
                              void* dgamma, void* dbeta, int M, int d, int accumulate,
                              void* scratch, size_t scratch_bytes, void* stream);
/* out[n] (+)= sum_m dy[m,n]   (bias gradients).  scratch >= colreduce_scratch_bytes(M,N,1) */
int b200_colsum(const void* dy, void* out, int M, int N, int ld, int accumulate, void* scratch,
                size_t scratch_bytes, void* stream);

/* Same result as b200_colsum, for callers that cannot allocate (no counterpart in the reference, which
 * has no native entry points at all): a plug-in inside a framework whose pool already owns the GPU's
 * memory, or a loop that must stay clear of cudaMalloc's driver locks and device synchronisation.
 * The scratch is a 512 KiB device array built into the library (held by every GPU that loads it until
 * exit), so no buffer is passed and nothing is allocated or freed.  The FIRST call on each device
 * looks up the array's device address; make it before any stream capture.  (The per-device address cache is
 * written without a lock: two host threads making that first call on one device at the same moment both
 * store the same value, which is harmless.)  All calls on a GPU share
 * the one array: keep them stream-ordered, overlapping calls race.  Limit: N <= 65536 columns and
 * scratch <= b200_colsum_static_max_bytes(); larger shapes return B200_ERR_STATIC_SCRATCH. */
#define B200_ERR_STATIC_SCRATCH (-8)
size_t b200_colsum_static_max_bytes(void);
int b200_colsum_static(const void* dy, void* out, int M, int N, int ld, int accumulate, void* stream);

/* ---- embedding (TH:85-89) ------------------------------------------------------ */
int b200_embedding_fwd(const int* tokens, const void* wte, const void* wpe, void* out, int B,
                       int T, int d, float drop_p, unsigned int seed, void* stream);
/* dwte[tok] += sum of dx over the positions holding tok: summed in fp32 by the one CTA that owns the token and added
 * once to the row (on top of the tied lm_head wgrad) - no atomics, deterministic; dwpe[t] (+)= sum_b dx */
int b200_embedding_bwd(const int* tokens, const void* dx, void* dwte, void* dwpe, int B, int T,
                       int d, float drop_p, unsigned int seed, int accumulate_wpe, void* stream);

/* ---- softmax cross-entropy forward+backward (TH:99-103) --------------------------
 * logits [M, ld] bf16 are overwritten by dlogits * grad_scale / n_valid when write_grad != 0.
 * loss_out (device scalar) = mean over rows with target != ignore_index, times loss_scale. */
int b200_cross_entropy(void* logits, const int* targets, float* loss_rows, int* n_valid,
                       float* loss_out, int M, int V, int ld, int ignore_index, float grad_scale,
                       int write_grad, int accumulate_loss, float loss_scale, void* stream);

/* ---- optimizer (TH:329, configs/deepspeed/zero2.json:2,27-44) -------------------
 * hp: 9 floats on the device {lr, beta1, beta2, eps, wd, 1/bc1, 1/sqrt(bc2), grad_scale, clip}.
 * gnorm_sq (nullable): squared global grad norm for clipping. n % 8 == 0. */
int b200_adamw_set_hp(float* hp, float lr, float beta1, float beta2, float eps, float wd,
                      float inv_bias_corr1, float inv_sqrt_bias_corr2, float grad_scale,
                      float clip_norm, void* stream);
int b200_adamw(float* master, float* m, float* v, const void* grad, void* w_out, size_t n,
               const float* hp, const float* gnorm_sq, void* stream);
/* *out += scale^2 * sum(grad^2) */
int b200_sumsq(const void* grad, size_t n, float scale, float* out, void* stream);
/* out = dropout_mask(seed) * in / (1-p): re-applies the mask a GEMM epilogue drew (same indexing) */
int b200_dropout_apply(const void* in, void* out, int M, int N, float drop_p, unsigned int seed,
                       void* stream);

/* ---- peer memory + fused collectives over NVLink / NVSwitch -----------
 * b200_peer_alloc cudaMallocs a zero-filled buffer and returns its 64-byte CUDA IPC handle;
 * b200_peer_open maps a peer's handle.  These two are the only entry points that allocate. */
int b200_peer_alloc(size_t bytes, void** ptr, void* handle64);
int b200_peer_open(const void* handle64, void** ptr);
int b200_peer_close(void* ptr);
int b200_peer_free(void* ptr);
size_t b200_comm_flag_bytes(void);
/* In the calls below `grads` / `weights` / `flags` are host arrays of ws device pointers: entry r is
 * rank r's buffer as mapped in THIS process.  All ranks must issue the same sequence of calls.
 * K13: grads[off, off+n) = scale * sum over ranks, in place on every rank (DDP, TH:217-222,374). */
int b200_comm_allreduce(void* const* grads, void* const* flags, int rank, int ws, size_t off,
                        size_t n, float scale, int ctas, void* stream);
/* K15: this rank's partition [lo, lo+n): g = sum over ranks of grads; optional global-norm clip
 * (hp[8]); AdamW on fp32 master/m/v; bf16 weights stored to every rank (push_all=1, ZeRO-2) or to
 * the own buffer only (ZeRO-3 / FSDP).  gshard: n bf16 of local scratch. hp as in b200_adamw.
 * gnorm_out (nullable, TWO floats): norm and squared norm; extra_norm_sq (nullable): squared norm of
 * gradients that are updated outside this kernel (the replicated small tensors under ZeRO-3). */
/* Every CTA count is clamped to what the device can hold at once (occupancy x SMs); K15 is launched cooperatively
 * because its clipping path has a barrier over the whole grid: a grid that cannot be co-resident is refused with a
 * CUDA error instead of spinning. */
int b200_comm_rs_adamw(void* const* grads, void* const* weights, void* const* flags, int rank,
                       int ws, size_t lo, size_t n, float* master, float* m, float* v, void* gshard,
                       const float* hp, int push_all, float* gnorm_out, const float* extra_norm_sq,
                       int ctas, void* stream);
/* K13 / K15 through an NVSwitch multicast (NVLS) mapping.  mc_base / weights_mc = THIS rank's multicast address of
 * the buffer (every rank's physical copy bound to one multicast object, e.g. by torch's symmetric memory):
 * multimem.ld_reduce sums the ranks' copies inside the switch, multimem.st writes all copies with one store. */
int b200_comm_allreduce_nvls(void* mc_base, void* const* flags, int rank, int ws, size_t off, size_t n,
                             float scale, int ctas, void* stream);
int b200_comm_rs_adamw_nvls(void* const* grads, void* const* weights, void* weights_mc, void* const* flags,
                            int rank, int ws, size_t lo, size_t n, float* master, float* m, float* v, void* gshard,
                            const float* hp, int push_all, float* gnorm_out, const float* extra_norm_sq, int ctas,
                            void* stream);
/* Single-GPU test mode of K13 / K15: ONE cooperative launch plays all ws ranks (gridDim.y = ws, rank = blockIdx.y)
 * over per-rank buffers that all live on the current device; grads[r] / weights[r] / flags[r] and the per-rank
 * arrays (length ws) belong to emulated rank r.  Separate launches per rank on one GPU are NOT an option: kernels
 * that wait for one another are not guaranteed to run at the same time.  Returns -9 if the device cannot hold
 * the whole grid at once. */
int b200_comm_allreduce_loop(void* const* grads, void* const* flags, int ws, size_t off, size_t n, float scale,
                             int ctas, void* stream);
int b200_comm_rs_adamw_loop(void* const* grads, void* const* weights, void* const* flags, int ws, const size_t* lo,
                            size_t n, float* const* master, float* const* m, float* const* v, void* const* gshard,
                            const float* const* hp, int push_all, float* const* gnorm_out,
                            const float* const* extra_norm_sq, int ctas, void* stream);
/* GEMM launches use programmatic dependent launch (their prologue overlaps the previous kernel's tail;
 * the kernel waits with griddepcontrol.wait before touching global memory).  0 turns it off. */
int b200_set_gemm_pdl(int on);
/* Extended GEMM arguments (all optional; a zeroed struct = b200_gemm_bf16).
 *  dbias / dbias_accum   weight-gradient GEMMs (epi NONE or ACCUM): dbias[m] (+)= sum_k A[m,k], the bias gradient,
 *                        from a 16-column row-sum MMA against an all-ones tile in the same main loop (replaces a
 *                        separate column-reduction kernel over dY).
 *  rowdot, rd_T, rd_hd   for B200_EPI_ROWDOT (rd_hd = 64 or 128 columns per head, M a multiple of rd_T).
 *  b_shards, b_ws, b_flat_off, b_shard_elems
 *                        K14: B is a weight matrix that is partitioned over b_ws ranks by flat offset
 *                        (partition size b_shard_elems, B[0,0] at flat offset b_flat_off; both multiples of ldb).
 *                        b_shards[r] = base of rank r's partition as mapped in THIS process (peer memory).  The B
 *                        argument itself is ignored: every B tile is fetched by TMA straight from its owner.
 *                        Replaces FSDP's pre-forward all-gather / ZeRO-3's per-module gathers (TH:229-237,366;
 *                        SURVEY N4,N8).  Returns -9 when no tile shape keeps every TMA box inside one owner.
 *  mc_units              tile units (CTAs / CTA pairs along M) per thread-block cluster that share each B tile by
 *                        TMA multicast: 0 / 1 = off, 2 / 4 / 8 = that many.  Local B only: with a partitioned B
 *                        (b_ws > 1) any value > 1 is rejected (-2).  TMA multicast loads here read only this GPU's
 *                        memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU. */
typedef struct B200GemmEx {
  void* dbias;
  int dbias_accum;
  float* rowdot;
  int rd_T, rd_hd;
  const void* const* b_shards;
  int b_ws;
  unsigned long long b_flat_off, b_shard_elems;
  int mc_units;
} B200GemmEx;
int b200_gemm_bf16_ex(const void* A, const void* B, void* D, int M, int N, int K, int lda, int ldb,
                      int ldd, int a_mn, int b_mn, int epi, const void* bias, const void* aux, void* D2,
                      float drop_p, unsigned int drop_seed, int bn_hint, int max_ctas, void* stream,
                      const B200GemmEx* ex);
/* Encoded TMA descriptors are cached by (address, shape, box); counters for tests and diagnostics. */
int b200_tmap_cache_stats(unsigned long long* hits, unsigned long long* misses);
/* Plain gather of the flat range [flat_off, flat_off+n) from the owners' partitions into dst.  The _from variant
 * takes the calling rank and starts rank/ws of the way into the range (wrapping around), so that ranks that gather
 * the same range at the same time read different owners at any moment instead of all hitting one GPU's links. */
int b200_comm_gather_range_from(void* const* shard_adj, int ws, int rank, void* dst, unsigned long long flat_off,
                                unsigned long long n, unsigned long long shard_elems, int ctas, void* stream);
int b200_comm_gather_range(void* const* shard_adj, int ws, void* dst, unsigned long long flat_off,
                           unsigned long long n, unsigned long long shard_elems, int ctas,
                           void* stream);
/* K16: the same gather as ONE collective.  Rank r reads piece r (1/ws) of the range from its owner(s) and writes it
 * at element offset dst_off into every rank's copy of the destination: one multimem.st through dst_mc (this rank's
 * NVSwitch multicast address of the destination buffer) or, with dst_mc == NULL, ws peer stores through dst_tab.
 * The owner sends the range once in total instead of once per reader.  Two rank barriers on flag plane 2: every
 * rank must issue the same sequence of K16 calls.  _loop: test mode, one cooperative launch plays all ranks. */
int b200_comm_gather_bcast(void* const* shard_adj, void* const* dst_tab, void* dst_mc, void* const* flags, int rank,
                           int ws, unsigned long long flat_off, unsigned long long dst_off, unsigned long long n,
                           unsigned long long shard_elems, int ctas, void* stream);
int b200_comm_gather_bcast_loop(void* const* shard_adj, void* const* dst_tab, void* const* flags, int ws,
                                unsigned long long flat_off, unsigned long long dst_off, unsigned long long n,
                                unsigned long long shard_elems, int ctas, void* stream);
/* dst[i] = src[i] with 16-byte volatile loads: P2P bandwidth probe when src is peer memory. */
int b200_peer_copy(const void* src, void* dst, size_t bytes, int ctas, void* stream);

/* Per-step dropout seed: a device word XORed into every dropout seed.  Register once, then launch
 * b200_step_seed_advance at the start of each micro-step (capturable in a CUDA graph). */
int b200_step_seed_register(unsigned int* device_word);
int b200_step_seed_advance(void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200BENCH_H_ */
