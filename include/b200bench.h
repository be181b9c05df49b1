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

/* D[M,N] = epi(A * B^T), tcgen05/TMEM/TMA.  a_mn / b_mn: 0 = operand stored [rows, K]
 * (K-major), 1 = stored [K, rows] (MN-major).  lda/ldb/ldd in elements, multiples of 8.
 * bn_hint: 128, 256 or 0 (auto).  max_ctas: 0 = one CTA per SM, else caps the persistent grid
 * (used to leave SMs free for an overlapped communication kernel).
 * Replaces cuBLAS under nn.Linear / MultiheadAttention projections (TH:58,113-123,127). */
int b200_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int lda, int ldb,
                   int ldd, int a_mn, int b_mn, int epi, const void* bias, const void* aux,
                   void* D2, float drop_p, unsigned int drop_seed, int bn_hint, int max_ctas,
                   void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200BENCH_H_ */
