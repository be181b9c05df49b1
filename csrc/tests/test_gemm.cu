// Standalone numerics + timing harness for b200_gemm_bf16 (no torch needed, starts in < 1 s).
// Usage: test_gemm [--time | small]   (small: numerics on the small shapes only, for compute-sanitizer)
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "../../include/b200bench.h"

#define CK(x)                                                                 \
  do {                                                                        \
    cudaError_t e_ = (x);                                                     \
    if (e_ != cudaSuccess) {                                                  \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(2);                                                                \
    }                                                                         \
  } while (0)

static float bf(const __nv_bfloat16& v) { return __bfloat162float(v); }
static uint32_t rng_state = 12345;
static float frand() {
  rng_state = rng_state * 1664525u + 1013904223u;
  return ((rng_state >> 8) & 0xFFFF) / 65536.0f - 0.5f;
}
static double gelu(double x) { return 0.5 * x * (1.0 + erf(x * 0.7071067811865476)); }
static double dgelu(double x) {
  return 0.5 * (1.0 + erf(x * 0.7071067811865476)) + x * 0.3989422804014327 * exp(-0.5 * x * x);
}

struct Case {
  int M, N, K, a_mn, b_mn, epi, bn;
};

static int run_case(const Case& c) {
  const int M = c.M, N = c.N, K = c.K;
  // storage shapes
  const size_t a_elems = (size_t)M * K, b_elems = (size_t)N * K, d_elems = (size_t)M * N;
  std::vector<__nv_bfloat16> hA(a_elems), hB(b_elems), hBias(N), hAux(d_elems), hD(d_elems),
      hD2(d_elems), hD0(d_elems);
  for (auto& v : hA) v = __float2bfloat16(frand());
  for (auto& v : hB) v = __float2bfloat16(frand());
  for (auto& v : hBias) v = __float2bfloat16(frand());
  for (auto& v : hAux) v = __float2bfloat16(frand() * 2.f);
  for (auto& v : hD0) v = __float2bfloat16(frand());
  __nv_bfloat16 *dA, *dB, *dBias, *dAux, *dD, *dD2;
  CK(cudaMalloc(&dA, a_elems * 2));
  CK(cudaMalloc(&dB, b_elems * 2));
  CK(cudaMalloc(&dBias, N * 2));
  CK(cudaMalloc(&dAux, d_elems * 2));
  CK(cudaMalloc(&dD, d_elems * 2));
  CK(cudaMalloc(&dD2, d_elems * 2));
  CK(cudaMemcpy(dA, hA.data(), a_elems * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, hB.data(), b_elems * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dBias, hBias.data(), N * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dAux, hAux.data(), d_elems * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dD, hD0.data(), d_elems * 2, cudaMemcpyHostToDevice));
  CK(cudaMemset(dD2, 0, d_elems * 2));
  const int lda = c.a_mn ? M : K, ldb = c.b_mn ? N : K;
  int rc = b200_gemm_bf16(dA, dB, dD, M, N, K, lda, ldb, N, c.a_mn, c.b_mn, c.epi, dBias, dAux,
                          dD2, 0.f, 0, c.bn, 0, nullptr);
  if (rc) {
    printf("  launch rc=%d\n", rc);
    return 1;
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("  kernel error: %s\n", cudaGetErrorString(e));
    exit(3);  // sticky error: stop the whole harness
  }
  CK(cudaMemcpy(hD.data(), dD, d_elems * 2, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hD2.data(), dD2, d_elems * 2, cudaMemcpyDeviceToHost));
  // sampled check
  const int samples = (d_elems <= 65536) ? (int)d_elems : 20000;
  double max_err = 0, max_ref = 0;
  int bad = 0;
  for (int s = 0; s < samples; ++s) {
    size_t idx = (d_elems <= 65536) ? (size_t)s
                                    : ((size_t)(rng_state = rng_state * 1664525u + 1013904223u) *
                                           2654435761ull % d_elems);
    int m = idx / N, n = idx % N;
    if (s < 64 && d_elems > 65536) {  // force corners / tile edges
      m = (s & 1) ? M - 1 - (s >> 1) % 7 : (s >> 1) % 130;
      n = (s & 2) ? N - 1 - (s >> 2) % 9 : (s >> 2) % 260 % N;
      idx = (size_t)m * N + n;
    }
    double acc = 0;
    for (int k = 0; k < K; ++k) {
      float a = c.a_mn ? bf(hA[(size_t)k * M + m]) : bf(hA[(size_t)m * K + k]);
      float b = c.b_mn ? bf(hB[(size_t)k * N + n]) : bf(hB[(size_t)n * K + k]);
      acc += (double)a * b;
    }
    double ref = acc, ref2 = 0;
    switch (c.epi) {
      case B200_EPI_BIAS: ref = acc + bf(hBias[n]); break;
      case B200_EPI_BIAS_GELU:
        ref = acc + bf(hBias[n]);
        ref2 = gelu(ref);
        break;
      case B200_EPI_BIAS_RESID: ref = acc + bf(hBias[n]) + bf(hAux[idx]); break;
      case B200_EPI_DGELU: ref = acc * dgelu(bf(hAux[idx])); break;
      case B200_EPI_ACCUM: ref = acc + bf(hD0[idx]); break;
      default: break;
    }
    double got = bf(hD[idx]);
    double err = fabs(got - ref);
    double tol = 0.02 + 0.01 * fabs(ref);
    if (err > tol) {
      if (bad < 5) printf("    mismatch (%d,%d): got %f ref %f\n", m, n, got, ref);
      ++bad;
    }
    if (c.epi == B200_EPI_BIAS_GELU) {
      double e2 = fabs(bf(hD2[idx]) - ref2);
      if (e2 > 0.02 + 0.01 * fabs(ref2)) {
        if (bad < 5) printf("    mismatch D2 (%d,%d): got %f ref %f\n", m, n, bf(hD2[idx]), ref2);
        ++bad;
      }
    }
    if (err > max_err) max_err = err;
    if (fabs(ref) > max_ref) max_ref = fabs(ref);
  }
  printf("  M=%d N=%d K=%d a_mn=%d b_mn=%d epi=%d bn=%d : max_err=%.4f (max_ref=%.2f) bad=%d %s\n",
         M, N, K, c.a_mn, c.b_mn, c.epi, c.bn, max_err, max_ref, bad, bad ? "FAIL" : "ok");
  cudaFree(dA); cudaFree(dB); cudaFree(dBias); cudaFree(dAux); cudaFree(dD); cudaFree(dD2);
  return bad ? 1 : 0;
}

static void time_case(int M, int N, int K, int a_mn, int b_mn, int epi, int bn, const char* name) {
  // rotate through enough distinct operand sets to exceed the 126 MB L2
  const size_t per = ((size_t)M * K + (size_t)N * K + 2 * (size_t)M * N) * 2 + N * 2;
  int nsets = (int)(300e6 / per) + 1;
  if (nsets > 64) nsets = 64;
  std::vector<__nv_bfloat16*> A(nsets), B(nsets), D(nsets), D2(nsets), X(nsets);
  __nv_bfloat16* bias;
  CK(cudaMalloc(&bias, N * 2));
  CK(cudaMemset(bias, 0, N * 2));
  for (int i = 0; i < nsets; ++i) {
    CK(cudaMalloc(&A[i], (size_t)M * K * 2));
    CK(cudaMalloc(&B[i], (size_t)N * K * 2));
    CK(cudaMalloc(&D[i], (size_t)M * N * 2));
    CK(cudaMalloc(&D2[i], (size_t)M * N * 2));
    CK(cudaMalloc(&X[i], (size_t)M * N * 2));
    CK(cudaMemset(A[i], 0x11, (size_t)M * K * 2));
    CK(cudaMemset(B[i], 0x11, (size_t)N * K * 2));
    CK(cudaMemset(X[i], 0, (size_t)M * N * 2));
  }
  const int lda = a_mn ? M : K, ldb = b_mn ? N : K;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int i = 0; i < 5; ++i)
    b200_gemm_bf16(A[i % nsets], B[i % nsets], D[i % nsets], M, N, K, lda, ldb, N, a_mn, b_mn, epi,
                   bias, X[i % nsets], D2[i % nsets], 0.f, 0, bn, 0, nullptr);
  CK(cudaDeviceSynchronize());
  const int iters = 40;
  CK(cudaEventRecord(e0));
  for (int i = 0; i < iters; ++i)
    b200_gemm_bf16(A[i % nsets], B[i % nsets], D[i % nsets], M, N, K, lda, ldb, N, a_mn, b_mn, epi,
                   bias, X[i % nsets], D2[i % nsets], 0.f, 0, bn, 0, nullptr);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double us = ms * 1000.0 / iters;
  const double tflops = 2.0 * M * N * K / (us * 1e-6) / 1e12;
  printf("  time %-28s M=%d N=%d K=%d bn=%d: %.2f us  %.1f TFLOP/s (%d operand sets, >L2)\n", name,
         M, N, K, bn, us, tflops, nsets);
  for (int i = 0; i < nsets; ++i) {
    cudaFree(A[i]); cudaFree(B[i]); cudaFree(D[i]); cudaFree(D2[i]); cudaFree(X[i]);
  }
  cudaFree(bias);
}

int main(int argc, char** argv) {
  int fails = 0;
  printf("== gemm numerics ==\n");
  const Case cases[] = {
      {128, 128, 64, 0, 0, B200_EPI_NONE, 128},
      {128, 128, 256, 0, 0, B200_EPI_NONE, 128},
      {256, 256, 512, 0, 0, B200_EPI_BIAS, 128},
      {256, 256, 512, 0, 0, B200_EPI_BIAS, 256},
      {256, 384, 192, 0, 1, B200_EPI_NONE, 128},
      {256, 384, 192, 1, 1, B200_EPI_NONE, 128},
      {256, 512, 192, 1, 1, B200_EPI_ACCUM, 256},
      {200, 136, 72, 0, 0, B200_EPI_BIAS_GELU, 128},
      {200, 136, 72, 0, 1, B200_EPI_DGELU, 128},
      {200, 136, 72, 1, 1, B200_EPI_BIAS_RESID, 128},
      {256, 256, 512, 0, 0, B200_EPI_BIAS, 2128},
      {512, 512, 256, 0, 0, B200_EPI_BIAS, 2256},
      {512, 384, 192, 0, 1, B200_EPI_NONE, 2128},
      {512, 512, 192, 1, 1, B200_EPI_ACCUM, 2256},
      {328, 264, 72, 0, 0, B200_EPI_BIAS_GELU, 2256},
      {2048, 3072, 1024, 0, 0, B200_EPI_BIAS, 2256},
      {2048, 1024, 4096, 0, 1, B200_EPI_DGELU, 2128},
      {4096, 1024, 2048, 1, 1, B200_EPI_NONE, 2256},
      {2048, 3072, 1024, 0, 0, B200_EPI_BIAS, 0},
      {2048, 1024, 4096, 0, 0, B200_EPI_BIAS_RESID, 0},
      {2048, 4096, 1024, 0, 0, B200_EPI_BIAS_GELU, 0},
      {2048, 1024, 4096, 0, 1, B200_EPI_DGELU, 0},
      {4096, 1024, 2048, 1, 1, B200_EPI_NONE, 0},
      {2048, 32000, 1024, 0, 0, B200_EPI_NONE, 0},
      {32000, 1024, 2048, 1, 1, B200_EPI_NONE, 0},
  };
  const bool small = argc > 1 && argv[1][0] == 's';
  for (const Case& c : cases) {
    if (small && (long long)c.M * c.N * c.K > (1ll << 27)) continue;
    fails += run_case(c);
  }
  printf("== gemm numerics: %d failing case(s) ==\n", fails);
  if (argc > 1 && !small) {
    printf("== gemm timing ==\n");
    time_case(2048, 3072, 1024, 0, 0, B200_EPI_BIAS, 0, "qkv fwd (auto)");
    time_case(2048, 3072, 1024, 0, 0, B200_EPI_BIAS, 2128, "qkv fwd pair");
    time_case(2048, 3072, 1024, 0, 0, B200_EPI_BIAS, 2256, "qkv fwd pair");
    time_case(2048, 1024, 4096, 0, 0, B200_EPI_BIAS_RESID, 2128, "fc2 fwd pair");
    time_case(2048, 1024, 4096, 0, 0, B200_EPI_BIAS_RESID, 2256, "fc2 fwd pair");
    time_case(2048, 4096, 1024, 0, 0, B200_EPI_BIAS_GELU, 2256, "fc1 fwd pair");
    time_case(2048, 4096, 1024, 0, 1, B200_EPI_DGELU, 2256, "fc2 dgrad pair");
    time_case(4096, 1024, 2048, 1, 1, B200_EPI_NONE, 2128, "fc1 wgrad pair");
    time_case(4096, 1024, 2048, 1, 1, B200_EPI_NONE, 2256, "fc1 wgrad pair");
    time_case(8192, 8192, 8192, 0, 0, B200_EPI_NONE, 2256, "square 8k pair");
    time_case(2048, 32000, 1024, 0, 0, B200_EPI_NONE, 2256, "lm_head fwd pair");
    time_case(2048, 3072, 1024, 0, 0, B200_EPI_BIAS, 128, "qkv fwd");
    time_case(2048, 3072, 1024, 0, 0, B200_EPI_BIAS, 256, "qkv fwd");
    time_case(2048, 1024, 1024, 0, 0, B200_EPI_BIAS_RESID, 128, "proj fwd");
    time_case(2048, 4096, 1024, 0, 0, B200_EPI_BIAS_GELU, 128, "fc1 fwd");
    time_case(2048, 4096, 1024, 0, 0, B200_EPI_BIAS_GELU, 256, "fc1 fwd");
    time_case(2048, 1024, 4096, 0, 0, B200_EPI_BIAS_RESID, 128, "fc2 fwd");
    time_case(2048, 4096, 1024, 0, 1, B200_EPI_DGELU, 128, "fc2 dgrad");
    time_case(1024, 4096, 2048, 1, 1, B200_EPI_NONE, 128, "fc2 wgrad");
    time_case(4096, 1024, 2048, 1, 1, B200_EPI_NONE, 128, "fc1 wgrad");
    time_case(2048, 32000, 1024, 0, 0, B200_EPI_NONE, 128, "lm_head fwd");
    time_case(2048, 32000, 1024, 0, 0, B200_EPI_NONE, 256, "lm_head fwd");
    time_case(8192, 8192, 8192, 0, 0, B200_EPI_NONE, 128, "square 8k");
    time_case(8192, 8192, 8192, 0, 0, B200_EPI_NONE, 256, "square 8k");
  }
  return fails ? 1 : 0;
}
