// Standalone numerics + timing harness for b200_gemm_bf16 (no torch needed, starts in < 1 s).
// Usage: test_gemm [--time | small]   (small: numerics on the small shapes only, for compute-sanitizer)
#include "clock_sampler.h"
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

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
  int units = 0;     // mc_units: tile units per cluster sharing B by TMA multicast (0 = auto)
  int owners = 0;    // > 1: B is split over this many separate allocations ("ranks") and fetched per owner (K14)
  int dbias = 0;     // 1: also produce dbias[m] = sum_k A[m,k]; 2: accumulate onto a previous value
  int rd_hd = 0;     // B200_EPI_ROWDOT head width (64 / 128); rd_T = M / 2 (two "batch" elements)
};

static int run_case(const Case& c) {
  const int M = c.M, N = c.N, K = c.K;
  // storage shapes
  const size_t a_elems = (size_t)M * K, b_elems = (size_t)N * K, d_elems = (size_t)M * N;
  std::vector<__nv_bfloat16> hA(a_elems), hB(b_elems), hBias(N), hAux(d_elems), hD(d_elems),
      hD2(d_elems), hD0(d_elems), hDb0(M), hDb(M);
  for (auto& v : hA) v = __float2bfloat16(frand());
  for (auto& v : hB) v = __float2bfloat16(frand());
  for (auto& v : hBias) v = __float2bfloat16(frand());
  for (auto& v : hAux) v = __float2bfloat16(frand() * 2.f);
  for (auto& v : hD0) v = __float2bfloat16(frand());
  for (auto& v : hDb0) v = __float2bfloat16(frand() * 4.f);
  __nv_bfloat16 *dA, *dB, *dBias, *dAux, *dD, *dD2, *dDb;
  float* dRd = nullptr;
  CK(cudaMalloc(&dA, a_elems * 2));
  CK(cudaMalloc(&dB, b_elems * 2));
  CK(cudaMalloc(&dBias, N * 2));
  CK(cudaMalloc(&dAux, d_elems * 2));
  CK(cudaMalloc(&dD, d_elems * 2));
  CK(cudaMalloc(&dD2, d_elems * 2));
  CK(cudaMalloc(&dDb, M * 2));
  CK(cudaMemcpy(dA, hA.data(), a_elems * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, hB.data(), b_elems * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dBias, hBias.data(), N * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dAux, hAux.data(), d_elems * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dD, hD0.data(), d_elems * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dDb, hDb0.data(), M * 2, cudaMemcpyHostToDevice));
  CK(cudaMemset(dD2, 0, d_elems * 2));
  const int lda = c.a_mn ? M : K, ldb = c.b_mn ? N : K;
  B200GemmEx ex;
  memset(&ex, 0, sizeof(ex));
  ex.mc_units = c.units;
  if (c.dbias) {
    ex.dbias = dDb;
    ex.dbias_accum = c.dbias == 2;
  }
  const int heads = c.rd_hd ? N / c.rd_hd : 0, rd_T = M / 2;
  if (c.rd_hd) {
    CK(cudaMalloc(&dRd, (size_t)2 * heads * rd_T * 4));
    CK(cudaMemset(dRd, 0, (size_t)2 * heads * rd_T * 4));
    ex.rowdot = dRd;
    ex.rd_T = rd_T;
    ex.rd_hd = c.rd_hd;
  }
  // K14 emulation on one GPU: the stored weight matrix [rows_w, ldb] sits at flat offset flat_off of a flat
  // parameter space that is cut into `owners` equal shards, each shard its own allocation (a "rank").
  std::vector<void*> shards;
  if (c.owners > 1) {
    const size_t rows_w = c.b_mn ? K : N;
    const size_t lead_rows = 64;                                   // the matrix does not start at flat offset 0
    size_t rpo = ((rows_w + lead_rows + c.owners - 1) / c.owners + 63) / 64 * 64;   // rows per owner
    ex.b_ws = c.owners;
    ex.b_flat_off = lead_rows * ldb;
    ex.b_shard_elems = rpo * ldb;
    std::vector<__nv_bfloat16> junk(rpo * ldb);
    for (int o = 0; o < c.owners; ++o) {
      void* sp;
      CK(cudaMalloc(&sp, rpo * ldb * 2));
      for (auto& v : junk) v = __float2bfloat16(77.f);              // anything read from outside the matrix shows
      for (size_t r = 0; r < rpo; ++r) {
        const long long g = (long long)o * rpo + r - lead_rows;     // row of the weight matrix
        if (g >= 0 && g < (long long)rows_w)
          memcpy(&junk[r * ldb], &hB[(size_t)g * ldb], (size_t)ldb * 2);
      }
      CK(cudaMemcpy(sp, junk.data(), rpo * ldb * 2, cudaMemcpyHostToDevice));
      shards.push_back(sp);
    }
    ex.b_shards = shards.data();
  }
  int rc = b200_gemm_bf16_ex(dA, c.owners > 1 ? nullptr : dB, dD, M, N, K, lda, ldb, N, c.a_mn, c.b_mn, c.epi, dBias,
                             dAux, dD2, 0.f, 0, c.bn, 0, nullptr, &ex);
  if (rc) {
    printf("  M=%d N=%d K=%d bn=%d units=%d owners=%d: launch rc=%d FAIL\n", M, N, K, c.bn, c.units, c.owners, rc);
    return 1;
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("  kernel error: %s\n", cudaGetErrorString(e));
    exit(3);  // sticky error: stop the whole harness
  }
  CK(cudaMemcpy(hD.data(), dD, d_elems * 2, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hD2.data(), dD2, d_elems * 2, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hDb.data(), dDb, M * 2, cudaMemcpyDeviceToHost));
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
  if (c.dbias) {   // bias gradient: row sums of A over K
    for (int m = 0; m < M; m += (M > 512 ? 7 : 1)) {
      double rs = 0;
      for (int k = 0; k < K; ++k) rs += c.a_mn ? bf(hA[(size_t)k * M + m]) : bf(hA[(size_t)m * K + k]);
      if (c.dbias == 2) rs += bf(hDb0[m]);
      const double got = bf(hDb[m]);
      if (fabs(got - rs) > 0.05 + 0.01 * fabs(rs)) {
        if (bad < 5) printf("    mismatch dbias[%d]: got %f ref %f\n", m, got, rs);
        ++bad;
      }
    }
  }
  if (c.rd_hd) {   // rowdot[(b*H + h)*T + t] = sum over the head's columns of D * aux
    std::vector<float> hRd((size_t)2 * heads * rd_T);
    CK(cudaMemcpy(hRd.data(), dRd, hRd.size() * 4, cudaMemcpyDeviceToHost));
    for (int m = 0; m < M; m += 5) {
      for (int h = 0; h < heads; ++h) {
        double dot = 0;
        for (int cc = 0; cc < c.rd_hd; ++cc) {
          const int n = h * c.rd_hd + cc;
          double acc = 0;
          for (int k = 0; k < K; ++k) {
            float a = c.a_mn ? bf(hA[(size_t)k * M + m]) : bf(hA[(size_t)m * K + k]);
            float b = c.b_mn ? bf(hB[(size_t)k * N + n]) : bf(hB[(size_t)n * K + k]);
            acc += (double)a * b;
          }
          dot += acc * bf(hAux[(size_t)m * N + n]);
        }
        const int b = m / rd_T, t = m % rd_T;
        const double got = hRd[((size_t)b * heads + h) * rd_T + t];
        if (fabs(got - dot) > 0.05 + 0.01 * fabs(dot)) {
          if (bad < 5) printf("    mismatch rowdot[m=%d,h=%d]: got %f ref %f\n", m, h, got, dot);
          ++bad;
        }
      }
    }
  }
  printf("  M=%d N=%d K=%d a_mn=%d b_mn=%d epi=%d bn=%d units=%d owners=%d dbias=%d rd=%d : max_err=%.4f (max_ref=%.2f) bad=%d %s\n",
         M, N, K, c.a_mn, c.b_mn, c.epi, c.bn, c.units, c.owners, c.dbias, c.rd_hd, max_err, max_ref, bad,
         bad ? "FAIL" : "ok");
  cudaFree(dA); cudaFree(dB); cudaFree(dBias); cudaFree(dAux); cudaFree(dD); cudaFree(dD2); cudaFree(dDb);
  if (dRd) cudaFree(dRd);
  for (void* sp : shards) cudaFree(sp);
  return bad ? 1 : 0;
}

static void time_case(int M, int N, int K, int a_mn, int b_mn, int epi, int bn, const char* name, int units = 0,
                      int dbias = 0) {
  B200GemmEx ex;
  memset(&ex, 0, sizeof(ex));
  ex.mc_units = units;
  __nv_bfloat16* dbias_buf = nullptr;
  if (dbias) {
    CK(cudaMalloc(&dbias_buf, (size_t)M * 2));
    ex.dbias = dbias_buf;
  }
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
    b200_gemm_bf16_ex(A[i % nsets], B[i % nsets], D[i % nsets], M, N, K, lda, ldb, N, a_mn, b_mn, epi,
                      bias, X[i % nsets], D2[i % nsets], 0.f, 0, bn, 0, nullptr, &ex);
  CK(cudaDeviceSynchronize());
  const int iters = 40;
  CK(cudaEventRecord(e0));
  for (int i = 0; i < iters; ++i)
    b200_gemm_bf16_ex(A[i % nsets], B[i % nsets], D[i % nsets], M, N, K, lda, ldb, N, a_mn, b_mn, epi,
                      bias, X[i % nsets], D2[i % nsets], 0.f, 0, bn, 0, nullptr, &ex);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double us = ms * 1000.0 / iters;
  const double tflops = 2.0 * M * N * K / (us * 1e-6) / 1e12;
  printf("  time %-28s M=%d N=%d K=%d bn=%d units=%d dbias=%d: %.2f us  %.1f TFLOP/s (%d operand sets, >L2)\n", name,
         M, N, K, bn, units, dbias, us, tflops, nsets);
  if (dbias_buf) cudaFree(dbias_buf);
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
      // ---- B shared inside a thread-block cluster by TMA multicast (units = tile units per cluster) ----
      {512, 256, 256, 0, 0, B200_EPI_BIAS, 128, 2},
      {512, 384, 192, 0, 1, B200_EPI_NONE, 128, 4},
      {1024, 512, 256, 0, 0, B200_EPI_BIAS, 256, 8},
      {1024, 512, 256, 0, 0, B200_EPI_BIAS, 2128, 2},
      {1024, 512, 192, 0, 1, B200_EPI_NONE, 2256, 4},
      {2048, 512, 256, 0, 0, B200_EPI_BIAS_GELU, 2256, 8},
      {2048, 1024, 512, 0, 1, B200_EPI_DGELU, 2256, 8},
      {1000, 520, 200, 0, 0, B200_EPI_BIAS_RESID, 2256, 4},
      // ---- K14: B fetched per owner from separate allocations (emulated ranks) ----
      {256, 384, 256, 0, 0, B200_EPI_BIAS, 2128, 1, 2},
      {256, 384, 256, 0, 1, B200_EPI_NONE, 128, 1, 4},
      {512, 512, 512, 0, 0, B200_EPI_BIAS, 2256, 1, 4},
      {2048, 3072, 1024, 0, 0, B200_EPI_BIAS, 0, 0, 8},
      {2048, 1024, 3072, 0, 1, B200_EPI_NONE, 0, 0, 8},
      {2048, 1024, 4096, 0, 0, B200_EPI_BIAS_RESID, 0, 0, 4},
      {2048, 1024, 4096, 0, 1, B200_EPI_DGELU, 0, 0, 2},
      // ---- weight gradient + bias gradient (row sums of A) in one main loop ----
      {256, 256, 192, 1, 1, B200_EPI_NONE, 128, 0, 0, 1},
      {384, 256, 320, 1, 1, B200_EPI_ACCUM, 256, 0, 0, 2},
      {512, 512, 256, 1, 1, B200_EPI_NONE, 2256, 0, 0, 1},
      {3072, 1024, 2048, 1, 1, B200_EPI_NONE, 0, 0, 0, 1},
      {1024, 4096, 2048, 1, 1, B200_EPI_ACCUM, 0, 0, 0, 2},
      // ---- D = acc plus per-head dot products with aux (attention-backward D) ----
      {256, 256, 128, 0, 1, B200_EPI_ROWDOT, 128, 0, 0, 0, 64},
      {512, 512, 256, 0, 1, B200_EPI_ROWDOT, 2256, 0, 0, 0, 128},
      {2048, 1024, 1024, 0, 1, B200_EPI_ROWDOT, 0, 0, 0, 0, 64},
  };
  const bool small = argc > 1 && argv[1][0] == 's';
  const bool time_only = argc > 1 && argv[1][0] == 't';     // "timeonly"
  for (const Case& c : cases) {
    if (time_only) break;
    if (small && (long long)c.M * c.N * c.K > (1ll << 27)) continue;
    fails += run_case(c);
  }
  printf("== gemm numerics: %d failing case(s) ==\n", fails);
  if (argc > 1 && !small) {
    printf("== gemm timing ==\n");
    clocks::Sampler sampler;
    sampler.start();
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
    sampler.stop_and_print("gemm timing, plain shapes");
    sampler.start();
    printf("== multicast / bias-gradient variants ==\n");
    for (int u : {1, 2, 4, 8}) time_case(2048, 3072, 1024, 0, 0, B200_EPI_BIAS, 2256, "qkv fwd pair mc", u);
    for (int u : {1, 2, 4, 8}) time_case(2048, 1024, 4096, 0, 1, B200_EPI_NONE, 2256, "fc1 dgrad pair mc", u);
    for (int u : {1, 4}) time_case(2048, 4096, 1024, 0, 0, B200_EPI_BIAS_GELU, 256, "fc1 fwd 1cta mc", u);
    time_case(3072, 1024, 2048, 1, 1, B200_EPI_NONE, 0, "qkv wgrad", 0, 0);
    time_case(3072, 1024, 2048, 1, 1, B200_EPI_NONE, 0, "qkv wgrad + dbias", 0, 1);
    time_case(4096, 1024, 2048, 1, 1, B200_EPI_NONE, 0, "fc1 wgrad", 0, 0);
    time_case(4096, 1024, 2048, 1, 1, B200_EPI_NONE, 0, "fc1 wgrad + dbias", 0, 1);
    time_case(1024, 1024, 2048, 1, 1, B200_EPI_NONE, 0, "proj wgrad + dbias", 0, 1);
    sampler.stop_and_print("gemm timing section");
  }
  return fails ? 1 : 0;
}
