// Standalone numerics + timing harness for the fused attention kernels (forward and backward).
#include "clock_sampler.h"
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
static uint32_t rng_state = 777;
static float frand() {
  rng_state = rng_state * 1664525u + 1013904223u;
  return ((rng_state >> 8) & 0xFFFF) / 65536.0f - 0.5f;
}

// CPU reference: forward (O, LSE) and backward (dQ, dK, dV) for one (b, h), no dropout.
static void ref_head(const std::vector<__nv_bfloat16>& qkv, const std::vector<__nv_bfloat16>& dO,
                     int T, int d, int b, int h, int causal, std::vector<double>& O,
                     std::vector<double>& LSE, std::vector<double>& dQ, std::vector<double>& dK,
                     std::vector<double>& dV, bool backward) {
  const int hd = 64;
  const double scale = 1.0 / sqrt((double)hd);
  O.assign((size_t)T * hd, 0);
  LSE.assign(T, 0);
  std::vector<double> P((size_t)T * T);
  auto Q = [&](int t, int c) { return (double)bf(qkv[(size_t)(b * T + t) * 3 * d + h * hd + c]); };
  auto K = [&](int t, int c) { return (double)bf(qkv[(size_t)(b * T + t) * 3 * d + d + h * hd + c]); };
  auto V = [&](int t, int c) { return (double)bf(qkv[(size_t)(b * T + t) * 3 * d + 2 * d + h * hd + c]); };
  for (int i = 0; i < T; ++i) {
    double mx = -1e300;
    for (int j = 0; j < T; ++j) {
      if (causal && j > i) { P[(size_t)i * T + j] = -1e300; continue; }
      double s = 0;
      for (int c = 0; c < hd; ++c) s += Q(i, c) * K(j, c);
      s *= scale;
      P[(size_t)i * T + j] = s;
      if (s > mx) mx = s;
    }
    double sum = 0;
    for (int j = 0; j < T; ++j) {
      double e = (causal && j > i) ? 0.0 : exp(P[(size_t)i * T + j] - mx);
      P[(size_t)i * T + j] = e;
      sum += e;
    }
    LSE[i] = mx + log(sum);
    for (int j = 0; j < T; ++j) {
      P[(size_t)i * T + j] /= sum;
      const double pj = P[(size_t)i * T + j];
      if (pj != 0)
        for (int c = 0; c < hd; ++c) O[(size_t)i * hd + c] += pj * V(j, c);
    }
  }
  if (!backward) return;
  dQ.assign((size_t)T * hd, 0);
  dK.assign((size_t)T * hd, 0);
  dV.assign((size_t)T * hd, 0);
  auto DO = [&](int t, int c) { return (double)bf(dO[(size_t)(b * T + t) * d + h * hd + c]); };
  for (int i = 0; i < T; ++i) {
    double D = 0;
    for (int c = 0; c < hd; ++c) D += DO(i, c) * O[(size_t)i * hd + c];
    for (int j = 0; j < T; ++j) {
      const double pj = P[(size_t)i * T + j];
      if (pj == 0) continue;
      double dp = 0;
      for (int c = 0; c < hd; ++c) dp += DO(i, c) * V(j, c);
      const double ds = pj * (dp - D) * scale;
      for (int c = 0; c < hd; ++c) {
        dV[(size_t)j * hd + c] += pj * DO(i, c);
        dQ[(size_t)i * hd + c] += ds * K(j, c);
        dK[(size_t)j * hd + c] += ds * Q(i, c);
      }
    }
  }
}

static int run_case(int B, int T, int H, int causal, bool backward) {
  const int hd = 64, d = H * hd;
  const size_t n_qkv = (size_t)B * T * 3 * d, n_o = (size_t)B * T * d;
  std::vector<__nv_bfloat16> hqkv(n_qkv), hO(n_o), hdO(n_o), hdqkv(n_qkv);
  for (auto& v : hqkv) v = __float2bfloat16(frand() * 4.f);
  for (auto& v : hdO) v = __float2bfloat16(frand());
  std::vector<float> hlse((size_t)B * H * T);
  __nv_bfloat16 *dqkv, *dO_, *ddO, *ddqkv;
  float *dlse, *dD, *dqacc;
  CK(cudaMalloc(&dqkv, n_qkv * 2));
  CK(cudaMalloc(&dO_, n_o * 2));
  CK(cudaMalloc(&ddO, n_o * 2));
  CK(cudaMalloc(&ddqkv, n_qkv * 2));
  CK(cudaMalloc(&dlse, hlse.size() * 4));
  CK(cudaMalloc(&dD, hlse.size() * 4));
  CK(cudaMalloc(&dqacc, n_o * 4));
  CK(cudaMemcpy(dqkv, hqkv.data(), n_qkv * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(ddO, hdO.data(), n_o * 2, cudaMemcpyHostToDevice));
  CK(cudaMemset(dO_, 0, n_o * 2));
  CK(cudaMemset(ddqkv, 0, n_qkv * 2));
  int rc = b200_attention_fwd(dqkv, dO_, dlse, B, T, H, hd, causal, 0.f, 0, nullptr);
  if (rc) { printf("  fwd launch rc=%d\n", rc); return 1; }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("  fwd kernel error: %s\n", cudaGetErrorString(e)); exit(3); }
  if (backward) {
    rc = b200_attention_bwd(dqkv, dO_, ddO, dlse, ddqkv, dD, dqacc, B, T, H, hd, causal, 0.f, 0,
                            nullptr);
    if (rc) { printf("  bwd launch rc=%d\n", rc); return 1; }
    e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("  bwd kernel error: %s\n", cudaGetErrorString(e)); exit(3); }
  }
  CK(cudaMemcpy(hO.data(), dO_, n_o * 2, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hlse.data(), dlse, hlse.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hdqkv.data(), ddqkv, n_qkv * 2, cudaMemcpyDeviceToHost));
  double eo = 0, el = 0, eq = 0, ek = 0, ev = 0, mq = 0, mk = 0, mv = 0;
  std::vector<double> O, LSE, dQ, dK, dV;
  for (int b = 0; b < B; ++b)
    for (int h = 0; h < H; ++h) {
      ref_head(hqkv, hdO, T, d, b, h, causal, O, LSE, dQ, dK, dV, backward);
      for (int t = 0; t < T; ++t) {
        el = fmax(el, fabs(LSE[t] - hlse[(size_t)(b * H + h) * T + t]));
        for (int c = 0; c < hd; ++c) {
          eo = fmax(eo, fabs(O[(size_t)t * hd + c] - bf(hO[(size_t)(b * T + t) * d + h * hd + c])));
          if (backward) {
            const size_t base = (size_t)(b * T + t) * 3 * d + h * hd + c;
            eq = fmax(eq, fabs(dQ[(size_t)t * hd + c] - bf(hdqkv[base])));
            ek = fmax(ek, fabs(dK[(size_t)t * hd + c] - bf(hdqkv[base + d])));
            ev = fmax(ev, fabs(dV[(size_t)t * hd + c] - bf(hdqkv[base + 2 * d])));
            mq = fmax(mq, fabs(dQ[(size_t)t * hd + c]));
            mk = fmax(mk, fabs(dK[(size_t)t * hd + c]));
            mv = fmax(mv, fabs(dV[(size_t)t * hd + c]));
          }
        }
      }
    }
  bool bad = eo > 0.03 || el > 0.02;
  if (backward) bad = bad || eq > 0.02 + 0.02 * mq || ek > 0.02 + 0.02 * mk || ev > 0.02 + 0.02 * mv;
  printf("  B=%d T=%d H=%d causal=%d: err O=%.4f LSE=%.4f", B, T, H, causal, eo, el);
  if (backward) printf(" dQ=%.4f/%.2f dK=%.4f/%.2f dV=%.4f/%.2f", eq, mq, ek, mk, ev, mv);
  printf(" %s\n", bad ? "FAIL" : "ok");
  cudaFree(dqkv); cudaFree(dO_); cudaFree(ddO); cudaFree(ddqkv); cudaFree(dlse); cudaFree(dD); cudaFree(dqacc);
  return bad ? 1 : 0;
}

static void time_case(int B, int T, int H, int causal, bool backward, float drop) {
  const int hd = 64, d = H * hd;
  const size_t n_qkv = (size_t)B * T * 3 * d, n_o = (size_t)B * T * d;
  const int nsets = 12;  // > L2 in aggregate for tier-A shapes
  std::vector<__nv_bfloat16*> qkv(nsets), O(nsets), dO(nsets), dqkv(nsets);
  std::vector<__nv_bfloat16> h(n_qkv);
  for (auto& v : h) v = __float2bfloat16(frand() * 2.f);
  float *lse, *D, *dqacc;
  CK(cudaMalloc(&lse, (size_t)B * H * T * 4));
  CK(cudaMalloc(&D, (size_t)B * H * T * 4));
  CK(cudaMalloc(&dqacc, n_o * 4));
  for (int i = 0; i < nsets; ++i) {
    CK(cudaMalloc(&qkv[i], n_qkv * 2));
    CK(cudaMalloc(&O[i], n_o * 2));
    CK(cudaMalloc(&dO[i], n_o * 2));
    CK(cudaMalloc(&dqkv[i], n_qkv * 2));
    CK(cudaMemcpy(qkv[i], h.data(), n_qkv * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dO[i], h.data(), n_o * 2, cudaMemcpyHostToDevice));
  }
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  auto run = [&](int i) {
    if (!backward)
      b200_attention_fwd(qkv[i % nsets], O[i % nsets], lse, B, T, H, hd, causal, drop, 1, nullptr);
    else
      b200_attention_bwd(qkv[i % nsets], O[i % nsets], dO[i % nsets], lse, dqkv[i % nsets], D, dqacc,
                         B, T, H, hd, causal, drop, 1, nullptr);
  };
  if (backward)
    for (int i = 0; i < nsets; ++i)
      b200_attention_fwd(qkv[i], O[i], lse, B, T, H, hd, causal, drop, 1, nullptr);
  for (int i = 0; i < 3; ++i) run(i);
  CK(cudaDeviceSynchronize());
  const int iters = 30;
  CK(cudaEventRecord(e0));
  for (int i = 0; i < iters; ++i) run(i);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double us = ms * 1000.0 / iters;
  double flops = 4.0 * B * H * (double)T * T * hd * (backward ? 2.5 : 1.0) * (causal ? 0.5 : 1.0);
  printf("  time attn %s B=%d T=%d H=%d causal=%d drop=%.1f: %.1f us  %.1f TFLOP/s\n",
         backward ? "bwd" : "fwd", B, T, H, causal, drop, us, flops / (us * 1e-6) / 1e12);
  for (int i = 0; i < nsets; ++i) { cudaFree(qkv[i]); cudaFree(O[i]); cudaFree(dO[i]); cudaFree(dqkv[i]); }
  cudaFree(lse); cudaFree(D); cudaFree(dqacc);
}

int main(int argc, char** argv) {
  const bool small = (argc > 1 && argv[1][0] == 's');   // numerics only (fwd+bwd), for compute-sanitizer
  const bool bwd = small || (argc > 1 && argv[1][0] == 'b');
  int fails = 0;
  printf("== attention numerics (%s) ==\n", bwd ? "fwd+bwd" : "fwd");
  fails += run_case(1, 128, 1, 0, bwd);
  fails += run_case(1, 256, 2, 0, bwd);
  fails += run_case(2, 384, 2, 0, bwd);
  fails += run_case(1, 512, 1, 1, bwd);
  fails += run_case(1, 200, 1, 0, bwd);
  fails += run_case(2, 328, 2, 1, bwd);
  printf("== attention numerics: %d failing case(s) ==\n", fails);
  if (small) return fails ? 1 : 0;
  printf("== attention timing ==\n");
  clocks::Sampler sampler;
  sampler.start();
  for (int variant : {0, 1}) {     // forward variants: 0 = exact per-block rescale, 1 = lazy rescale with O in TMEM
    b200_set_attention_fwd_variant(variant);
    printf("  forward variant %d (%s)\n", variant, variant ? "lazy, default" : "exact");
    time_case(1, 2048, 16, 0, false, 0.f);
    time_case(1, 2048, 16, 0, false, 0.1f);
    time_case(1, 2048, 16, 1, false, 0.f);
  }
  time_case(1, 2048, 16, 0, false, 0.f);
  time_case(1, 2048, 16, 0, false, 0.1f);
  time_case(1, 2048, 16, 1, false, 0.f);
  time_case(8, 2048, 16, 0, false, 0.f);
  time_case(1, 8192, 16, 0, false, 0.f);
  if (bwd) {
    time_case(1, 2048, 16, 0, true, 0.f);
    time_case(1, 2048, 16, 0, true, 0.1f);
    time_case(8, 2048, 16, 0, true, 0.f);
  }
  sampler.stop_and_print("attention timing section");
  return fails ? 1 : 0;
}
