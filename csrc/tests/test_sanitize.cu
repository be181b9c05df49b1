// One binary that drives every single-GPU kernel of the library at small shapes, meant to be run
// under compute-sanitizer (the reference has no sanitizer targets at all):
//
//   compute-sanitizer --tool memcheck  build/bin/test_sanitize
//   compute-sanitizer --tool synccheck build/bin/test_sanitize
//
// The GEMM and attention parts are the numerics harnesses themselves (their `small` mode), pulled in
// under namespaces; the bandwidth kernels are called once each with poisoned guard regions around
// every output so that an out-of-bounds store that memcheck's allocation granularity would hide is
// still caught.  Exit code 0 = numerics pass and all guards intact.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../include/b200bench.h"
#include "clock_sampler.h"   // before the namespaces below: the harness sources include it too

namespace gemm_t {
#define main harness_main
#include "test_gemm.cu"
#undef main
}  // namespace gemm_t
namespace attn_t {
#define main harness_main
#include "test_attn.cu"
#undef main
}  // namespace attn_t

namespace {

constexpr size_t kGuard = 256;          // bytes of 0xA5 before and after every tracked buffer
struct Guarded {
  unsigned char* base = nullptr;
  size_t bytes = 0;
  const char* name = "";
  void* ptr() const { return base + kGuard; }
};
std::vector<Guarded> g_bufs;

void* galloc(size_t bytes, const char* name, int fill = 0) {
  Guarded g;
  g.bytes = (bytes + 255) / 256 * 256;
  g.name = name;
  CK(cudaMalloc(&g.base, g.bytes + 2 * kGuard));
  CK(cudaMemset(g.base, 0xA5, g.bytes + 2 * kGuard));
  CK(cudaMemset(g.base + kGuard, fill, bytes));
  g_bufs.push_back(g);
  return g.ptr();
}

int check_guards() {
  int bad = 0;
  std::vector<unsigned char> h(kGuard);
  for (const Guarded& g : g_bufs) {
    for (int side = 0; side < 2; ++side) {
      CK(cudaMemcpy(h.data(), g.base + (side ? kGuard + g.bytes : 0), kGuard, cudaMemcpyDeviceToHost));
      for (size_t i = 0; i < kGuard; ++i)
        if (h[i] != 0xA5) {
          printf("  GUARD HIT: %s %s byte %zu\n", g.name, side ? "after" : "before", i);
          ++bad;
          break;
        }
    }
  }
  return bad;
}

void fill_bf16(void* d, size_t n, float scale, uint32_t seed) {
  std::vector<__nv_bfloat16> h(n);
  uint32_t s = seed;
  for (size_t i = 0; i < n; ++i) {
    s = s * 1664525u + 1013904223u;
    h[i] = __float2bfloat16(((s >> 8) * (1.0f / 16777216.0f) - 0.5f) * 2.f * scale);
  }
  CK(cudaMemcpy(d, h.data(), n * 2, cudaMemcpyHostToDevice));
}

#define RC(call)                                             \
  do {                                                       \
    int rc_ = (call);                                        \
    if (rc_ != 0) {                                          \
      printf("  %s -> error %d\n", #call, rc_);              \
      ++fails;                                               \
    }                                                        \
  } while (0)

// Odd sizes on purpose: M not a multiple of the row tile, d the smallest / an unusual multiple of 8.
int run_elementwise(int M, int d, int V, int B, int T) {
  int fails = 0;
  printf("  elementwise M=%d d=%d V=%d (B=%d T=%d)\n", M, d, V, B, T);
  const size_t md = (size_t)M * d;
  void* x = galloc(md * 2, "x");           fill_bf16(x, md, 1.f, 1);
  void* dy = galloc(md * 2, "dy");         fill_bf16(dy, md, 1.f, 2);
  void* y = galloc(md * 2, "y");
  void* dx = galloc(md * 2, "dx");
  void* gamma = galloc(d * 2, "gamma");    fill_bf16(gamma, d, 1.f, 3);
  void* beta = galloc(d * 2, "beta");      fill_bf16(beta, d, 1.f, 4);
  float* mean = (float*)galloc(M * 4, "mean");
  float* rstd = (float*)galloc(M * 4, "rstd");
  void* dgamma = galloc(d * 2, "dgamma");
  void* dbeta = galloc(d * 2, "dbeta");
  void* dbias = galloc(d * 2, "dbias");
  const size_t sb = b200_colreduce_scratch_bytes(M, d, 2);
  void* scratch = galloc(sb, "colreduce scratch");
  RC(b200_layernorm_fwd(x, gamma, beta, y, mean, rstd, M, d, 1e-5f, nullptr));
  RC(b200_layernorm_bwd(dy, x, gamma, mean, rstd, x, dx, M, d, nullptr));
  RC(b200_layernorm_bwd_params(dy, x, mean, rstd, dgamma, dbeta, M, d, 0, scratch, sb, nullptr));
  RC(b200_layernorm_bwd_params(dy, x, mean, rstd, dgamma, dbeta, M, d, 1, scratch, sb, nullptr));
  RC(b200_colsum(dy, dbias, M, d, d, 0, scratch, sb, nullptr));
  RC(b200_colsum_static(dy, dbias, M, d, d, 1, nullptr));
  RC(b200_dropout_apply(dy, dx, M, d, 0.1f, 99u, nullptr));

  // embedding forward / backward
  std::vector<int> htok((size_t)B * T);
  for (size_t i = 0; i < htok.size(); ++i) htok[i] = (int)((i * 2654435761u) % (uint32_t)V);
  int* tok = (int*)galloc(htok.size() * 4, "tokens");
  CK(cudaMemcpy(tok, htok.data(), htok.size() * 4, cudaMemcpyHostToDevice));
  void* wte = galloc((size_t)V * d * 2, "wte");   fill_bf16(wte, (size_t)V * d, 0.02f, 5);
  void* wpe = galloc((size_t)T * d * 2, "wpe");   fill_bf16(wpe, (size_t)T * d, 0.02f, 6);
  void* emb = galloc((size_t)B * T * d * 2, "emb");
  void* dwte = galloc((size_t)V * d * 2, "dwte");
  void* dwpe = galloc((size_t)T * d * 2, "dwpe");
  RC(b200_embedding_fwd(tok, wte, wpe, emb, B, T, d, 0.1f, 7u, nullptr));
  RC(b200_embedding_bwd(tok, emb, dwte, dwpe, B, T, d, 0.1f, 7u, 0, nullptr));

  // cross-entropy, with some ignored rows
  const int Mc = B * T;
  void* logits = galloc((size_t)Mc * V * 2, "logits");   fill_bf16(logits, (size_t)Mc * V, 4.f, 8);
  std::vector<int> htgt(Mc);
  for (int i = 0; i < Mc; ++i) htgt[i] = (i % 7 == 3) ? -1 : (int)((i * 40503u) % (uint32_t)V);
  int* tgt = (int*)galloc(Mc * 4, "targets");
  CK(cudaMemcpy(tgt, htgt.data(), Mc * 4, cudaMemcpyHostToDevice));
  float* loss_rows = (float*)galloc(Mc * 4, "loss_rows");
  int* n_valid = (int*)galloc(4, "n_valid");
  float* loss = (float*)galloc(4, "loss");
  RC(b200_cross_entropy(logits, tgt, loss_rows, n_valid, loss, Mc, V, V, -1, 1.f, 1, 0, 1.f, nullptr));

  // AdamW + squared norm over an n that is a multiple of 8 but of nothing larger
  const size_t n = 8 * 1237;
  float* master = (float*)galloc(n * 4, "master");
  float* m1 = (float*)galloc(n * 4, "exp_avg");
  float* m2 = (float*)galloc(n * 4, "exp_avg_sq");
  void* grad = galloc(n * 2, "grad");       fill_bf16(grad, n, 1.f, 9);
  void* wout = galloc(n * 2, "w_out");
  float* hp = (float*)galloc(9 * 4, "hp");
  float* gn = (float*)galloc(4, "gnorm_sq");
  RC(b200_adamw_set_hp(hp, 1e-4f, 0.9f, 0.999f, 1e-8f, 0.01f, 10.f, 31.6f, 1.f, 1.f, nullptr));
  RC(b200_sumsq(grad, n, 1.f, gn, nullptr));
  RC(b200_adamw(master, m1, m2, grad, wout, n, hp, gn, nullptr));
  unsigned int* seed_word = (unsigned int*)galloc(4, "step seed");
  RC(b200_step_seed_register(seed_word));
  RC(b200_step_seed_advance(nullptr));
  CK(cudaDeviceSynchronize());
  float hl = 0.f;
  CK(cudaMemcpy(&hl, loss, 4, cudaMemcpyDeviceToHost));
  if (!(hl > 0.f && hl < 100.f)) {
    printf("  cross-entropy loss out of range: %f\n", hl);
    ++fails;
  }
  return fails;
}

}  // namespace

int main() {
  char a0[] = "test_sanitize", a1[] = "small";
  char* argv[] = {a0, a1, nullptr};
  int fails = 0;
  fails += gemm_t::harness_main(2, argv);
  fails += attn_t::harness_main(2, argv);
  printf("== bandwidth kernels ==\n");
  fails += run_elementwise(200, 1024, 1000, 2, 100);
  fails += run_elementwise(77, 256, 520, 1, 77);
  const int guard_hits = check_guards();
  printf("== bandwidth kernels: %d error(s), %d guard hit(s) ==\n", fails, guard_hits);
  CK(cudaDeviceSynchronize());
  printf("== test_sanitize: %s ==\n", (fails || guard_hits) ? "FAIL" : "PASS");
  return (fails || guard_hits) ? 1 : 0;
}
