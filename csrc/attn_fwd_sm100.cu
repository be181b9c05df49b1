// Fused attention forward for sm_100a: S = Q K^T and O = P V on tcgen05 tensor cores with the
// accumulators in tensor memory, online softmax in registers, K/V streamed by TMA.  The T x T score
// matrix is never written to memory.
//
// Replaces the reference's unfused math path (benchmarking/train_harness.py:127 ->
// nn.MultiheadAttention with need_weights=True: bmm, fp32 softmax, dropout, bmm and a discarded
// head-mean).  Non-causal by default because the reference passes no mask
// (SURVEY Q5); `causal` switches the mask on.  Dropout is applied to P after the softmax with a
// counter-based RNG keyed on (seed, b, h, q, k) so that backward regenerates the same mask.
//
// Input  qkv [B*T, 3*d]  bf16, row m = b*T + t, columns [q | k | v], head h at h*64 inside each.
// Output out [B*T, d]    bf16, head h at columns h*64 .. h*64+63;   lse [B*H, T] fp32 (natural log).
//
// CTA = one (batch, head, 128-row query block), 384 threads (3 warpgroups, registers re-balanced
// with setmaxnreg: 40 for the control warpgroup, 232 for the softmax warpgroups):
//   warp 0      TMA producer: Q once, then K/V tiles of 128 keys through 3-stage rings
//   warp 1      MMA issuer:   S[w] = Q K_j^T  (M128 N128 K64),  O_blk[w] = P[w] V_j  (M128 N64 K128)
//   warps 4-7   softmax warpgroup 0 : even key blocks      (thread = one query row)
//   warps 8-11  softmax warpgroup 1 : odd key blocks
// Each warpgroup owns one S buffer (128 TMEM columns), one O_blk buffer (64 columns) and one P
// buffer in shared memory (bf16, 128B-swizzled K-major so it is directly the A operand of P V), and
// keeps its own running (max, sum, O) in registers; the two partial results are merged at the end.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <type_traits>

#include "b200_internal.h"
#include "sm100_ptx.cuh"

namespace b200 {

constexpr int AT_BM = 128;       // query rows per CTA
constexpr int AT_BN = 128;       // keys per block
constexpr int AT_HD = 64;        // head dim
constexpr int AT_KV_STAGES = 3;
constexpr int AT_THREADS = 384;  // warpgroup 0: TMA + MMA (+2 idle warps), warpgroups 1-2: softmax
constexpr int AT_TILE = AT_BN * AT_HD * 2;   // 16 KiB: one [128 x 64] bf16 tile
constexpr int AT_PBYTES = AT_BM * AT_BN * 2; // 32 KiB: one P buffer
constexpr int AT_SMEM = AT_TILE * (1 + 2 * AT_KV_STAGES) + 2 * AT_PBYTES + 1280 /*bars+merge*/ +
                        2048 /*dropout column keys: 2 warpgroups x 2 buffers x 128*/;
constexpr float NEG_BIG = -1.0e30f;

struct AttnFwdParams {
  __nv_bfloat16* out;
  float* lse;
  int B, T, H, d;
  float scale_log2;       // softmax_scale * log2(e)
  int causal;
  uint32_t drop_threshold;
  float drop_scale;
  uint32_t drop_seed;
  const uint32_t* seed_ptr;
};

// LAZY = false: every softmax thread keeps its O row in registers and folds each block's P V product in with the
//   exact running-max rescale (the round-1 kernel).
// LAZY = true: O accumulates in tensor memory across the blocks of a warpgroup (the P V MMAs accumulate), the S row
//   stays in registers between the max and the exponentials (one TMEM read instead of two), and the reference
//   maximum m_ref only moves - with a rescale of the O row in TMEM - when a block's maximum exceeds it by more than 8
//   (log2 domain): probabilities are then at most 2^8 before normalisation, exact in fp32 and harmless in bf16, and
//   after the first blocks almost no block pays for a rescale.
template <bool LAZY>
__global__ void __launch_bounds__(AT_THREADS, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap map_qkv, const AttnFwdParams p) {
  // SWIZZLE_128B tiles need 1024-byte alignment; declaring the alignment (instead of rounding the
  // pointer up through an integer) keeps the address space known, so accesses are LDS/STS.
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + AT_TILE;
  uint8_t* sV = sK + AT_KV_STAGES * AT_TILE;
  uint8_t* sP = sV + AT_KV_STAGES * AT_TILE;  // 2 x 32 KiB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * AT_PBYTES);
  uint64_t* q_full = bars;                         // 1
  uint64_t* k_full = bars + 1;                     // [3]
  uint64_t* k_empty = bars + 4;                    // [3]
  uint64_t* v_full = bars + 7;                     // [3]
  uint64_t* v_empty = bars + 10;                   // [3]
  uint64_t* s_full = bars + 13;                    // [2]
  uint64_t* p_full = bars + 15;                    // [2]
  uint64_t* o_full = bars + 17;                    // [2]
  uint64_t* s_free = bars + 19;                    // [2]  LAZY: the S row is in registers, the buffer can be refilled
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 21);
  float* merge_ml = reinterpret_cast<float*>(bars + 22);  // [2][128] m, l of warpgroup 1
  uint32_t* sColKey = reinterpret_cast<uint32_t*>(merge_ml + 256);   // [2 warpgroups][2 buffers][128]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qblocks = (p.T + AT_BM - 1) / AT_BM;
  const int qb = blockIdx.x % qblocks;
  const int bh = blockIdx.x / qblocks;
  const int b = bh / p.H, h = bh % p.H;
  const int q0 = qb * AT_BM;
  const int row_base = b * p.T;  // first row of this batch element in qkv
  int nkv = (p.T + AT_BN - 1) / AT_BN;
  if (p.causal) nkv = min(nkv, qb + 1);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_qkv);
    mbar_init(q_full, 1);
    for (int s = 0; s < AT_KV_STAGES; ++s) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], 1);
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], 1);
    }
    for (int w = 0; w < 2; ++w) {
      mbar_init(&s_full[w], 1);
      mbar_init(&p_full[w], 128);
      mbar_init(&o_full[w], 1);
      mbar_init(&s_free[w], 128);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();   // dependents may start their prologue; they wait for this grid's completion themselves
  pdl_wait();      // nothing below may run before the previous kernel's writes are visible
  // TMEM columns: S0 [0,128) S1 [128,256) O0 [256,320) O1 [320,384)

  // LAZY with dropout: release the S buffer as soon as the row is in registers (measured: 58.6 -> 55.4 us with
  // dropout 0.1 at the headline shape, but 45.3 -> 47.6 us without dropout, where the softmax is too short to gain)
  const bool early_s = LAZY && p.drop_threshold != 0;
  if (warp < 4) {
    setmaxnreg_dec<40>();
  if (warp == 0) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      const int qcol = h * AT_HD, kcol = p.d + h * AT_HD, vcol = 2 * p.d + h * AT_HD;
      mbar_expect_tx(q_full, AT_TILE);
      tma_load_2d(&map_qkv, q_full, sQ, qcol, row_base + q0);
      int ks = 0, vs = 0;
      uint32_t kph = 0, vph = 0;
      auto load_k = [&](int j) {
        mbar_wait(&k_empty[ks], kph ^ 1);
        mbar_expect_tx(&k_full[ks], AT_TILE);
        tma_load_2d(&map_qkv, &k_full[ks], sK + ks * AT_TILE, kcol, row_base + j * AT_BN);
        if (++ks == AT_KV_STAGES) { ks = 0; kph ^= 1; }
      };
      auto load_v = [&](int j) {
        mbar_wait(&v_empty[vs], vph ^ 1);
        mbar_expect_tx(&v_full[vs], AT_TILE);
        tma_load_2d(&map_qkv, &v_full[vs], sV + vs * AT_TILE, vcol, row_base + j * AT_BN);
        if (++vs == AT_KV_STAGES) { vs = 0; vph ^= 1; }
      };
      load_k(0);
      if (nkv > 1) load_k(1);
      for (int j = 0; j < nkv; ++j) {
        load_v(j);
        if (j + 2 < nkv) load_k(j + 2);
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    if (lane == 0) {
      const uint32_t idesc_s = umma_idesc_bf16(AT_BM, AT_BN, false, false);
      const uint32_t idesc_o = umma_idesc_bf16(AT_BM, AT_HD, false, true);
      const uint32_t aQ = smem_u32(sQ);
      int ks = 0, vs = 0;
      uint32_t kph = 0, vph = 0;
      auto issue_s = [&](int w) {
        mbar_wait(&k_full[ks], kph);
        tc_fence_after();
        const uint32_t aK = smem_u32(sK + ks * AT_TILE);
#pragma unroll
        for (int k = 0; k < AT_HD / 16; ++k)
          umma_bf16_ss(tmem_base + w * 128, umma_smem_desc(aQ + k * 32, 16, 1024),
                       umma_smem_desc(aK + k * 32, 16, 1024), idesc_s, k ? 1u : 0u);
        umma_commit(&k_empty[ks]);
        umma_commit(&s_full[w]);
        if (++ks == AT_KV_STAGES) { ks = 0; kph ^= 1; }
      };
      mbar_wait(q_full, 0);
      tc_fence_after();
      issue_s(0);
      if (nkv > 1) issue_s(1);
      uint32_t pph[2] = {0, 0}, sfph[2] = {0, 0};
      for (int j = 0; j < nkv; ++j) {
        const int w = j & 1;
        if (early_s) {
          // the warpgroup holds the whole S(j) row in registers: S(j+2) is computed while it is still busy with
          // the exponentials of block j, so it never waits for scores
          if (j + 2 < nkv) {
            mbar_wait(&s_free[w], sfph[w]);
            sfph[w] ^= 1;
            tc_fence_after();
            issue_s(w);
          }
        }
        mbar_wait(&p_full[w], pph[w]);  // P(j) in smem, S[w] fully read
        pph[w] ^= 1;
        tc_fence_after();
        if (!early_s) {
          if (j + 2 < nkv) issue_s(w);  // next scores for this warpgroup first
        }
        mbar_wait(&v_full[vs], vph);
        tc_fence_after();
        const uint32_t aP = smem_u32(sP + w * AT_PBYTES);
        const uint32_t aV = smem_u32(sV + vs * AT_TILE);
#pragma unroll
        for (int k = 0; k < AT_BN / 16; ++k)
          umma_bf16_ss(tmem_base + 256 + w * 64,
                       umma_smem_desc(aP + (k >> 2) * (AT_BM * 128) + (k & 3) * 32, 16, 1024),
                       umma_smem_desc(aV + k * 2048, AT_BN * 128, 1024), idesc_o,
                       (k || (LAZY && j >= 2)) ? 1u : 0u);   // LAZY: later blocks add onto the warpgroup's O
        umma_commit(&v_empty[vs]);
        umma_commit(&o_full[w]);
        if (++vs == AT_KV_STAGES) { vs = 0; vph ^= 1; }
      }
    }
  }
  } else {
    setmaxnreg_inc<232>();
    // ================================ softmax warpgroups ==========================
    const int w = (warp - 4) >> 2;        // softmax warpgroup 0 / 1
    const uint32_t seed = p.drop_seed ^ ((p.drop_threshold && p.seed_ptr) ? __ldg(p.seed_ptr) : 0u);
    const int quarter = warp & 3;         // TMEM lane quarter
    const int r = quarter * 32 + lane;    // query row inside the block
    const int qrow = q0 + r;
    const uint32_t t_lane = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t tS = tmem_base + t_lane + w * 128;
    const uint32_t tO = tmem_base + t_lane + 256 + w * 64;
    uint8_t* myP = sP + w * AT_PBYTES + r * 128;

    float o[AT_HD];
#pragma unroll
    for (int i = 0; i < AT_HD; ++i) o[i] = 0.f;
    float m_run = NEG_BIG, l_run = 0.f, m_ref = NEG_BIG, m_prev = NEG_BIG;
    uint32_t sph = 0, oph = 0;
    int n_done = 0;

    auto accumulate_prev = [&]() {
      // O_blk of this warpgroup's previous block is relative to m_prev; o[] is relative to m_ref.
      mbar_wait(&o_full[w], oph);
      oph ^= 1;
      tc_fence_after();
      const float alpha = exp2f(m_ref - m_prev);
      uint32_t v0[32], v1[32];
      tmem_ld_32x32b_x32(tO, v0);
      tmem_ld_32x32b_x32(tO + 32, v1);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        o[i] = o[i] * alpha + __uint_as_float(v0[i]);
        o[32 + i] = o[32 + i] * alpha + __uint_as_float(v1[i]);
      }
      m_ref = m_prev;
    };

    const uint32_t rowkey = drop_rowkey(seed, static_cast<uint32_t>(bh) * p.T + qrow);
    const bool drop = p.drop_threshold != 0;
    const uint32_t thr32 = p.drop_threshold;   // p * 2^32

    // One key block for this thread's query row.  MASK is a compile-time flag so that the common
    // (interior, non-causal) blocks carry no per-element predicate; limit = first masked key.
    auto block_body = [&](auto mask_c, int k0, int limit) {
      constexpr bool MASK = decltype(mask_c)::value;
      const uint32_t* colkeys = sColKey + (w * 2 + (n_done & 1)) * 128;
      if (drop) {
        // b(k) for the 128 keys of this block, one per thread; double-buffered by block parity so one
        // warpgroup barrier per block is enough
        sColKey[(w * 2 + (n_done & 1)) * 128 + r] = drop_colkey(seed, static_cast<uint32_t>(bh) * p.T + k0 + r);
        asm volatile("bar.sync %0, 128;" ::"r"(2 + w) : "memory");
      }
      // ---- pass 1: row maximum (4 independent chains, two 64-column loads) ----------------
      float mx[4] = {NEG_BIG, NEG_BIG, NEG_BIG, NEG_BIG};
#pragma unroll 1
      for (int hlf = 0; hlf < 2; ++hlf) {
        uint32_t v0[32], v1[32];
        tmem_ld_32x32b_x32(tS + hlf * 64, v0);
        tmem_ld_32x32b_x32(tS + hlf * 64 + 32, v1);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float s0 = __uint_as_float(v0[i]), s1 = __uint_as_float(v1[i]);
          if (MASK) {
            if (k0 + hlf * 64 + i >= limit) s0 = NEG_BIG;
            if (k0 + hlf * 64 + 32 + i >= limit) s1 = NEG_BIG;
          }
          mx[i & 3] = fmaxf(mx[i & 3], fmaxf(s0, s1));
        }
      }
      const float mxx = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      const float m_new = fmaxf(m_run, mxx * p.scale_log2);
      // previous block of this warpgroup: its P buffer must be free before we overwrite it
      if (n_done > 0) accumulate_prev();
      // ---- pass 2: P = exp2(s*scale - m_new), row sum, bf16 -> swizzled smem -----------------
      float rs[4] = {0.f, 0.f, 0.f, 0.f};
      const float neg_m = -m_new;
      uint32_t va[32], vb[32];
      tmem_ld_32x32b_x32(tS, va);
#pragma unroll
      for (int c = 0; c < AT_BN / 32; ++c) {
        uint32_t(&v)[32] = (c & 1) ? vb : va;
        tmem_ld_wait();
        if (c + 1 < AT_BN / 32) tmem_ld_32x32b_x32(tS + (c + 1) * 32, (c & 1) ? va : vb);
        float pv[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float e = exp2f(fmaf(__uint_as_float(v[i]), p.scale_log2, neg_m));
          if (MASK) e = (k0 + c * 32 + i < limit) ? e : 0.f;
          rs[i & 3] += e;
          pv[i] = e;
        }
        if (drop) {
          // keep / drop only: the 1/(1-p) rescale of the kept probabilities is applied once, to the final O
#pragma unroll
          for (int t4 = 0; t4 < 32; t4 += 4) {
            const uint4 b4 = *reinterpret_cast<const uint4*>(colkeys + c * 32 + t4);
            pv[t4 + 0] = drop_keep2d(rowkey, b4.x, thr32) ? pv[t4 + 0] : 0.f;
            pv[t4 + 1] = drop_keep2d(rowkey, b4.y, thr32) ? pv[t4 + 1] : 0.f;
            pv[t4 + 2] = drop_keep2d(rowkey, b4.z, thr32) ? pv[t4 + 2] : 0.f;
            pv[t4 + 3] = drop_keep2d(rowkey, b4.w, thr32) ? pv[t4 + 3] : 0.f;
          }
        }
        // 32 keys = 4 x 16-byte chunks of the 128-byte swizzled row; chunk index XOR (row & 7)
        uint8_t* dst = myP + (c >> 1) * (AT_BM * 128);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 u;
          u.x = pack_bf16x2(pv[g * 8 + 0], pv[g * 8 + 1]);
          u.y = pack_bf16x2(pv[g * 8 + 2], pv[g * 8 + 3]);
          u.z = pack_bf16x2(pv[g * 8 + 4], pv[g * 8 + 5]);
          u.w = pack_bf16x2(pv[g * 8 + 6], pv[g * 8 + 7]);
          const int chunk = ((c & 1) * 4 + g) ^ (r & 7);
          *reinterpret_cast<uint4*>(dst + chunk * 16) = u;
        }
      }
      l_run = l_run * exp2f(m_run - m_new) + ((rs[0] + rs[1]) + (rs[2] + rs[3]));
      m_run = m_new;
      m_prev = m_new;
    };

    // LAZY variant of one key block (see the kernel comment).  Uses m_ref as the only maximum.
    auto block_body_lazy = [&](auto mask_c, int k0, int limit) {
      constexpr bool MASK = decltype(mask_c)::value;
      const uint32_t* colkeys = sColKey + (w * 2 + (n_done & 1)) * 128;
      if (drop) {
        sColKey[(w * 2 + (n_done & 1)) * 128 + r] = drop_colkey(seed, static_cast<uint32_t>(bh) * p.T + k0 + r);
        asm volatile("bar.sync %0, 128;" ::"r"(2 + w) : "memory");
      }
      // ---- the whole S row of this thread: TMEM -> registers, once ----------------------------
      uint32_t sv[4][32];
      tmem_ld_32x32b_x32(tS, sv[0]);
      tmem_ld_32x32b_x32(tS + 32, sv[1]);
      tmem_ld_32x32b_x32(tS + 64, sv[2]);
      tmem_ld_32x32b_x32(tS + 96, sv[3]);
      tmem_ld_wait();
      if (early_s) {
        tc_fence_before();
        mbar_arrive(&s_free[w]);                        // the MMA warp may refill this S buffer with block j+2
      }
      float mx[4] = {NEG_BIG, NEG_BIG, NEG_BIG, NEG_BIG};
#pragma unroll
      for (int c = 0; c < 4; ++c) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          if (MASK) {
            if (k0 + c * 32 + i >= limit) sv[c][i] = __float_as_uint(NEG_BIG);   // exp2 of it is 0 below
          }
          mx[i & 3] = fmaxf(mx[i & 3], __uint_as_float(sv[c][i]));
        }
      }
      const float m_blk = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * p.scale_log2;
      // the previous P V of this warpgroup read the P buffer written below and produced the O that may be rescaled
      if (n_done > 0) {
        mbar_wait(&o_full[w], oph);
        oph ^= 1;
        tc_fence_after();
      }
      const bool grow = m_blk > m_ref + 8.f;            // first block: m_ref = NEG_BIG
      if (__any_sync(0xffffffffu, grow)) {              // tcgen05.ld / st are warp-wide: rows that stay use alpha = 1
        const float m_new = grow ? m_blk : m_ref;
        const float alpha = exp2f(m_ref - m_new);       // first block: 0
        if (n_done > 0) {
#pragma unroll
          for (int ch = 0; ch < AT_HD / 16; ++ch) {
            uint32_t v[16];
            tmem_ld_32x32b_x16(tO + ch * 16, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
            tmem_st_32x32b_x16(tO + ch * 16, v);
          }
          tmem_st_wait();
        }
        l_run *= alpha;
        m_ref = m_new;
      }
      // ---- P = exp2(s*scale - m_ref), row sum, bf16 -> swizzled smem --------------------------
      float rs[4] = {0.f, 0.f, 0.f, 0.f};
      const float neg_m = -m_ref;
#pragma unroll
      for (int c = 0; c < AT_BN / 32; ++c) {
        float pv[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float e = exp2f(fmaf(__uint_as_float(sv[c][i]), p.scale_log2, neg_m));
          rs[i & 3] += e;
          pv[i] = e;
        }
        if (drop) {
#pragma unroll
          for (int t4 = 0; t4 < 32; t4 += 4) {
            const uint4 b4 = *reinterpret_cast<const uint4*>(colkeys + c * 32 + t4);
            pv[t4 + 0] = drop_keep2d(rowkey, b4.x, thr32) ? pv[t4 + 0] : 0.f;
            pv[t4 + 1] = drop_keep2d(rowkey, b4.y, thr32) ? pv[t4 + 1] : 0.f;
            pv[t4 + 2] = drop_keep2d(rowkey, b4.z, thr32) ? pv[t4 + 2] : 0.f;
            pv[t4 + 3] = drop_keep2d(rowkey, b4.w, thr32) ? pv[t4 + 3] : 0.f;
          }
        }
        uint8_t* dst = myP + (c >> 1) * (AT_BM * 128);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 u;
          u.x = pack_bf16x2(pv[g * 8 + 0], pv[g * 8 + 1]);
          u.y = pack_bf16x2(pv[g * 8 + 2], pv[g * 8 + 3]);
          u.z = pack_bf16x2(pv[g * 8 + 4], pv[g * 8 + 5]);
          u.w = pack_bf16x2(pv[g * 8 + 6], pv[g * 8 + 7]);
          const int chunk = ((c & 1) * 4 + g) ^ (r & 7);
          *reinterpret_cast<uint4*>(dst + chunk * 16) = u;
        }
      }
      l_run += (rs[0] + rs[1]) + (rs[2] + rs[3]);
    };

    for (int j = w; j < nkv; j += 2) {
      mbar_wait(&s_full[w], sph);
      sph ^= 1;
      tc_fence_after();
      const int k0 = j * AT_BN;
      const bool need_mask = (k0 + AT_BN > p.T) || (p.causal && j == qb);
      if (need_mask) {
        const int limit = p.causal ? min(p.T, qrow + 1) : p.T;
        if constexpr (LAZY) block_body_lazy(std::true_type{}, k0, limit);
        else block_body(std::true_type{}, k0, limit);
      } else {
        if constexpr (LAZY) block_body_lazy(std::false_type{}, k0, 0);
        else block_body(std::false_type{}, k0, 0);
      }
      tc_fence_before();
      fence_proxy_async();       // generic-proxy smem writes -> visible to the tensor core
      mbar_arrive(&p_full[w]);
      ++n_done;
    }
    if constexpr (LAZY) {
      if (n_done > 0) {          // the warpgroup's O (relative to m_ref) leaves tensor memory once
        mbar_wait(&o_full[w], oph);
        oph ^= 1;
        tc_fence_after();
        uint32_t v0[32], v1[32];
        tmem_ld_32x32b_x32(tO, v0);
        tmem_ld_32x32b_x32(tO + 32, v1);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          o[i] = __uint_as_float(v0[i]);
          o[32 + i] = __uint_as_float(v1[i]);
        }
      }
    } else {
      if (n_done > 0) accumulate_prev();
    }

    // ---- merge the two warpgroups and write out -------------------------------------------
    // warpgroup 1 parks (m, l, O) in shared memory (its own P buffer is idle by now).
    float* park = reinterpret_cast<float*>(sP + AT_PBYTES);  // [64][128] fp32, column-major by row
    if (w == 1) {
      merge_ml[r] = m_ref;
      merge_ml[128 + r] = l_run;
#pragma unroll
      for (int i = 0; i < AT_HD; ++i) park[i * 128 + r] = o[i];
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (w == 0) {
      const float m1 = merge_ml[r], l1 = merge_ml[128 + r];
      const float m = fmaxf(m_ref, m1);
      const float a0 = exp2f(m_ref - m), a1 = exp2f(m1 - m);
      const float l = l_run * a0 + l1 * a1;
      const float inv = l > 0.f ? p.drop_scale / l : 0.f;   // drop_scale = 1/(1-p): dropout rescale of P, folded in here
      if (qrow < p.T) {
        __nv_bfloat16* dst = p.out + static_cast<size_t>(row_base + qrow) * p.d + h * AT_HD;
#pragma unroll
        for (int g = 0; g < AT_HD / 8; ++g) {
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i)
            f[i] = (o[g * 8 + i] * a0 + park[(g * 8 + i) * 128 + r] * a1) * inv;
          uint4 u;
          u.x = pack_bf16x2(f[0], f[1]);
          u.y = pack_bf16x2(f[2], f[3]);
          u.z = pack_bf16x2(f[4], f[5]);
          u.w = pack_bf16x2(f[6], f[7]);
          *reinterpret_cast<uint4*>(dst + g * 8) = u;
        }
        p.lse[static_cast<size_t>(bh) * p.T + qrow] = (m + log2f(l)) * 0.6931471805599453f;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace b200

using namespace b200;

// 1 = lazy rescale with O in tensor memory (default), 0 = exact per-block rescale with O in registers.
// B200BENCH_ATTN_FWD=exact picks 0 at start-up; b200_set_attention_fwd_variant switches at run time (tests, A/B).
static int& attn_fwd_variant() {
  static int v = [] {
    const char* e = getenv("B200BENCH_ATTN_FWD");
    return (e && e[0] == 'e') ? 0 : 1;
  }();
  return v;
}
extern "C" int b200_set_attention_fwd_variant(int lazy) {
  const int old = attn_fwd_variant();
  if (lazy >= 0) attn_fwd_variant() = lazy ? 1 : 0;
  return old;
}

extern "C" int b200_attention_fwd(const void* qkv, void* out, float* lse, int B, int T, int H,
                                  int head_dim, int causal, float drop_p, unsigned int seed,
                                  void* stream) {
  if (head_dim != AT_HD) return -5;
  const int d = H * head_dim;
  CUtensorMap map;
  int rc = make_tmap_2d_bf16(&map, qkv, static_cast<uint64_t>(B) * T, 3 * d, 3 * d, AT_BN, 64);
  if (rc) return rc;
  static bool attr_done[64] = {false};
  const bool lazy = attn_fwd_variant() != 0;
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_done[dev & 63]) {
    cudaError_t e = cudaFuncSetAttribute(attn_fwd_kernel<false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(attn_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AT_SMEM);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done[dev & 63] = true;
  }
  AttnFwdParams p;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.lse = lse;
  p.B = B;
  p.T = T;
  p.H = H;
  p.d = d;
  p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(head_dim));
  p.causal = causal;
  if (drop_p > 0.f && drop_p < 1.f) {
    p.drop_threshold = static_cast<uint32_t>(static_cast<double>(drop_p) * 4294967296.0);
    p.drop_scale = 1.f / (1.f - drop_p);
  } else {
    p.drop_threshold = 0;
    p.drop_scale = 1.f;
  }
  p.drop_seed = seed;
  p.seed_ptr = step_seed_ptr();
  const int grid = B * H * ((T + AT_BM - 1) / AT_BM);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(AT_THREADS);
  cfg.dynamicSmemBytes = AT_SMEM;
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // PDL: prologue overlaps the previous kernel's tail
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return static_cast<int>(lazy ? cudaLaunchKernelEx(&cfg, attn_fwd_kernel<true>, map, p)
                              : cudaLaunchKernelEx(&cfg, attn_fwd_kernel<false>, map, p));
}
