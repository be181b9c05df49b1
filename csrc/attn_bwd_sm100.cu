// Fused attention backward for sm_100a (recompute-based, FlashAttention-2 style loop order) with
// all five matrix products on tcgen05 tensor cores and their accumulators in tensor memory.
//
// The reference gets this through autograd over bmm/softmax/dropout/bmm and keeps the fp32
// [B*H, T, T] probabilities alive for it (benchmarking/train_harness.py:127,374; SURVEY.md §2.6
// K4, §5.7); here only O and the log-sum-exp are saved and P is recomputed tile by tile.
//
// CTA = one (batch, head, 128-key block j); it keeps K_j, V_j resident and streams the query
// blocks i.  Everything is computed TRANSPOSED (rows = keys, columns = queries) so that each
// thread owns one key row in tensor memory and writes its row of P^T / dS^T straight into the
// 128B-swizzled K-major shared-memory layout the next MMAs consume:
//   S^T  = K_j  Q_i^T          A = K_j (K-major)    B = Q_i  (K-major)    M128 N128 K64
//   dP^T = V_j  dO_i^T         A = V_j (K-major)    B = dO_i (K-major)    M128 N128 K64
//   dV_j += P^T  dO_i          A = P^T (K-major)    B = dO_i (MN-major)   M128 N64  K128
//   dK_j += dS^T Q_i           A = dS^T (K-major)   B = Q_i  (MN-major)   M128 N64  K128
//   dQ_i  = dS   K_j           A = dS^T buffer read MN-major,  B = K_j (MN-major)  M128 N64 K128
// dQ_i is added to an fp32 accumulator in global memory (vector red.add) because every key block
// contributes to it; a small kernel converts it to bf16 afterwards.
//
// 384 threads: warpgroup 0 = TMA producer (warp 0) + MMA issuer (warp 1), warpgroups 1-2 = compute
// (two threads per key row, each covering 64 of the 128 query columns); registers re-balanced
// with setmaxnreg (40 / 232).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <type_traits>

#include "b200_internal.h"
#include "sm100_ptx.cuh"

namespace b200 {

constexpr int AB_BLK = 128;
constexpr int AB_HD = 64;
constexpr int AB_THREADS = 384;  // warpgroup 0: TMA + MMA (+2 idle warps), warpgroups 1-2: compute
constexpr int AB_TILE = AB_BLK * AB_HD * 2;    // 16 KiB
constexpr int AB_PBYTES = AB_BLK * AB_BLK * 2; // 32 KiB
constexpr int AB_SMEM = 2 * AB_TILE /*K,V*/ + 4 * AB_TILE /*Q,dO x2*/ + 4 * AB_PBYTES /*P,dS x2*/ +
                        3 * AB_BLK * 4 /*lse,D,keys*/ + 256 /*bars*/;   // 231168 B of the 232448 B limit

struct AttnBwdParams {
  const float* lse;
  const float* D;
  float* dq_acc;
  __nv_bfloat16* dqkv;
  int B, T, H, d;
  float scale, scale_log2;
  int causal;
  uint32_t drop_threshold;
  float drop_scale;
  uint32_t drop_seed;
  const uint32_t* seed_ptr;
  int diag_skip_dq;   // diagnostics only (B200BENCH_ATTN_BWD=noreduce): leave out the dQ reductions, result is wrong
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c),
               "f"(d)
               : "memory");
}

// EARLY: the S^T / dP^T columns are released as soon as the compute warps hold the tile in registers, so the next
// block's scores are computed during this block's exponentials (B200BENCH_ATTN_BWD=serial selects the other order).
template <bool EARLY>
__global__ void __launch_bounds__(AB_THREADS, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap map_qkv,
                const __grid_constant__ CUtensorMap map_do, const AttnBwdParams p) {
  // SWIZZLE_128B tiles need 1024-byte alignment; declaring the alignment (instead of rounding the
  // pointer up through an integer) keeps the address space known, so accesses are LDS/STS.
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* sK = smem;
  uint8_t* sV = sK + AB_TILE;
  uint8_t* sQ = sV + AB_TILE;            // [2]
  uint8_t* sDO = sQ + 2 * AB_TILE;       // [2]
  uint8_t* sP = sDO + 2 * AB_TILE;       // P^T  [2 buffers][128 keys][128 queries] bf16, two 64-query chunks each
  uint8_t* sDS = sP + 2 * AB_PBYTES;     // dS^T same layout; buffer = query-block index & 1
  float* sLSE = reinterpret_cast<float*>(sDS + 2 * AB_PBYTES);  // [128] (already times log2e)
  float* sD = sLSE + AB_BLK;                                 // [128]
  uint32_t* sKey = reinterpret_cast<uint32_t*>(sD + AB_BLK); // [128] dropout row keys
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKey + AB_BLK);
  uint64_t* kv_full = bars;          // 1
  uint64_t* qdo_full = bars + 1;     // [2]
  uint64_t* qdo_empty = bars + 3;    // [2]
  uint64_t* sdp_full = bars + 5;     // 1
  uint64_t* pds_full = bars + 6;     // 1 (256 arrivals)
  uint64_t* dq_full = bars + 7;      // 1
  uint64_t* dq_empty = bars + 8;     // 1 (256 arrivals)
  uint64_t* sdp_free = bars + 9;     // 1 (256 arrivals): S^T / dP^T of the block are in registers
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nblk = (p.T + AB_BLK - 1) / AB_BLK;
  const int jb = blockIdx.x % nblk;
  const int bh = blockIdx.x / nblk;
  const int b = bh / p.H, h = bh % p.H;
  const int k0 = jb * AB_BLK;
  const int row_base = b * p.T;
  const int i_begin = p.causal ? jb : 0;   // causal: only query blocks at or after the key block

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_qkv);
    tma_prefetch_desc(&map_do);
    mbar_init(kv_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&qdo_full[s], 1);
      mbar_init(&qdo_empty[s], 1);
    }
    mbar_init(sdp_full, 1);
    mbar_init(pds_full, 256);
    mbar_init(dq_full, 1);
    mbar_init(dq_empty, 256);
    mbar_init(sdp_free, 256);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();   // dependents may start their prologue; they wait for this grid's completion themselves
  pdl_wait();      // nothing below may run before the previous kernel's writes are visible
  constexpr uint32_t TC_S = 0, TC_DP = 128, TC_DV = 256, TC_DK = 320, TC_DQ = 384;

  if (warp < 4) {
    setmaxnreg_dec<40>();
  if (warp == 0) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      const int qcol = h * AB_HD, kcol = p.d + h * AB_HD, vcol = 2 * p.d + h * AB_HD;
      mbar_expect_tx(kv_full, 2 * AB_TILE);
      tma_load_2d(&map_qkv, kv_full, sK, kcol, row_base + k0);
      tma_load_2d(&map_qkv, kv_full, sV, vcol, row_base + k0);
      int st = 0;
      uint32_t ph = 0;
      for (int i = i_begin; i < nblk; ++i) {
        mbar_wait(&qdo_empty[st], ph ^ 1);
        mbar_expect_tx(&qdo_full[st], 2 * AB_TILE);
        tma_load_2d(&map_qkv, &qdo_full[st], sQ + st * AB_TILE, qcol, row_base + i * AB_BLK);
        tma_load_2d(&map_do, &qdo_full[st], sDO + st * AB_TILE, h * AB_HD, row_base + i * AB_BLK);
        if (++st == 2) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    if (lane == 0) {
      const uint32_t id_sdp = umma_idesc_bf16(AB_BLK, AB_BLK, false, false);
      const uint32_t id_dvk = umma_idesc_bf16(AB_BLK, AB_HD, false, true);
      const uint32_t id_dq = umma_idesc_bf16(AB_BLK, AB_HD, true, true);
      const uint32_t aK = smem_u32(sK), aV = smem_u32(sV);
      mbar_wait(kv_full, 0);
      tc_fence_after();
      int st = 0;               // Q/dO stage of the block whose S,dP are issued next
      uint32_t ph = 0, it_ph = 0;
      auto issue_sdp = [&]() {  // S^T and dP^T of the next query block (its Q/dO are in stage st)
        mbar_wait(&qdo_full[st], ph);
        tc_fence_after();
        const uint32_t aQ = smem_u32(sQ + st * AB_TILE), aDO = smem_u32(sDO + st * AB_TILE);
#pragma unroll
        for (int k = 0; k < AB_HD / 16; ++k)
          umma_bf16_ss(tmem_base + TC_S, umma_smem_desc(aK + k * 32, 16, 1024),
                       umma_smem_desc(aQ + k * 32, 16, 1024), id_sdp, k ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < AB_HD / 16; ++k)
          umma_bf16_ss(tmem_base + TC_DP, umma_smem_desc(aV + k * 32, 16, 1024),
                       umma_smem_desc(aDO + k * 32, 16, 1024), id_sdp, k ? 1u : 0u);
        umma_commit(sdp_full);
        if (++st == 2) { st = 0; ph ^= 1; }
      };
      if (i_begin < nblk) issue_sdp();
      for (int i = i_begin; i < nblk; ++i) {
        const int buf = (i - i_begin) & 1;      // P/dS buffer and Q/dO stage of block i
        if (EARLY && i + 1 < nblk) {
          // the compute warps hold S^T(i) / dP^T(i) in registers: the next block's scores are computed while they
          // are still busy with the exponentials of this one, so they never wait for the tensor cores
          mbar_wait(sdp_free, it_ph);
          tc_fence_after();
          issue_sdp();
        }
        mbar_wait(pds_full, it_ph);             // P^T(i), dS^T(i) are in shared memory; S/dP TMEM is free
        tc_fence_after();
        if (!EARLY && i + 1 < nblk) issue_sdp();   // next block's scores first: the compute warps start
                                                         // on them while the three products below run
        mbar_wait(dq_empty, it_ph ^ 1);         // previous dQ tile has been drained
        tc_fence_after();
        const uint32_t aQ = smem_u32(sQ + buf * AB_TILE), aDO = smem_u32(sDO + buf * AB_TILE);
        const uint32_t aP = smem_u32(sP + buf * AB_PBYTES), aDS = smem_u32(sDS + buf * AB_PBYTES);
        const uint32_t acc = (i > i_begin) ? 1u : 0u;
#pragma unroll
        for (int k = 0; k < AB_BLK / 16; ++k) {
          const uint32_t a_off = (k >> 2) * (AB_BLK * 128) + (k & 3) * 32;
          umma_bf16_ss(tmem_base + TC_DV, umma_smem_desc(aP + a_off, 16, 1024),
                       umma_smem_desc(aDO + k * 2048, AB_BLK * 128, 1024), id_dvk, (k ? 1u : acc));
        }
#pragma unroll
        for (int k = 0; k < AB_BLK / 16; ++k) {
          const uint32_t a_off = (k >> 2) * (AB_BLK * 128) + (k & 3) * 32;
          umma_bf16_ss(tmem_base + TC_DK, umma_smem_desc(aDS + a_off, 16, 1024),
                       umma_smem_desc(aQ + k * 2048, AB_BLK * 128, 1024), id_dvk, (k ? 1u : acc));
        }
#pragma unroll
        for (int k = 0; k < AB_BLK / 16; ++k)
          umma_bf16_ss(tmem_base + TC_DQ, umma_smem_desc(aDS + k * 2048, AB_BLK * 128, 1024),
                       umma_smem_desc(aK + k * 2048, AB_BLK * 128, 1024), id_dq, k ? 1u : 0u);
        umma_commit(&qdo_empty[buf]);
        umma_commit(dq_full);
        it_ph ^= 1;
      }
    }
  }
  } else {
    setmaxnreg_inc<232>();
    // ================================ compute warps ===============================
    const int wg = (warp - 4) >> 2;      // which 64-query half of the block this thread covers
    const uint32_t seed = p.drop_seed ^ ((p.drop_threshold && p.seed_ptr) ? __ldg(p.seed_ptr) : 0u);
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;   // key row (S^T, dP^T, dV, dK) / query row (dQ)
    const int krow = k0 + r;
    const uint32_t t_lane = static_cast<uint32_t>(quarter * 32) << 16;
    const int tid = threadIdx.x - 128;   // 0..255
    const bool drop = p.drop_threshold != 0;
    const uint32_t thr32 = p.drop_threshold;   // p * 2^32
    const float lg_rescale = drop ? log2f(p.drop_scale) : 0.f;   // drop_scale = 1/(1-p)
    const float inv_rescale = drop ? 1.f / p.drop_scale : 1.f;
    const uint32_t colkey = drop_colkey(seed, static_cast<uint32_t>(bh) * p.T + krow);   // b(k) of this thread's key
    // 32-bit shared addresses, taken once (see lds_v4 / sts_v4 in sm100_ptx.cuh for why)
    const uint32_t aLSE = smem_u32(sLSE), aD = smem_u32(sD), aKey = smem_u32(sKey);
    const uint32_t aP0 = smem_u32(sP), aDS0 = smem_u32(sDS);
    uint32_t rowP = 0;         // set per query block: P/dS buffer (i & 1)
    uint32_t rowS = 0;
    uint32_t it_ph = 0;

    // One (key block, query block) tile for this thread's key row and its 64 query columns.
    // EDGE (compile-time) adds the sequence-tail / causal predicates; interior tiles carry none.
    auto tile_body = [&](auto edge_c, auto drop_c, int q0) {
      constexpr bool EDGE = decltype(edge_c)::value;
      constexpr bool DROP = decltype(drop_c)::value;
      const bool row_ok = krow < p.T;
      const int q_lo = p.causal ? krow : 0;   // valid queries: q_lo <= q < T
      uint32_t vs[2][32], vd[2][32];
      tmem_ld_32x32b_x32(tmem_base + t_lane + TC_S + wg * 64, vs[0]);
      tmem_ld_32x32b_x32(tmem_base + t_lane + TC_DP + wg * 64, vd[0]);
      if constexpr (EARLY) {
        // the whole tile goes to registers at once; the S / dP columns in tensor memory are free from here on
        tmem_ld_32x32b_x32(tmem_base + t_lane + TC_S + wg * 64 + 32, vs[1]);
        tmem_ld_32x32b_x32(tmem_base + t_lane + TC_DP + wg * 64 + 32, vd[1]);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(sdp_free);
      }
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int col0 = wg * 64 + c * 32;
        if constexpr (!EARLY) {
          tmem_ld_wait();
          if (c == 0) {
            tmem_ld_32x32b_x32(tmem_base + t_lane + TC_S + wg * 64 + 32, vs[1]);
            tmem_ld_32x32b_x32(tmem_base + t_lane + TC_DP + wg * 64 + 32, vd[1]);
          }
        }
        float pt[32], ds[32];
#pragma unroll
        for (int e4 = 0; e4 < 32; e4 += 4) {
          const uint4 l4 = lds_v4(aLSE + (col0 + e4) * 4);
          const uint4 d4 = lds_v4(aD + (col0 + e4) * 4);
          const float ls[4] = {__uint_as_float(l4.x), __uint_as_float(l4.y), __uint_as_float(l4.z), __uint_as_float(l4.w)};
          const float dd[4] = {__uint_as_float(d4.x), __uint_as_float(d4.y), __uint_as_float(d4.z), __uint_as_float(d4.w)};
          uint32_t kk[4] = {0, 0, 0, 0};
          if (DROP) {
            const uint4 k4 = lds_v4(aKey + (col0 + e4) * 4);
            kk[0] = k4.x; kk[1] = k4.y; kk[2] = k4.z; kk[3] = k4.w;
          }
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const int e = e4 + t;
            float pv = exp2f(fmaf(__uint_as_float(vs[c][e]), p.scale_log2, -ls[t]));
            if (EDGE) {
              const int q = q0 + col0 + e;
              pv = (row_ok && q < p.T && q >= q_lo) ? pv : 0.f;
            }
            float dp = __uint_as_float(vd[c][e]);
            float pd = pv;
            if (DROP) {
              // dS = P * (keep ? dP/(1-p) : 0 - D) and the dropped, rescaled P feeds dV.  Here pv already is
              // P/(1-p) (the rescale rides on the exponent: staged LSE minus log2(1/(1-p))) and dd is D*(1-p), so
              // dS = pv * (keep ? dP : 0 - dd) and the per-element work is two selects, no multiply.
              const bool keep = drop_keep2d(kk[t], colkey, thr32);
              dp = keep ? dp : 0.f;
              pd = keep ? pv : 0.f;
            }
            pt[e] = pd;
            ds[e] = pv * (dp - dd[t]);
          }
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int chunk = (c * 4 + g) ^ (r & 7);
          uint4 u;
          u.x = pack_bf16x2(pt[g * 8 + 0], pt[g * 8 + 1]);
          u.y = pack_bf16x2(pt[g * 8 + 2], pt[g * 8 + 3]);
          u.z = pack_bf16x2(pt[g * 8 + 4], pt[g * 8 + 5]);
          u.w = pack_bf16x2(pt[g * 8 + 6], pt[g * 8 + 7]);
          sts_v4(rowP + chunk * 16, u);
          u.x = pack_bf16x2(ds[g * 8 + 0], ds[g * 8 + 1]);
          u.y = pack_bf16x2(ds[g * 8 + 2], ds[g * 8 + 3]);
          u.z = pack_bf16x2(ds[g * 8 + 4], ds[g * 8 + 5]);
          u.w = pack_bf16x2(ds[g * 8 + 6], ds[g * 8 + 7]);
          sts_v4(rowS + chunk * 16, u);
        }
      }
    };

    // dQ of query block qi: rows = queries, this thread adds 32 of the 64 head-dim columns
    auto drain_dq = [&](int qi, uint32_t phase) {
      mbar_wait(dq_full, phase);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld_32x32b_x32(tmem_base + t_lane + TC_DQ + wg * 32, v);
      tmem_ld_wait();
      const int q = qi * AB_BLK + r;
      // Every key block adds into dQ: red.global.add.v4.f32 from registers.  Measured at the headline shape: leaving
      // the reductions out takes the kernel from 126 to 111 us; handing them to the copy engine instead was slower
      // (one 128-byte bulk reduce-add per thread from shared memory: 157 us - its addresses are uniform registers, so
      // a warp issues them lane by lane; one [32 x 32] fp32 tile reduce-add per warp through a tensor map: 141 us).
      if (q < p.T && !p.diag_skip_dq) {
        float* dst = p.dq_acc + static_cast<size_t>(row_base + q) * p.d + h * AB_HD + wg * 32;
#pragma unroll
        for (int e = 0; e < 32; e += 4)
          red_add_v4(dst + e, __uint_as_float(v[e]) * p.scale, __uint_as_float(v[e + 1]) * p.scale,
                     __uint_as_float(v[e + 2]) * p.scale, __uint_as_float(v[e + 3]) * p.scale);
      }
      tc_fence_before();
      mbar_arrive(dq_empty);
    };

    // per-query statistics of block qi for this thread: LSE (threads 0..127) or D (threads 128..255)
    auto load_stat = [&](int qi) -> float {
      const int q = qi * AB_BLK + (tid & (AB_BLK - 1));
      if (qi >= nblk || q >= p.T) return 0.f;
      return (tid < AB_BLK) ? p.lse[static_cast<size_t>(bh) * p.T + q] : p.D[static_cast<size_t>(bh) * p.T + q];
    };
    float pre_val = load_stat(i_begin);

    for (int i = i_begin; i < nblk; ++i) {
      const int q0 = i * AB_BLK;
      const int buf = (i - i_begin) & 1;
      rowP = aP0 + buf * AB_PBYTES + wg * (AB_BLK * 128) + r * 128;
      rowS = aDS0 + buf * AB_PBYTES + wg * (AB_BLK * 128) + r * 128;
      // stage LSE (times log2e), D and the dropout row keys for the 128 queries of this block; the
      // global loads were issued one iteration ahead (pre_val), so nobody stalls on them here
      if (tid < AB_BLK) {
        sts_u32(aLSE + tid * 4, __float_as_uint(pre_val * 1.4426950408889634f - lg_rescale));
      } else {
        sts_u32(aD + (tid - AB_BLK) * 4, __float_as_uint(pre_val * inv_rescale));
        if (drop) sts_u32(aKey + (tid - AB_BLK) * 4, drop_rowkey(seed, static_cast<uint32_t>(bh) * p.T + q0 + tid - AB_BLK));
      }
      pre_val = load_stat(i + 1);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      mbar_wait(sdp_full, it_ph);
      tc_fence_after();
      // buffer (i & 1) was last read by the products of block i-2, whose completion (dq_full) every
      // thread observed when it drained dQ(i-2) during the previous iteration
      const bool edge = (k0 + AB_BLK > p.T) || (q0 + AB_BLK > p.T) || (p.causal && i == jb);
      if (drop) {
        if (edge) tile_body(std::true_type{}, std::true_type{}, q0); else tile_body(std::false_type{}, std::true_type{}, q0);
      } else {
        if (edge) tile_body(std::true_type{}, std::false_type{}, q0); else tile_body(std::false_type{}, std::false_type{}, q0);
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");   // everyone is done with sLSE / sD / sKey
      tc_fence_before();
      fence_proxy_async();
      mbar_arrive(pds_full);
      // the previous block's dQ is (about to be) complete: drain it while S,dP of the next block and the
      // dV/dK/dQ products of this one run on the tensor cores
      if (i > i_begin) drain_dq(i - 1, it_ph ^ 1);
      it_ph ^= 1;
    }
    if (i_begin < nblk) drain_dq(nblk - 1, it_ph ^ 1);
    // ---- epilogue: dV (warpgroup 0) and dK (warpgroup 1), rows = keys ------------------------
    // The last dq_full wait (final drain above) also covers the final dV / dK MMAs (same commit).
    if (i_begin < nblk) {
      const uint32_t tcol = (wg == 0) ? TC_DV : TC_DK;
      const float oscale = (wg == 0) ? 1.f : p.scale;   // softmax scale folded out of dS
      const bool ok = krow < p.T;   // tcgen05.ld is warp-collective: only the stores are predicated
      __nv_bfloat16* dst = p.dqkv + static_cast<size_t>(row_base + (ok ? krow : k0)) * 3 * p.d +
                           (wg == 0 ? 2 * p.d : p.d) + h * AB_HD;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + t_lane + tcol + c * 32, v);
        tmem_ld_wait();
        if (ok) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4 u;
            u.x = pack_bf16x2(__uint_as_float(v[g * 8 + 0]) * oscale, __uint_as_float(v[g * 8 + 1]) * oscale);
            u.y = pack_bf16x2(__uint_as_float(v[g * 8 + 2]) * oscale, __uint_as_float(v[g * 8 + 3]) * oscale);
            u.z = pack_bf16x2(__uint_as_float(v[g * 8 + 4]) * oscale, __uint_as_float(v[g * 8 + 5]) * oscale);
            u.w = pack_bf16x2(__uint_as_float(v[g * 8 + 6]) * oscale, __uint_as_float(v[g * 8 + 7]) * oscale);
            *reinterpret_cast<uint4*>(dst + c * 32 + g * 8) = u;
          }
        }
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

// D[b,h,t] = sum_c dO[b,t,h,c] * O[b,t,h,c]   (one warp per (row, head))
__global__ void __launch_bounds__(256)
attn_bwd_prep_kernel(const __nv_bfloat16* __restrict__ O, const __nv_bfloat16* __restrict__ dO,
                     float* __restrict__ D, int B, int T, int H) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (gw >= B * T * H) return;
  const int m = gw / H, h = gw % H;
  const size_t off = static_cast<size_t>(m) * H * AB_HD + h * AB_HD + lane * 2;
  const float2 o = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(O + off));
  const float2 g = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(dO + off));
  float s = o.x * g.x + o.y * g.y;
#pragma unroll
  for (int k = 16; k > 0; k >>= 1) s += __shfl_xor_sync(0xffffffffu, s, k);
  if (lane == 0) {
    const int b = m / T, t = m % T;
    D[(static_cast<size_t>(b) * H + h) * T + t] = s;
  }
}

// dq_acc fp32 [M, d] -> bf16 into dqkv[:, 0:d] (row stride 3d).  The accumulator is zeroed again on the way
// out, so the next backward call finds it ready and no memset pass is needed.
__global__ void __launch_bounds__(256)
attn_bwd_dq_convert_kernel(float* __restrict__ acc, __nv_bfloat16* __restrict__ dqkv, int M,
                           int d) {
  pdl_wait();
  const int nvec = d >> 3;
  const size_t total = static_cast<size_t>(M) * nvec;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t m = i / nvec;
    const int c = static_cast<int>(i % nvec);
    float4* src = reinterpret_cast<float4*>(acc + m * d) + 2 * c;
    const float4 a = src[0];
    const float4 bq = src[1];
    src[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    src[1] = make_float4(0.f, 0.f, 0.f, 0.f);
    uint4 u;
    u.x = pack_bf16x2(a.x, a.y);
    u.y = pack_bf16x2(a.z, a.w);
    u.z = pack_bf16x2(bq.x, bq.y);
    u.w = pack_bf16x2(bq.z, bq.w);
    reinterpret_cast<uint4*>(dqkv + m * 3 * d)[c] = u;
  }
}

}  // namespace b200

using namespace b200;

static int attention_bwd_impl(const void* qkv, const void* out, const void* dout,
                              const float* lse, void* dqkv, float* D, float* dq_acc, int B,
                              int T, int H, int head_dim, int causal, float drop_p,
                              unsigned int seed, void* stream, int flags) {
  if (head_dim != AB_HD) return -5;
  const int d = H * head_dim;
  const int M = B * T;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  CUtensorMap map_qkv, map_do;
  int rc = make_tmap_2d_bf16(&map_qkv, qkv, M, 3 * d, 3 * d, AB_BLK, 64);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&map_do, dout, M, d, d, AB_BLK, 64);
  if (rc) return rc;
  static bool attr_done[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_done[dev & 63]) {
    cudaError_t e = cudaFuncSetAttribute(attn_bwd_kernel<true>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, AB_SMEM);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(attn_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AB_SMEM);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done[dev & 63] = true;
  }
  if (!(flags & B200_ATTN_BWD_ACC_IS_ZERO)) cudaMemsetAsync(dq_acc, 0, static_cast<size_t>(M) * d * sizeof(float), s);
  if (!(flags & B200_ATTN_BWD_HAVE_D)) {
    const long warps = static_cast<long>(M) * H;
    const int blocks = static_cast<int>((warps * 32 + 255) / 256);
    attn_bwd_prep_kernel<<<blocks, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(out),
                                                static_cast<const __nv_bfloat16*>(dout), D, B, T, H);
  }
  AttnBwdParams p;
  p.lse = lse;
  p.D = D;
  p.dq_acc = dq_acc;
  p.dqkv = static_cast<__nv_bfloat16*>(dqkv);
  p.B = B;
  p.T = T;
  p.H = H;
  p.d = d;
  p.scale = 1.f / sqrtf(static_cast<float>(head_dim));
  p.scale_log2 = p.scale * 1.4426950408889634f;
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
  static const int early = [] {            // B200BENCH_ATTN_BWD=serial: S / dP of the next block only after P / dS
    const char* e = getenv("B200BENCH_ATTN_BWD");
    return (e && e[0] == 's') ? 0 : 1;
  }();
  static const int skip_dq = [] {
    const char* e = getenv("B200BENCH_ATTN_BWD");
    return (e && e[0] == 'n') ? 1 : 0;
  }();
  p.diag_skip_dq = skip_dq;
  const int grid = B * H * ((T + AB_BLK - 1) / AB_BLK);
  {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(AB_THREADS);
    cfg.dynamicSmemBytes = AB_SMEM;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // PDL: prologue overlaps the prep kernel's tail
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t le = early ? cudaLaunchKernelEx(&cfg, attn_bwd_kernel<true>, map_qkv, map_do, p)
                           : cudaLaunchKernelEx(&cfg, attn_bwd_kernel<false>, map_qkv, map_do, p);
    if (le != cudaSuccess) return static_cast<int>(le);
  }
  {
    const size_t work = static_cast<size_t>(M) * (d / 8);
    size_t blocks = (work + 255) / 256;
    const size_t cap = static_cast<size_t>(device_sm_count()) * 8;
    if (blocks > cap) blocks = cap;
    launch_pdl(attn_bwd_dq_convert_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, s, dq_acc,
               static_cast<__nv_bfloat16*>(dqkv), M, d);
  }
  return static_cast<int>(cudaGetLastError());
}

extern "C" int b200_attention_bwd(const void* qkv, const void* out, const void* dout,
                                  const float* lse, void* dqkv, float* D, float* dq_acc, int B,
                                  int T, int H, int head_dim, int causal, float drop_p,
                                  unsigned int seed, void* stream) {
  return attention_bwd_impl(qkv, out, dout, lse, dqkv, D, dq_acc, B, T, H, head_dim, causal, drop_p, seed, stream, 0);
}

extern "C" int b200_attention_bwd_ex(const void* qkv, const void* out, const void* dout,
                                     const float* lse, void* dqkv, float* D, float* dq_acc, int B,
                                     int T, int H, int head_dim, int causal, float drop_p,
                                     unsigned int seed, void* stream, int flags) {
  return attention_bwd_impl(qkv, out, dout, lse, dqkv, D, dq_acc, B, T, H, head_dim, causal, drop_p, seed, stream,
                            flags);
}
