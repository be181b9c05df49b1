// tcgen05 / TMEM / TMA bf16 GEMM for sm_100a with fused epilogues.
//
//   D[M,N] = epilogue( A[M,K] * B[N,K]^T )      fp32 accumulation in tensor memory
//
// Each operand may be K-major (row-major [rows, K]) or MN-major (row-major [K, rows]), which
// gives the three layouts a linear layer needs without any transpose pass:
//   forward  Y  = X  W^T         A K-major,  B K-major
//   dgrad    dX = dY W           A K-major,  B MN-major
//   wgrad    dW = dY^T X         A MN-major, B MN-major
// Replaces the cuBLAS calls the reference reaches through nn.Linear / F.multi_head_attention
// (benchmarking/train_harness.py:58,113-123,127) and their autograd backward GEMMs
// (SURVEY.md §2.6 K3,K5-K8,K10).
//
// Structure: persistent, warp-specialised CTA of 192 threads, one CTA per SM.
//   warp 0      TMA producer  (cp.async.bulk.tensor -> 128B-swizzled smem ring, mbarrier tx)
//   warp 1      MMA issuer    (one elected lane: tcgen05.mma, M=128 (256 for a CTA pair), N=BN, K=16)
//   warps 2..5  epilogue      (tcgen05.ld 32x32b -> registers -> fused epilogue -> smem -> TMA store)
// Two accumulator stages in TMEM (2*BN columns) let the epilogue of tile i overlap the main
// loop of tile i+1.
//
// K14 (the FSDP / ZeRO-3 parameter all-gather, TH:229-237,366) lives in the producer warp:
// when B is a weight matrix that is partitioned over ranks, there is one tensor map per OWNING rank over
// that rank's shard as mapped into this process (peer memory over NVLink), and every B tile is fetched by
// TMA straight from its owner into shared memory - no gather pass, no staging copy in HBM, no flags.
// For a LOCAL B the kernel can also share B inside a thread-block cluster: the CTAs (or CTA pairs) that work on the
// same N-tile for different M-tiles each fetch 1/P of the B tile and multicast it to the others
// (.multicast::cluster).  TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load
// from another GPU's memory hung the GPU: a partitioned B is fetched unicast by every CTA (host and kernel both
// refuse the combination).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <type_traits>

#include "b200_internal.h"
#include "sm100_ptx.cuh"

// Diagnostic build switch: -DGEMM_LANE0_LOOPS runs the producer / MMA loops in lane 0 only (round-1 style).
#ifdef GEMM_LANE0_LOOPS
#define GEMM_LOOP_LANES (lane == 0)
#define GEMM_ELECT() true
#define GEMM_SYNCWARP() ((void)0)
#else
#define GEMM_LOOP_LANES true
#define GEMM_ELECT() elect_one()
#define GEMM_SYNCWARP() __syncwarp()
#endif
#ifndef GEMM_LAUNCH_BOUND
#define GEMM_LAUNCH_BOUND 192
#endif

namespace b200 {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_THREADS = 192;
constexpr int GEMM_MAX_OWNERS = 8;
constexpr int GEMM_MAX_UNITS = 8;       // tile units (CTAs or CTA pairs) per cluster that share B by multicast

template <int BN, int CTAS>
struct GemmCfg {
  static constexpr int kBNLocal = BN / CTAS;             // B rows held by each CTA of the pair
  static constexpr int kABytes = GEMM_BM * GEMM_BK * 2;  // 16 KiB
  static constexpr int kBBytes = kBNLocal * GEMM_BK * 2;
  static constexpr int kStagingBytes = 32768;            // TMA-store staging of the epilogue
  // barriers (256 B) + 2 KiB that hold either the double-buffered fp32 bias slice (2 x BN floats) or, for a
  // weight-gradient GEMM that also produces the bias gradient, the all-ones B tile of the row-sum MMA
  static constexpr int kBarBytes = 256 + 2048;
  static constexpr int kRing = 232448 - kStagingBytes - kBarBytes;   // what is left of the 227 KiB
  static constexpr int kStages = (kRing / (kABytes + kBBytes)) > 8 ? 8 : (kRing / (kABytes + kBBytes));
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTmemCols = 2 * BN;  // 256 or 512, power of two
  static constexpr int kBiasCol = 2 * BN - 32;   // row-sum accumulator (16 columns) of the bias-gradient mode
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + kBarBytes;
  static_assert(2 * BN * 4 <= 2048, "bias slice must fit its 2 KiB");
};

struct BMaps {
  CUtensorMap m[GEMM_MAX_OWNERS];
};

struct GemmParams {
  int M, N, K;
  int a_mn, b_mn;  // 1 = MN-major operand
  int epi;         // B200_EPI_*
  int ldd;         // leading dimension of D / out2 / aux / resid (elements)
  __nv_bfloat16* D;
  __nv_bfloat16* D2;            // second output (GELU activation)
  const __nv_bfloat16* bias;    // [N]
  const __nv_bfloat16* aux;     // residual or pre-activation, [M, ldd]
  float drop_scale;             // 1/(1-p)
  uint32_t drop_threshold;      // p * 2^32, 0 = no dropout
  uint32_t drop_seed;
  const uint32_t* seed_ptr;     // per-step seed word in device memory (nullable)
  int num_m_tiles, num_n_tiles; // tiles of a whole CLUSTER: (units * CTAS * 128) rows x BN columns
  int units;                    // P: tile units per cluster along M; B is multicast across them (1 = off)
  // ---- K14: B partitioned over owning ranks.  Row r of B lives in owner (b_row0 + r) / b_rows_per_owner at
  // local row (b_row0 + r) % b_rows_per_owner of that owner's shard, viewed as a [rows_per_owner, ldb] matrix.
  int b_ws;                     // number of owners (<= 1: B is one local matrix)
  uint32_t b_row0, b_rows_per_owner;
  // ---- weight-gradient extras: dbias[m] (+)= sum_k A[m,k]  (the bias gradient is the row sum of dY^T)
  __nv_bfloat16* dbias;
  int dbias_accum;
  // ---- B200_EPI_ROWDOT: D = acc and rowdot[(b*H + h)*T + t] = sum_{c in head h} acc[m,c] * aux[m,c], m = b*T + t
  float* rowdot;
  int rd_T, rd_hd;
};

// CTAS = 1: one CTA per 128 x BN tile (cta_group::1).
// CTAS = 2: a CTA pair per 256 x BN tile (cta_group::2): each CTA stages its own 128 rows
//           of A and HALF of B, the leader issues M=256 MMAs that read both halves, so the shared-memory
//           traffic per MAC halves for B - at 128x128 per CTA the single-CTA kernel is shared-memory-bound
//           (operand reads + TMA fills = 2 x 128 B/clk), which caps it near 50 % of the tensor peak.
// A cluster holds p.units such tile units (consecutive M-tiles of one N-tile).
template <int BN, int CTAS>
__global__ void __launch_bounds__(GEMM_LAUNCH_BOUND, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a,
                         const __grid_constant__ BMaps maps_b,
                         const __grid_constant__ CUtensorMap map_d,
                         const __grid_constant__ CUtensorMap map_d2, const GemmParams p) {
  using Cfg = GemmCfg<BN, CTAS>;
  constexpr int BNL = Cfg::kBNLocal;
  const int P = p.units;
  if (P > 1 && p.b_ws > 1) __trap();   // multicast loads only from local memory (see the header comment); never reached
  const int CL = P * CTAS;                                  // CTAs per cluster
  const uint32_t crank = (CL > 1) ? cluster_ctarank() : 0u;
  const uint32_t cta_rank = crank % CTAS;                   // which half of a pair
  const uint32_t unit_in_cl = crank / CTAS;
  const bool is_leader = cta_rank == 0;
  const int cluster_id = blockIdx.x / CL;
  const int num_clusters = gridDim.x / CL;
  const uint16_t mask_all = static_cast<uint16_t>((1u << CL) - 1u);
  const uint16_t mask_unit = static_cast<uint16_t>(CTAS == 2 ? (3u << (crank & ~1u)) : (1u << crank));
  uint16_t mask_half = 0;                                   // every CTA that holds the same half of B as this one
  for (int u = 0; u < P; ++u) mask_half |= static_cast<uint16_t>(1u << (u * CTAS + cta_rank));
  // SWIZZLE_128B tiles need 1024-byte alignment; declaring the alignment (instead of rounding the
  // pointer up through an integer) keeps the address space known, so accesses are LDS/STS.
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* tiles = smem;
  uint8_t* stg = smem + Cfg::kStages * Cfg::kStageBytes;   // epilogue staging: 4 warps x 2 x 4 KiB
  uint64_t* bars = reinterpret_cast<uint64_t*>(stg + Cfg::kStagingBytes);
  uint64_t* full_bar = bars;                         // [kStages]
  uint64_t* empty_bar = bars + Cfg::kStages;         // [kStages]
  uint64_t* tfull_bar = bars + 2 * Cfg::kStages;     // [2]
  uint64_t* tempty_bar = bars + 2 * Cfg::kStages + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * Cfg::kStages + 4);
  uint8_t* aux_smem = reinterpret_cast<uint8_t*>(bars) + 256;   // 2 KiB: bias slice or the all-ones tile

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = (p.K + GEMM_BK - 1) / GEMM_BK;
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;
  const bool bias_grad = p.dbias != nullptr;
  const int nacc = bias_grad ? 1 : 2;   // bias-gradient mode gives the second accumulator stage to the row sums

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    for (int o = 0; o < (p.b_ws > 1 ? p.b_ws : 1); ++o) tma_prefetch_desc(&maps_b.m[o]);
    tma_prefetch_desc(&map_d);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], P);     // one commit per tile unit of the cluster that read the slot
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 4 * CTAS);  // one arrive per epilogue warp (of both CTAs of a pair)
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    if (CTAS == 2) tmem_alloc_2cta(tmem_slot, Cfg::kTmemCols); else tmem_alloc(tmem_slot, Cfg::kTmemCols);
  }
  if (bias_grad) {   // B operand of the row-sum MMA: bf16 1.0 everywhere, so layout and swizzle do not matter
    for (int i = threadIdx.x; i < 512; i += GEMM_THREADS) reinterpret_cast<uint32_t*>(aux_smem)[i] = 0x3F803F80u;
    fence_proxy_async();
  }
  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();   // the peers' barriers are initialised before anyone signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Everything above (barrier init, TMEM allocation, descriptor prefetch) may overlap the tail of the
  // previous kernel in the stream; nothing below may start before that kernel's writes are visible.
  pdl_trigger();
  pdl_wait();

  if (warp == 0) {
    // ================================ TMA producer ================================
    // (the whole warp walks the loop with warp-uniform values; one elected lane issues the copies)
    if (GEMM_LOOP_LANES) {
      int stage = 0;
      uint32_t phase = 0;
      // This thread's instruction count per k-block is on the critical path of the whole pipeline (it has to stay
      // under the ~512 cycles the tensor cores need per k-block), so everything that does not change with k is
      // resolved per tile: slice offsets, and for a partitioned B the owner and the local row in its shard.
      const bool remote = p.b_ws > 1;
      const int sr = BNL / P;            // K-major B: rows of this CTA's half that this unit fetches for everybody
      const int kr = GEMM_BK / P;        // MN-major B: k-rows of every 64-column chunk that this unit fetches
      const int b_dst_off = static_cast<int>(unit_in_cl) * (p.b_mn ? kr : sr) * 128;
      const int b_row_add = static_cast<int>(unit_in_cl) * (p.b_mn ? kr : sr);
      auto issue_b = [&](const CUtensorMap* m, void* dst, int c0, int lrow) {
        if (P > 1) {
          if (CTAS == 2) tma_load_2d_2cta_mc(m, &full_bar[stage], dst, c0, lrow, mask_half);
          else tma_load_2d_mc(m, &full_bar[stage], dst, c0, lrow, mask_half);
        } else {
          if (CTAS == 2) tma_load_2d_2cta(m, &full_bar[stage], dst, c0, lrow);
          else tma_load_2d(m, &full_bar[stage], dst, c0, lrow);
        }
      };
      // owner of B row `row` and the row's index inside that owner's shard
      auto resolve = [&](int row, uint32_t& owner, int& lrow) {
        const uint32_t g = p.b_row0 + static_cast<uint32_t>(row);
        owner = g / p.b_rows_per_owner;
        if (owner >= static_cast<uint32_t>(p.b_ws)) owner = p.b_ws - 1;    // past the end: out of the map -> zeros
        lrow = static_cast<int>(g - owner * p.b_rows_per_owner);
      };
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const int m0 = (tile % p.num_m_tiles) * (GEMM_BM * CL) + static_cast<int>(crank) * GEMM_BM;
        const int n0 = (tile / p.num_m_tiles) * BN + static_cast<int>(cta_rank) * BNL;   // this CTA's B rows
        const CUtensorMap* mb = &maps_b.m[0];
        uint32_t owner = 0;
        int brow = (p.b_mn ? 0 : n0) + b_row_add;       // K-major: fixed per tile; MN-major: advances with k
        if (remote) {
          resolve(brow, owner, brow);
          mb = &maps_b.m[owner];
        }
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = tiles + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          const int k0 = kb * GEMM_BK;
          if (GEMM_ELECT()) {
            // pair mode: all bytes (of both CTAs) are credited to the leader's barrier; only it arrives
            if (is_leader) mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes * CTAS);
            auto load_a = [&](void* dst, int c0, int c1) {
              if (CTAS == 2) tma_load_2d_2cta(&map_a, &full_bar[stage], dst, c0, c1);
              else tma_load_2d(&map_a, &full_bar[stage], dst, c0, c1);
            };
            if (!p.a_mn) {
              load_a(sa, k0, m0);
            } else {
#pragma unroll
              for (int c = 0; c < GEMM_BM / 64; ++c) load_a(sa + c * (GEMM_BK * 128), m0 + 64 * c, k0);
            }
            if (!p.b_mn) {
              // K-major B: this CTA's BNL rows; with multicast each unit fetches BNL/P of them for everybody
              issue_b(mb, sb + b_dst_off, k0, brow);
            } else {
              // MN-major B: 64 k-rows x 64 n-columns per chunk; with multicast each unit fetches 64/P k-rows
#pragma unroll
              for (int c = 0; c < BNL / 64; ++c)
                issue_b(mb, sb + c * (GEMM_BK * 128) + b_dst_off, n0 + 64 * c, brow);
            }
          }
          GEMM_SYNCWARP();
          if (p.b_mn) {
            brow += GEMM_BK;
            if (remote && brow >= static_cast<int>(p.b_rows_per_owner) && owner + 1 < static_cast<uint32_t>(p.b_ws)) {
              brow -= static_cast<int>(p.b_rows_per_owner);     // the next k-block belongs to the next owner
              mb = &maps_b.m[++owner];
            }
          }
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    if (is_leader && GEMM_LOOP_LANES) {
      // The whole warp walks the loop (warp-uniform control flow and values, so the compiler keeps descriptors and
      // counters in uniform registers); one elected lane issues.
      // The issue rate of this single thread bounds the whole kernel (four MMAs of 128 tensor-core cycles each per
      // k-block), so the loop body is kept to a few adds per MMA: the shared-memory descriptors of stage 0 are
      // built once, and stage / k-step offsets are added to their address field (bits [0,14), units of 16 bytes;
      // a CTA's shared memory is < 2^18 bytes, so the add never carries out of the field).
      const uint32_t idesc = umma_idesc_bf16(GEMM_BM * CTAS, BN, p.a_mn != 0, p.b_mn != 0);
      const uint32_t idesc_rowsum = umma_idesc_bf16(GEMM_BM * CTAS, 16, p.a_mn != 0, false);
      // K-major: +32 B per 16-element K step inside the 128 B swizzle row.
      // MN-major: +16 k-rows * 128 B = 2048 B per K step; MN chunks are BK*128 B apart.
      const uint32_t a_k = (p.a_mn ? 2048u : 32u) >> 4;
      const uint32_t b_k = (p.b_mn ? 2048u : 32u) >> 4;
      const uint64_t adesc0 = umma_smem_desc(smem_u32(tiles), p.a_mn ? GEMM_BK * 128u : 16u, 1024);
      const uint64_t bdesc0 = umma_smem_desc(smem_u32(tiles) + Cfg::kABytes, p.b_mn ? GEMM_BK * 128u : 16u, 1024);
      const uint64_t odesc0 = umma_smem_desc(smem_u32(aux_smem), 16, 1024);   // all-ones tile, +2 units per k-step
      constexpr uint32_t kStageUnits = Cfg::kStageBytes >> 4;
      const uint16_t empty_mask = (P > 1) ? mask_all : mask_unit;
      auto commit_empty = [&](uint64_t* bar) {
        // smem slot reusable (in every CTA of the cluster that fed it) once the MMAs issued so far retire
        if (CTAS == 2) umma_commit_2cta(bar, empty_mask);
        else if (P > 1) umma_commit_mc(bar, empty_mask);
        else umma_commit(bar);
      };
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const bool rowsum = bias_grad && (tile / p.num_m_tiles) == 0;   // one N-tile is enough for the row sums
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
        if (!rowsum) {
          for (int kb = 0; kb < num_kb; ++kb) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint64_t ad = adesc0 + static_cast<uint32_t>(stage) * kStageUnits;
            const uint64_t bd = bdesc0 + static_cast<uint32_t>(stage) * kStageUnits;
            if (GEMM_ELECT()) {
              if (CTAS == 2) {
                umma_bf16_ss_2cta(tmem_d, ad, bd, idesc, kb ? 1u : 0u);
                umma_bf16_ss_2cta(tmem_d, ad + a_k, bd + b_k, idesc, 1u);
                umma_bf16_ss_2cta(tmem_d, ad + 2 * a_k, bd + 2 * b_k, idesc, 1u);
                umma_bf16_ss_2cta(tmem_d, ad + 3 * a_k, bd + 3 * b_k, idesc, 1u);
              } else {
                umma_bf16_ss(tmem_d, ad, bd, idesc, kb ? 1u : 0u);
                umma_bf16_ss(tmem_d, ad + a_k, bd + b_k, idesc, 1u);
                umma_bf16_ss(tmem_d, ad + 2 * a_k, bd + 2 * b_k, idesc, 1u);
                umma_bf16_ss(tmem_d, ad + 3 * a_k, bd + 3 * b_k, idesc, 1u);
              }
              commit_empty(&empty_bar[stage]);
            }
            GEMM_SYNCWARP();
            if (++stage == Cfg::kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
        } else {
          // weight gradient + bias gradient: [128 x 16] += A[128 x 16k] * ones, every column = the row sum of A
          const uint32_t tmem_rs = tmem_base + Cfg::kBiasCol;
          for (int kb = 0; kb < num_kb; ++kb) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint64_t ad = adesc0 + static_cast<uint32_t>(stage) * kStageUnits;
            const uint64_t bd = bdesc0 + static_cast<uint32_t>(stage) * kStageUnits;
            if (GEMM_ELECT()) {
#pragma unroll
              for (uint32_t k = 0; k < GEMM_BK / 16; ++k) {
                const uint32_t accf = (kb | k) ? 1u : 0u;
                if (CTAS == 2) {
                  umma_bf16_ss_2cta(tmem_d, ad + k * a_k, bd + k * b_k, idesc, accf);
                  umma_bf16_ss_2cta(tmem_rs, ad + k * a_k, odesc0 + 2 * k, idesc_rowsum, accf);
                } else {
                  umma_bf16_ss(tmem_d, ad + k * a_k, bd + k * b_k, idesc, accf);
                  umma_bf16_ss(tmem_rs, ad + k * a_k, odesc0 + 2 * k, idesc_rowsum, accf);
                }
              }
              commit_empty(&empty_bar[stage]);
            }
            GEMM_SYNCWARP();
            if (++stage == Cfg::kStages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        // accumulator complete -> epilogue (of both CTAs)
        if (GEMM_ELECT()) {
          if (CTAS == 2) umma_commit_2cta(&tfull_bar[acc], mask_unit); else umma_commit(&tfull_bar[acc]);
        }
        GEMM_SYNCWARP();
        if (++acc == nacc) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else {
    // ================================ epilogue ====================================
    const int quarter = warp & 3;  // TMEM lane quarter this warp may access
    const int et = threadIdx.x - 64;   // 0..127 among the epilogue threads
    const uint32_t seed = p.drop_seed ^ ((p.drop_threshold && p.seed_ptr) ? __ldg(p.seed_ptr) : 0u);
    float* s_bias = reinterpret_cast<float*>(aux_smem);   // [2][BN] floats, double-buffered per tile
    int acc = 0;
    uint32_t acc_phase = 0;
    int tile_no = 0;

    // One output tile.  EPI is a compile-time constant inside, so each epilogue kind gets a tight loop:
    // aux (residual / pre-activation / old D) for chunk c+1 is requested before chunk c is processed,
    // the bias comes from shared memory, TMEM is read 32 columns at a time.  Results do not go to global
    // memory from the registers (that would be 32 half-filled sectors per warp store): each warp packs
    // its 32 rows x 64 columns into a 128B-swizzled 4 KiB staging tile and one lane hands it to the TMA
    // store engine, which also clips rows >= M and columns >= N.
    uint8_t* stg_w = stg + quarter * 8192;    // this warp's two 4 KiB staging tiles
    uint32_t nst = 0;
    auto run_tile = [&](auto epi_c, int m0, int n0, const float* bias_s) {
      constexpr int EPI = decltype(epi_c)::value;
      constexpr bool HAS_BIAS = (EPI == B200_EPI_BIAS || EPI == B200_EPI_BIAS_GELU || EPI == B200_EPI_BIAS_RESID ||
                                 EPI == B200_EPI_BIAS_GELU_GRAD);
      constexpr bool HAS_AUX = (EPI == B200_EPI_BIAS_RESID || EPI == B200_EPI_DGELU || EPI == B200_EPI_ACCUM ||
                                EPI == B200_EPI_ROWDOT || EPI == B200_EPI_MUL);
      constexpr bool DUAL = (EPI == B200_EPI_BIAS_GELU || EPI == B200_EPI_BIAS_GELU_GRAD);
      const int row = m0 + quarter * 32 + lane;
      const bool row_ok = row < p.M;
      const size_t row_off = static_cast<size_t>(row_ok ? row : 0) * p.ldd;
      const __nv_bfloat16* auxp = (EPI == B200_EPI_ACCUM) ? p.D : p.aux;
      const uint32_t taddr_row =
          tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(acc * BN);
      uint4 aux_cur[4], aux_nxt[4];
      auto load_aux = [&](int c, uint4 (&dst)[4]) {
        if (HAS_AUX) {
          const int col0 = n0 + c * 32;
#pragma unroll
          for (int g = 0; g < 4; ++g)
            if (row_ok && col0 + g * 8 < p.N)
              dst[g] = *reinterpret_cast<const uint4*>(auxp + row_off + col0 + g * 8);
        }
      };
      load_aux(0, aux_cur);
      uint8_t* buf0 = nullptr;
      uint8_t* buf1 = nullptr;
      float dot = 0.f;                       // B200_EPI_ROWDOT: running sum over the columns of one head
      const int rd_mask = (EPI == B200_EPI_ROWDOT) ? (p.rd_hd >> 5) - 1 : 0;    // 64 / 128 columns = 2 / 4 chunks
      const int rd_shift = (p.rd_hd == 128) ? 7 : 6;
      float* rd_dst = nullptr;
      if (EPI == B200_EPI_ROWDOT) {
        const int r = row_ok ? row : 0;
        const int b = r / p.rd_T, t = r - b * p.rd_T;
        rd_dst = p.rowdot + (static_cast<size_t>(b) * (p.N >> rd_shift)) * p.rd_T + t;   // + head * T
      }
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr_row + c * 32, r);
        if (c + 1 < BN / 32) load_aux(c + 1, aux_nxt);
        if ((c & 1) == 0) {
          // claim staging for this 64-column chunk: a tile is free once the store issued from it has
          // finished READING shared memory
          if (DUAL) {
            if (lane == 0) tma_store_wait_read<0>();
            buf0 = stg_w;
            buf1 = stg_w + 4096;
          } else {
            if (lane == 0) tma_store_wait_read<1>();
            buf0 = stg_w + (nst & 1) * 4096;
            ++nst;
          }
          __syncwarp();
        }
        tmem_ld_wait();
        uint8_t* row0 = buf0 + lane * 128;
        uint8_t* row1 = DUAL ? buf1 + lane * 128 : nullptr;
#pragma unroll
        for (int g = 0; g < 4; ++g) {  // 8 columns (16 B of bf16) per group
          const int sw = ((((c & 1) * 4 + g) ^ (lane & 7)) << 4);   // 128B-swizzled 16-byte slot
          uint4 o;
          const int col = n0 + c * 32 + g * 8;
          float v[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[g * 8 + j]);
          if (HAS_BIAS) {
            const float4 b0 = *reinterpret_cast<const float4*>(bias_s + c * 32 + g * 8);
            const float4 b1 = *reinterpret_cast<const float4*>(bias_s + c * 32 + g * 8 + 4);
            v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
            v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
          }
          if (DUAL) {   // pre-activation (or the GELU derivative) -> D, activation -> D2
            float ga[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float dg;
              gelu_erf_both(v[j], ga[j], dg);
              if (EPI == B200_EPI_BIAS_GELU_GRAD) v[j] = dg;
            }
            o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
            o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
            *reinterpret_cast<uint4*>(row0 + sw) = o;
            o.x = pack_bf16x2(ga[0], ga[1]); o.y = pack_bf16x2(ga[2], ga[3]);
            o.z = pack_bf16x2(ga[4], ga[5]); o.w = pack_bf16x2(ga[6], ga[7]);
            *reinterpret_cast<uint4*>(row1 + sw) = o;
          } else {
            if (EPI == B200_EPI_BIAS_RESID && p.drop_threshold) {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const uint64_t idx = static_cast<uint64_t>(row) * p.N + (col + j);
                v[j] = dropout_keep(seed, idx, p.drop_threshold) ? v[j] * p.drop_scale : 0.0f;
              }
            }
            if (HAS_AUX) {
              const bool ok = row_ok && col < p.N;     // out-of-range aux was never loaded
              const uint32_t xw[4] = {aux_cur[g].x, aux_cur[g].y, aux_cur[g].z, aux_cur[g].w};
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float2 f = ok ? unpack_bf16x2(xw[j]) : make_float2(0.f, 0.f);
                if (EPI == B200_EPI_DGELU) {
                  v[2 * j] *= gelu_erf_grad(f.x);
                  v[2 * j + 1] *= gelu_erf_grad(f.y);
                } else if (EPI == B200_EPI_MUL) {
                  v[2 * j] *= f.x;
                  v[2 * j + 1] *= f.y;
                } else if (EPI == B200_EPI_ROWDOT) {
                  dot = fmaf(v[2 * j], f.x, dot);
                  dot = fmaf(v[2 * j + 1], f.y, dot);
                } else {   // residual add / accumulate
                  v[2 * j] += f.x;
                  v[2 * j + 1] += f.y;
                }
              }
            }
            o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
            o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
            *reinterpret_cast<uint4*>(row0 + sw) = o;
          }
        }
        if (EPI == B200_EPI_ROWDOT && ((c + 1) & rd_mask) == 0) {
          const int col_end = n0 + (c + 1) * 32;             // first column after this head
          if (row_ok && col_end <= p.N) rd_dst[static_cast<size_t>((col_end >> rd_shift) - 1) * p.rd_T] = dot;
          dot = 0.f;
        }
        if (c & 1) {
          fence_proxy_async();     // generic-proxy smem writes -> visible to the TMA engine
          __syncwarp();
          if (lane == 0) {
            const int col64 = n0 + (c >> 1) * 64;
            if (col64 < p.N) {
              tma_store_2d(&map_d, buf0, col64, m0 + quarter * 32);
              tma_store_commit();
              if (DUAL) {
                tma_store_2d(&map_d2, buf1, col64, m0 + quarter * 32);
                tma_store_commit();
              }
            }
          }
        }
        if (HAS_AUX) {
#pragma unroll
          for (int g = 0; g < 4; ++g) aux_cur[g] = aux_nxt[g];
        }
      }
    };

    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      const int m0 = (tile % p.num_m_tiles) * (GEMM_BM * CL) + static_cast<int>(crank) * GEMM_BM;
      const int nt = tile / p.num_m_tiles;
      const int n0 = nt * BN;
      float* bias_s = s_bias + (tile_no & 1) * BN;
      ++tile_no;
      if (p.epi == B200_EPI_BIAS || p.epi == B200_EPI_BIAS_GELU || p.epi == B200_EPI_BIAS_RESID ||
          p.epi == B200_EPI_BIAS_GELU_GRAD) {
        // stage this tile's bias slice once (overlaps the main loop still running for this tile)
        for (int i = et; i < BN; i += 128)
          bias_s[i] = (n0 + i < p.N) ? __bfloat162float(p.bias[n0 + i]) : 0.f;
        asm volatile("bar.sync 3, 128;" ::: "memory");
      }
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      switch (p.epi) {
        case B200_EPI_BIAS: run_tile(std::integral_constant<int, B200_EPI_BIAS>{}, m0, n0, bias_s); break;
        case B200_EPI_BIAS_GELU: run_tile(std::integral_constant<int, B200_EPI_BIAS_GELU>{}, m0, n0, bias_s); break;
        case B200_EPI_BIAS_RESID: run_tile(std::integral_constant<int, B200_EPI_BIAS_RESID>{}, m0, n0, bias_s); break;
        case B200_EPI_DGELU: run_tile(std::integral_constant<int, B200_EPI_DGELU>{}, m0, n0, bias_s); break;
        case B200_EPI_ACCUM: run_tile(std::integral_constant<int, B200_EPI_ACCUM>{}, m0, n0, bias_s); break;
        case B200_EPI_ROWDOT: run_tile(std::integral_constant<int, B200_EPI_ROWDOT>{}, m0, n0, bias_s); break;
        case B200_EPI_BIAS_GELU_GRAD: run_tile(std::integral_constant<int, B200_EPI_BIAS_GELU_GRAD>{}, m0, n0, bias_s); break;
        case B200_EPI_MUL: run_tile(std::integral_constant<int, B200_EPI_MUL>{}, m0, n0, bias_s); break;
        default: run_tile(std::integral_constant<int, B200_EPI_NONE>{}, m0, n0, bias_s); break;
      }
      if (bias_grad && nt == 0) {
        // row sums of A (= the bias gradient): any of the 16 identical columns of the row-sum accumulator
        uint32_t rs;
        tmem_ld_32x32b_x1(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + Cfg::kBiasCol, rs);
        tmem_ld_wait();
        const int row = m0 + quarter * 32 + lane;
        if (row < p.M) {
          float s = __uint_as_float(rs);
          if (p.dbias_accum) s += __bfloat162float(p.dbias[row]);
          p.dbias[row] = __float2bfloat16(s);
        }
      }
      // all tcgen05.ld of this warp are complete (wait::ld above): release the accumulator
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CTAS == 2) mbar_arrive_leader(&tempty_bar[acc]); else mbar_arrive(&tempty_bar[acc]);
      }
      if (++acc == nacc) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (lane == 0) tma_store_wait<0>();   // staging must stay valid until the last store has drained
  }

  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync_all();   // no CTA leaves (or frees TMEM) while a cluster peer may still touch it
  if (warp == 1) {
    tc_fence_after();
    if (CTAS == 2) tmem_dealloc_2cta(tmem_base, Cfg::kTmemCols); else tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ----------------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) !=
            cudaSuccess ||
        qres != cudaDriverEntryPointSuccess) {
      return nullptr;
    }
    fn = reinterpret_cast<PFN_encodeTiled>(ptr);
  }
  return fn;
}

// Tensor maps depend only on (address, shape, leading dimension, box): the engine calls the same GEMMs on the
// same buffers every step, so encoded maps are kept in a small direct-mapped cache instead of paying four
// driver calls per launch (eager mode, the phase profiler and the tests do not replay a CUDA graph).
struct TmapKey {
  const void* ptr;
  uint64_t rows, cols, ld;
  uint32_t box_rows, box_cols;
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows &&
           box_cols == o.box_cols;
  }
};
struct TmapSlot {
  TmapKey key;
  bool valid;
  CUtensorMap map;
};
constexpr int kTmapCacheSlots = 2048;
static TmapSlot g_tmap_cache[kTmapCacheSlots];
static std::mutex g_tmap_mutex;
static unsigned long long g_tmap_hits = 0, g_tmap_misses = 0;

// 2-D bf16 tensor map over a row-major [rows, cols] matrix with leading dimension ld (elements),
// box = [box_rows, box_cols], 128-byte swizzle (box_cols * 2 bytes must be <= 128).
int make_tmap_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_rows, uint32_t box_cols) {
  const TmapKey key{ptr, rows, cols, ld, box_rows, box_cols};
  uint64_t h = reinterpret_cast<uint64_t>(ptr) * 0x9E3779B97F4A7C15ull;
  h ^= (rows * 0xC2B2AE3D27D4EB4Full) ^ (cols << 17) ^ (ld << 31) ^ (static_cast<uint64_t>(box_rows) << 7) ^ box_cols;
  h ^= h >> 29;
  TmapSlot& slot = g_tmap_cache[h % kTmapCacheSlots];
  {
    std::lock_guard<std::mutex> lock(g_tmap_mutex);
    if (slot.valid && slot.key == key) {
      memcpy(map, &slot.map, sizeof(CUtensorMap));
      ++g_tmap_hits;
      return 0;
    }
  }
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return -1;
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * sizeof(__nv_bfloat16)};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return -static_cast<int>(r) - 1000;
  std::lock_guard<std::mutex> lock(g_tmap_mutex);
  slot.key = key;
  memcpy(&slot.map, map, sizeof(CUtensorMap));
  slot.valid = true;
  ++g_tmap_misses;
  return 0;
}

int device_sm_count() {
  static int sms[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!sms[dev]) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
  return sms[dev];
}

static int g_gemm_pdl = 1;   // b200_set_gemm_pdl(0) turns programmatic dependent launch off

template <int BN, int CTAS>
static cudaError_t gemm_func_attrs() {
  using Cfg = GemmCfg<BN, CTAS>;
  static bool attr_done[64] = {false};   // one flag set per template instantiation
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_done[dev & 63]) {
    cudaError_t e = cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN, CTAS>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    if (e != cudaSuccess) return e;
    attr_done[dev & 63] = true;
  }
  return cudaSuccess;
}

// Clusters of more than 8 CTAs (8 CTA pairs sharing one B tile) are "non-portable": opt in the first time one
// is asked for, so that ordinary launches keep the default scheduling.
template <int BN, int CTAS>
static cudaError_t gemm_allow_big_clusters() {
  static bool done[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!done[dev & 63]) {
    cudaError_t e = cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN, CTAS>,
                                         cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return e;
    done[dev & 63] = true;
  }
  return cudaSuccess;
}

// How many clusters of (units * CTAS) CTAs of this kernel can be resident at once (one CTA per SM, every CTA
// of a cluster inside one GPC).  Cached per device and configuration.
template <int BN, int CTAS>
static int max_resident_clusters(int units) {
  using Cfg = GemmCfg<BN, CTAS>;
  static int cache[64][GEMM_MAX_UNITS + 1] = {{0}};
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (units < 1 || units > GEMM_MAX_UNITS) return 0;
  if (cache[dev][units]) return cache[dev][units] < 0 ? 0 : cache[dev][units];
  if (gemm_func_attrs<BN, CTAS>() != cudaSuccess) return 0;
  if (units * CTAS > 8 && gemm_allow_big_clusters<BN, CTAS>() != cudaSuccess) return 0;
  int n = 0;
  if (units == 1) {
    n = device_sm_count() / CTAS;
  } else {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(units * CTAS * 8);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = Cfg::kSmemBytes;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = units * CTAS;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (cudaOccupancyMaxActiveClusters(&n, gemm_bf16_tcgen05_kernel<BN, CTAS>, &cfg) != cudaSuccess) {
      cudaGetLastError();
      n = 0;
    }
  }
  cache[dev][units] = n > 0 ? n : -1;
  return n;
}

static int max_resident_clusters_rt(int BN, int CTAS, int units) {
  if (CTAS == 2) return BN == 256 ? max_resident_clusters<256, 2>(units) : max_resident_clusters<128, 2>(units);
  return BN == 256 ? max_resident_clusters<256, 1>(units) : max_resident_clusters<128, 1>(units);
}

template <int BN, int CTAS>
static int launch_gemm(const CUtensorMap& ma, const BMaps& mb, const CUtensorMap& md,
                       const CUtensorMap& md2, GemmParams& p, int max_ctas, cudaStream_t stream) {
  using Cfg = GemmCfg<BN, CTAS>;
  cudaError_t e = gemm_func_attrs<BN, CTAS>();
  if (e != cudaSuccess) return static_cast<int>(e);
  const int CL = p.units * CTAS;
  p.num_m_tiles = (p.M + GEMM_BM * CL - 1) / (GEMM_BM * CL);
  p.num_n_tiles = (p.N + BN - 1) / BN;
  const int tiles = p.num_m_tiles * p.num_n_tiles;
  int clusters = max_resident_clusters<BN, CTAS>(p.units);
  if (clusters < 1) return -9;
  if (max_ctas > 0 && max_ctas / CL < clusters) clusters = max_ctas / CL > 0 ? max_ctas / CL : 1;
  if (tiles < clusters) clusters = tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(clusters * CL);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // PDL: see pdl_wait() in the kernel
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_gemm_pdl ? 2 : 1;
  return static_cast<int>(cudaLaunchKernelEx(&cfg, gemm_bf16_tcgen05_kernel<BN, CTAS>, ma, mb, md, md2, p));
}

}  // namespace b200

using namespace b200;

static int gemm_impl(const void* A, const void* B, void* D, int M, int N, int K, int lda, int ldb,
                     int ldd, int a_mn, int b_mn, int epi, const void* bias, const void* aux,
                     void* D2, float drop_p, unsigned int drop_seed, int bn_hint, int max_ctas,
                     void* stream, const B200GemmEx* ex) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((N % 8) || (K % 8) || (lda % 8) || (ldb % 8) || (ldd % 8)) return -2;
  if (a_mn && (M % 8)) return -2;
  const int b_ws = (ex && ex->b_ws > 1) ? ex->b_ws : 1;
  uint32_t b_row0 = 0, b_rpo = 0;
  if (b_ws > 1) {
    // the owners' shards are viewed as [rows_per_owner, ldb] matrices: offset and shard size must be whole rows
    if (b_ws > GEMM_MAX_OWNERS || !ex->b_shards || ex->b_shard_elems == 0 || (ex->b_shard_elems % ldb) ||
        (ex->b_flat_off % ldb))
      return -2;
    b_row0 = static_cast<uint32_t>(ex->b_flat_off / ldb);
    b_rpo = static_cast<uint32_t>(ex->b_shard_elems / ldb);
  }
  const int b_rows = b_mn ? K : N;      // rows of the stored weight matrix
  // bn_hint: 0 = auto, 128 / 256 = single-CTA tile width, 2128 / 2256 = CTA-pair (cta_group::2) mode
  int BN = bn_hint % 1000, CTAS = bn_hint >= 2000 ? 2 : 1;
  // TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's
  // memory hung the GPU.  A partitioned B (tensor maps over peer memory) is therefore always fetched unicast.
  if (b_ws > 1 && ex && ex->mc_units > 1) return -2;
  int want_units = (b_ws > 1) ? 1 : (ex ? ex->mc_units : 0);   // 0 = auto = off
  // Units per cluster for a (bn, ctas) candidate: as many consecutive M-tiles as there are (<= 8, power of two),
  // every TMA box of B a multiple of 8 rows, and no box may straddle two owners.
  auto units_for = [&](int bn, int ctas) {
    int want = want_units > 0 ? want_units : 1;
    int u = 1;
    while (u * 2 <= want && u * 2 <= GEMM_MAX_UNITS && u * 2 * 128 * ctas <= ((M + 127) / 128) * 128) u *= 2;
    for (; u >= 1; u >>= 1) {
      const int box = b_mn ? GEMM_BK / u : (bn / ctas) / u;
      if (box < 8 || (box % 8)) continue;
      if (b_ws > 1 && ((b_row0 % box) || (b_rpo % box))) continue;
      if (u > 1 && max_resident_clusters_rt(bn, ctas, u) < 1) continue;
      return u;
    }
    return 0;
  };
  int units = 1;
  if (BN != 128 && BN != 256) {
    // Pick (pair mode, tile width) by a wave model: rounds of tiles over the scheduling units times the
    // per-tile main-loop time, with the shared-memory-bandwidth efficiency of each shape
    // (128x128: operand reads + TMA fills = 2 x 128 B/clk -> ~0.5; 128x256 and pair 256x128: ~0.68;
    //  pair 256x256: ~0.95 - see profiles/README.md).
    struct Cand { int bn, ctas; double eff; };
    const Cand cands[4] = {{128, 1, 0.50}, {256, 1, 0.68}, {128, 2, 0.68}, {256, 2, 0.95}};
    double best = 1e30;
    BN = 0;
    for (const Cand& c : cands) {
      if (c.bn > 128 && N < 256) continue;
      if (c.ctas == 2 && M <= 128) continue;
      const int u = units_for(c.bn, c.ctas);
      if (u < 1) continue;
      const long mt = (M + 128 * c.ctas * u - 1) / (128 * c.ctas * u);
      const long nt = (N + c.bn - 1) / c.bn;
      const long slots = max_resident_clusters_rt(c.bn, c.ctas, u);
      if (slots < 1) continue;
      const long rounds = (mt * nt + slots - 1) / slots;
      double cost = rounds * (c.bn / 128.0) / c.eff + 0.15 * rounds;   // + per-tile epilogue/issue overhead
      // remote B: every unit of M that is NOT covered by the cluster fetches the weight over NVLink again
      if (b_ws > 1) cost *= 1.0 + 0.25 * (mt - 1);
      if (cost < best - 1e-9) {
        best = cost;
        BN = c.bn;
        CTAS = c.ctas;
        units = u;
      }
    }
    if (!BN) return -9;
  } else {
    units = units_for(BN, CTAS);
    if (units < 1) return -9;
  }
  const int b_box_rows = b_mn ? GEMM_BK / units : (BN / CTAS) / units;
  if (b_ws > 1 && ((b_row0 % b_box_rows) || (b_rpo % b_box_rows))) return -9;
  (void)b_rows;

  CUtensorMap ma;
  BMaps mb;
  int rc;
  if (!a_mn)
    rc = make_tmap_2d_bf16(&ma, A, M, K, lda, GEMM_BM, 64);
  else
    rc = make_tmap_2d_bf16(&ma, A, K, M, lda, GEMM_BK, 64);
  if (rc) return rc;
  if (b_ws > 1) {
    // one map per owner over its whole shard [rows_per_owner, ldb]; the kernel picks the owner per box
    for (int o = 0; o < b_ws; ++o) {
      rc = make_tmap_2d_bf16(&mb.m[o], ex->b_shards[o], b_rpo, ldb, ldb, b_box_rows, 64);
      if (rc) return rc;
    }
    for (int o = b_ws; o < GEMM_MAX_OWNERS; ++o) mb.m[o] = mb.m[0];
  } else {
    if (!b_mn)
      rc = make_tmap_2d_bf16(&mb.m[0], B, N, K, ldb, b_box_rows, 64);
    else
      rc = make_tmap_2d_bf16(&mb.m[0], B, K, N, ldb, b_box_rows, 64);
    if (rc) return rc;
    for (int o = 1; o < GEMM_MAX_OWNERS; ++o) mb.m[o] = mb.m[0];
  }
  // outputs: 32-row x 64-column boxes, one per epilogue warp and 64-column chunk
  CUtensorMap md, md2;
  rc = make_tmap_2d_bf16(&md, D, M, N, ldd, 32, 64);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&md2, D2 ? D2 : D, M, N, ldd, 32, 64);
  if (rc) return rc;

  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = M;
  p.N = N;
  p.K = K;
  p.a_mn = a_mn;
  p.b_mn = b_mn;
  p.epi = epi;
  p.ldd = ldd;
  p.D = static_cast<__nv_bfloat16*>(D);
  p.D2 = static_cast<__nv_bfloat16*>(D2);
  p.bias = static_cast<const __nv_bfloat16*>(bias);
  p.aux = static_cast<const __nv_bfloat16*>(aux);
  if (drop_p > 0.f && drop_p < 1.f) {
    p.drop_scale = 1.0f / (1.0f - drop_p);
    p.drop_threshold = static_cast<uint32_t>(static_cast<double>(drop_p) * 4294967296.0);
  } else {
    p.drop_scale = 1.0f;
    p.drop_threshold = 0;
  }
  p.drop_seed = drop_seed;
  p.seed_ptr = step_seed_ptr();
  p.units = units;
  p.b_ws = b_ws;
  p.b_row0 = b_row0;
  p.b_rows_per_owner = b_rpo ? b_rpo : 1;
  if (ex) {
    p.dbias = static_cast<__nv_bfloat16*>(ex->dbias);
    p.dbias_accum = ex->dbias_accum;
    p.rowdot = ex->rowdot;
    p.rd_T = ex->rd_T;
    p.rd_hd = ex->rd_hd;
  }
  if (epi == B200_EPI_ROWDOT) {
    if (!p.rowdot || !p.aux || p.rd_T <= 0 || (p.rd_hd != 64 && p.rd_hd != 128) || (N % p.rd_hd) || (M % p.rd_T))
      return -2;
  }
  if (p.dbias && epi != B200_EPI_NONE && epi != B200_EPI_ACCUM) return -2;   // the ones tile uses the bias slice
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (CTAS == 2)
    return (BN == 256) ? launch_gemm<256, 2>(ma, mb, md, md2, p, max_ctas, s) : launch_gemm<128, 2>(ma, mb, md, md2, p, max_ctas, s);
  return (BN == 256) ? launch_gemm<256, 1>(ma, mb, md, md2, p, max_ctas, s) : launch_gemm<128, 1>(ma, mb, md, md2, p, max_ctas, s);
}

extern "C" int b200_set_gemm_pdl(int on) {
  g_gemm_pdl = on ? 1 : 0;
  return 0;
}

extern "C" int b200_tmap_cache_stats(unsigned long long* hits, unsigned long long* misses) {
  std::lock_guard<std::mutex> lock(g_tmap_mutex);
  if (hits) *hits = g_tmap_hits;
  if (misses) *misses = g_tmap_misses;
  return 0;
}

extern "C" int b200_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int lda,
                              int ldb, int ldd, int a_mn, int b_mn, int epi, const void* bias,
                              const void* aux, void* D2, float drop_p, unsigned int drop_seed,
                              int bn_hint, int max_ctas, void* stream) {
  return gemm_impl(A, B, D, M, N, K, lda, ldb, ldd, a_mn, b_mn, epi, bias, aux, D2, drop_p, drop_seed,
                   bn_hint, max_ctas, stream, nullptr);
}

extern "C" int b200_gemm_bf16_ex(const void* A, const void* B, void* D, int M, int N, int K, int lda,
                                 int ldb, int ldd, int a_mn, int b_mn, int epi, const void* bias,
                                 const void* aux, void* D2, float drop_p, unsigned int drop_seed,
                                 int bn_hint, int max_ctas, void* stream, const B200GemmEx* ex) {
  return gemm_impl(A, B, D, M, N, K, lda, ldb, ldd, a_mn, b_mn, epi, bias, aux, D2, drop_p, drop_seed,
                   bn_hint, max_ctas, stream, ex);
}
