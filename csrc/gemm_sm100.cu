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
//   warp 1      MMA issuer    (one elected lane: tcgen05.mma cta_group::1, M=128, N=BN, K=16)
//   warps 2..5  epilogue      (tcgen05.ld 32x32b -> registers -> fused epilogue -> global)
// Two accumulator stages in TMEM (2*BN columns) let the epilogue of tile i overlap the main
// loop of tile i+1.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <type_traits>

#include "b200_internal.h"
#include "sm100_ptx.cuh"

namespace b200 {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_THREADS = 192;
constexpr int GEMM_GATHER_THREADS = 64;    // 2 extra warps in gather mode (256 threads total: no register cap below 255)

__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t* addr) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu_u32(uint32_t* addr, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 ld_volatile_v4(const uint4* addr) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(addr)
               : "memory");
  return v;
}
// generic-proxy global writes/reads <-> async-proxy (TMA) accesses of the same global memory
__device__ __forceinline__ void fence_proxy_async_global() {
  asm volatile("fence.proxy.async.global;" ::: "memory");
}

template <int BN, int CTAS>
struct GemmCfg {
  static constexpr int kBNLocal = BN / CTAS;             // B rows held by each CTA of the pair
  static constexpr int kABytes = GEMM_BM * GEMM_BK * 2;  // 16 KiB
  static constexpr int kBBytes = kBNLocal * GEMM_BK * 2;
  static constexpr int kStagingBytes = 32768;            // TMA-store staging of the epilogue
  static constexpr int kBarBytes = 256 + 2 * BN * 4;     // barriers + double-buffered fp32 bias slice
  static constexpr int kRing = 232448 - kStagingBytes - kBarBytes;   // what is left of the 227 KiB
  static constexpr int kStages = (kRing / (kABytes + kBBytes)) > 8 ? 8 : (kRing / (kABytes + kBBytes));
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTmemCols = 2 * BN;  // 256 or 512, power of two
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + kBarBytes;
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
  int num_m_tiles, num_n_tiles;
  // ---- K14: all-gather of the B operand fused into this GEMM (gather_ws == 0: off) -------------
  // B (a weight matrix, row-major [g_rows, g_cols]) is partitioned over ranks by flat offset; four
  // extra warps per CTA pull it chunk by chunk (g_chunk_rows rows each) from the owners' shards over
  // NVLink into the local staging buffer that the TMA descriptor of B points at, and publish one flag
  // per chunk; the TMA producer waits for the chunks a tile needs.  Transfer overlaps the math.
  int gather_ws;
  int g_rows, g_cols, g_chunk_rows;
  unsigned long long g_flat_off;     // flat-layout offset of B[0,0]
  unsigned long long g_shard;        // elements per rank partition
  const __nv_bfloat16* g_src[8];     // per rank: partition base MINUS rank*g_shard (flat-indexed)
  __nv_bfloat16* g_dst;              // staging buffer (== B)
  uint32_t* g_flags;                 // [chunks] ready flags
  const uint32_t* g_epoch;           // device word E, bumped after every gather-GEMM
  int g_self;                        // 1: gather B in this kernel (flags <- E+1); 0: a previous kernel prefetched it
  // prefetch job: the NEXT GEMM's weight, pulled into its staging buffer while this GEMM computes
  // (flags <- E+2).  pf_rows == 0: none.   [FSDP forward_prefetch / backward_prefetch]
  int pf_rows, pf_cols;
  unsigned long long pf_flat_off;
  __nv_bfloat16* pf_dst;
  uint32_t* pf_flags;
};

// CTAS = 1: one CTA per 128 x BN tile (cta_group::1).
// CTAS = 2: a CTA pair (2-CTA cluster) per 256 x BN tile (cta_group::2): each CTA stages its own 128 rows
//           of A and HALF of B, the leader issues M=256 MMAs that read both halves, so the shared-memory
//           traffic per MAC halves for B — at 128x128 per CTA the single-CTA kernel is shared-memory-bound
//           (operand reads + TMA fills = 2 x 128 B/clk), which caps it near 50 % of the tensor peak.
template <int BN, int CTAS>
__global__ void __launch_bounds__(GEMM_THREADS + GEMM_GATHER_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a,
                         const __grid_constant__ CUtensorMap map_b,
                         const __grid_constant__ CUtensorMap map_d,
                         const __grid_constant__ CUtensorMap map_d2, const GemmParams p) {
  using Cfg = GemmCfg<BN, CTAS>;
  constexpr int BNL = Cfg::kBNLocal;
  const uint32_t cta_rank = (CTAS == 2) ? cluster_ctarank() : 0u;
  const bool is_leader = cta_rank == 0;
  const int unit = blockIdx.x / CTAS;          // tile-scheduling unit: a CTA or a CTA pair
  const int num_units = gridDim.x / CTAS;
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

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = (p.K + GEMM_BK - 1) / GEMM_BK;
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    tma_prefetch_desc(&map_d);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
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
  tc_fence_before();
  __syncthreads();
  if (CTAS == 2) cluster_sync_all();   // the peer's barriers are initialised before anyone signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Everything above (barrier init, TMEM allocation, descriptor prefetch) may overlap the tail of the
  // previous kernel in the stream; nothing below may start before that kernel's writes are visible.
  pdl_trigger();
  pdl_wait();

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const bool gather = p.gather_ws > 0;
      // own weight: published as E+1 either by this kernel's gather warps or by the previous kernel's
      // prefetch (whose epoch was E-1).  Zero-filled flags can never satisfy the wait.
      const uint32_t g_want = gather ? *p.g_epoch + 1u : 0u;
      bool g_all_ready = false;
      auto wait_chunk = [&](int c) {
        uint32_t spins = 0;
        while (static_cast<int32_t>(ld_acquire_gpu_u32(p.g_flags + c) - g_want) < 0) {
          if (++spins > (1u << 28)) __trap();
        }
      };
      for (int tile = unit; tile < num_tiles; tile += num_units) {
        const int m0 = (tile % p.num_m_tiles) * (GEMM_BM * CTAS) + static_cast<int>(cta_rank) * GEMM_BM;
        const int n0 = (tile / p.num_m_tiles) * BN + static_cast<int>(cta_rank) * BNL;   // this CTA's B rows
        if (gather && !p.b_mn) {   // B rows = n: this CTA needs the chunks covering [n0, n0+BNL)
          const int n1 = (n0 + BNL < p.g_rows ? n0 + BNL : p.g_rows) - 1;
          for (int c = n0 / p.g_chunk_rows; c <= n1 / p.g_chunk_rows; ++c) wait_chunk(c);
          fence_proxy_async_global();
        }
        for (int kb = 0; kb < num_kb; ++kb) {
          if (gather && p.b_mn && !g_all_ready) {   // B rows = k: one chunk per k-block, first tile only
            const int r0 = kb * GEMM_BK;
            const int r1 = (r0 + GEMM_BK < p.g_rows ? r0 + GEMM_BK : p.g_rows) - 1;
            for (int c = r0 / p.g_chunk_rows; c <= r1 / p.g_chunk_rows; ++c) wait_chunk(c);
            fence_proxy_async_global();
          }
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = tiles + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          // pair mode: all bytes (of both CTAs) are credited to the leader's barrier; only it arrives
          if (is_leader) mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes * CTAS);
          const int k0 = kb * GEMM_BK;
          auto load = [&](const CUtensorMap* m, void* dst, int c0, int c1) {
            if (CTAS == 2) tma_load_2d_2cta(m, &full_bar[stage], dst, c0, c1);
            else tma_load_2d(m, &full_bar[stage], dst, c0, c1);
          };
          if (!p.a_mn) {
            load(&map_a, sa, k0, m0);
          } else {
#pragma unroll
            for (int c = 0; c < GEMM_BM / 64; ++c) load(&map_a, sa + c * (GEMM_BK * 128), m0 + 64 * c, k0);
          }
          if (!p.b_mn) {
            load(&map_b, sb, k0, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BNL / 64; ++c) load(&map_b, sb + c * (GEMM_BK * 128), n0 + 64 * c, k0);
          }
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        g_all_ready = true;
      }
    }
  } else if (warp >= 6) {
    // ================================ K14 gather warps =============================
    if (p.gather_ws > 0) {
      const int gt = threadIdx.x - GEMM_THREADS;   // 0..63
      const uint32_t E = *p.g_epoch;
      auto gather_job = [&](unsigned long long flat_off, int rows, int cols, __nv_bfloat16* out,
                            uint32_t* flags, uint32_t publish) {
        const size_t chunk_elems = static_cast<size_t>(p.g_chunk_rows) * cols;
        const size_t total = static_cast<size_t>(rows) * cols;
        const int nchunks = (rows + p.g_chunk_rows - 1) / p.g_chunk_rows;
        for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
          size_t e = c * chunk_elems;
          const size_t e1 = (e + chunk_elems < total) ? e + chunk_elems : total;
          while (e < e1) {
            const size_t flat = flat_off + e;
            const int owner = static_cast<int>(flat / p.g_shard);
            const size_t seg_end = static_cast<size_t>(owner + 1) * p.g_shard - flat_off;
            const size_t e_stop = seg_end < e1 ? seg_end : e1;
            const uint4* src = reinterpret_cast<const uint4*>(p.g_src[owner] + flat);
            uint4* dst = reinterpret_cast<uint4*>(out + e);
            const size_t nvec = (e_stop - e) >> 3;
            // 8 independent 16-byte loads per thread in flight: NVLink latency is hidden by depth
            for (size_t i = gt; i < nvec; i += 8 * GEMM_GATHER_THREADS) {
              uint4 v[8];
#pragma unroll
              for (int u = 0; u < 8; ++u)
                if (i + u * GEMM_GATHER_THREADS < nvec) v[u] = ld_volatile_v4(src + i + u * GEMM_GATHER_THREADS);
#pragma unroll
              for (int u = 0; u < 8; ++u)
                if (i + u * GEMM_GATHER_THREADS < nvec) dst[i + u * GEMM_GATHER_THREADS] = v[u];
            }
            e = e_stop;
          }
          __threadfence();
          fence_proxy_async_global();
          asm volatile("bar.sync 2, 64;" ::: "memory");
          if (gt == 0) st_release_gpu_u32(flags + c, publish);
        }
      };
      if (p.g_self) gather_job(p.g_flat_off, p.g_rows, p.g_cols, p.g_dst, p.g_flags, E + 1u);  // urgent: this GEMM waits on it
      if (p.pf_rows > 0) gather_job(p.pf_flat_off, p.pf_rows, p.pf_cols, p.pf_dst, p.pf_flags, E + 2u);
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==================================
    if (lane == 0 && is_leader) {
      const uint32_t idesc = umma_idesc_bf16(GEMM_BM * CTAS, BN, p.a_mn != 0, p.b_mn != 0);
      // K-major: +32 B per 16-element K step inside the 128 B swizzle row.
      // MN-major: +16 k-rows * 128 B = 2048 B per K step; MN chunks are BK*128 B apart.
      const uint32_t a_kstep = p.a_mn ? 2048u : 32u;
      const uint32_t b_kstep = p.b_mn ? 2048u : 32u;
      const uint32_t a_lbo = p.a_mn ? GEMM_BK * 128u : 16u;
      const uint32_t b_lbo = p.b_mn ? GEMM_BK * 128u : 16u;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = unit; tile < num_tiles; tile += num_units) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(tiles + stage * Cfg::kStageBytes);
          const uint32_t sb = sa + Cfg::kABytes;
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k) {
            const uint64_t adesc = umma_smem_desc(sa + k * a_kstep, a_lbo, 1024);
            const uint64_t bdesc = umma_smem_desc(sb + k * b_kstep, b_lbo, 1024);
            if (CTAS == 2) umma_bf16_ss_2cta(tmem_d, adesc, bdesc, idesc, (kb | k) ? 1u : 0u);
            else umma_bf16_ss(tmem_d, adesc, bdesc, idesc, (kb | k) ? 1u : 0u);
          }
          // smem slot reusable (in both CTAs of a pair) once these MMAs retire
          if (CTAS == 2) umma_commit_2cta(&empty_bar[stage]); else umma_commit(&empty_bar[stage]);
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        // accumulator complete -> epilogue (of both CTAs)
        if (CTAS == 2) umma_commit_2cta(&tfull_bar[acc]); else umma_commit(&tfull_bar[acc]);
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp < 6) {
    // ================================ epilogue ====================================
    const int quarter = warp & 3;  // TMEM lane quarter this warp may access
    const int et = threadIdx.x - 64;   // 0..127 among the epilogue threads
    const uint32_t seed = p.drop_seed ^ ((p.drop_threshold && p.seed_ptr) ? __ldg(p.seed_ptr) : 0u);
    float* s_bias = reinterpret_cast<float*>(tmem_slot + 4);   // [2][BN] floats, double-buffered per tile
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
      constexpr bool HAS_BIAS = (EPI == B200_EPI_BIAS || EPI == B200_EPI_BIAS_GELU || EPI == B200_EPI_BIAS_RESID);
      constexpr bool HAS_AUX = (EPI == B200_EPI_BIAS_RESID || EPI == B200_EPI_DGELU || EPI == B200_EPI_ACCUM);
      constexpr bool DUAL = (EPI == B200_EPI_BIAS_GELU);
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
          if (DUAL) {   // pre-activation -> D, activation -> D2
            o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
            o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
            *reinterpret_cast<uint4*>(row0 + sw) = o;
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = gelu_erf(v[j]);
            o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
            o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
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

    for (int tile = unit; tile < num_tiles; tile += num_units) {
      const int m0 = (tile % p.num_m_tiles) * (GEMM_BM * CTAS) + static_cast<int>(cta_rank) * GEMM_BM;
      const int n0 = (tile / p.num_m_tiles) * BN;
      float* bias_s = s_bias + (tile_no & 1) * BN;
      ++tile_no;
      if (p.epi == B200_EPI_BIAS || p.epi == B200_EPI_BIAS_GELU || p.epi == B200_EPI_BIAS_RESID) {
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
        default: run_tile(std::integral_constant<int, B200_EPI_NONE>{}, m0, n0, bias_s); break;
      }
      // all tcgen05.ld of this warp are complete (wait::ld above): release the accumulator
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CTAS == 2) mbar_arrive_leader(&tempty_bar[acc]); else mbar_arrive(&tempty_bar[acc]);
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (lane == 0) tma_store_wait<0>();   // staging must stay valid until the last store has drained
  }

  tc_fence_before();
  __syncthreads();
  if (CTAS == 2) cluster_sync_all();   // neither CTA leaves (or frees TMEM) while the pair may still touch it
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

// 2-D bf16 tensor map over a row-major [rows, cols] matrix with leading dimension ld (elements),
// box = [box_rows, box_cols], 128-byte swizzle (box_cols * 2 bytes must be <= 128).
int make_tmap_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_rows, uint32_t box_cols) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return -1;
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * sizeof(__nv_bfloat16)};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -static_cast<int>(r) - 1000;
}

int device_sm_count() {
  static int sms[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!sms[dev]) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
  return sms[dev];
}

static int g_gemm_pdl = 1;   // b200_set_option("gemm_pdl", 0) turns programmatic dependent launch off

template <int BN, int CTAS>
static int launch_gemm(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& md,
                       const CUtensorMap& md2, GemmParams& p, int max_ctas, cudaStream_t stream) {
  using Cfg = GemmCfg<BN, CTAS>;
  static bool attr_done[64] = {false};   // one flag set per template instantiation
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_done[dev & 63]) {
    cudaError_t e = cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN, CTAS>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes);
    if (e != cudaSuccess) return static_cast<int>(e);
    attr_done[dev & 63] = true;
  }
  p.num_m_tiles = (p.M + GEMM_BM * CTAS - 1) / (GEMM_BM * CTAS);
  p.num_n_tiles = (p.N + BN - 1) / BN;
  const int tiles = p.num_m_tiles * p.num_n_tiles;
  int grid = device_sm_count();
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
  if (tiles * CTAS < grid) grid = tiles * CTAS;
  grid -= grid % CTAS;
  if (grid < CTAS) grid = CTAS;
  const int threads = GEMM_THREADS + (p.gather_ws > 0 ? GEMM_GATHER_THREADS : 0);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CTAS;
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

struct GatherArgs {
  int ws = 0;
  const void* const* shard_adj = nullptr;
  unsigned long long flat_off = 0, shard = 0;
  int rows = 0, cols = 0;
  void* flags = nullptr;
  void* epoch = nullptr;
  int self_gather = 1;
  unsigned long long pf_flat_off = 0;
  int pf_rows = 0, pf_cols = 0;
  void* pf_dst = nullptr;
  void* pf_flags = nullptr;
};

__global__ void gather_epoch_bump_kernel(uint32_t* w) { *w += 1; }

static int gemm_impl(const void* A, const void* B, void* D, int M, int N, int K, int lda, int ldb,
                     int ldd, int a_mn, int b_mn, int epi, const void* bias, const void* aux,
                     void* D2, float drop_p, unsigned int drop_seed, int bn_hint, int max_ctas,
                     void* stream, const GatherArgs& ga) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((N % 8) || (K % 8) || (lda % 8) || (ldb % 8) || (ldd % 8)) return -2;
  if (a_mn && (M % 8)) return -2;
  // bn_hint: 0 = auto, 128 / 256 = single-CTA tile width, 2128 / 2256 = CTA-pair (cta_group::2) mode
  int BN = bn_hint % 1000, CTAS = bn_hint >= 2000 ? 2 : 1;
  if (BN != 128 && BN != 256) {
    // Pick (pair mode, tile width) by a wave model: rounds of tiles over the scheduling units times the
    // per-tile main-loop time, with the shared-memory-bandwidth efficiency of each shape
    // (128x128: operand reads + TMA fills = 2 x 128 B/clk -> ~0.5; 128x256 and pair 256x128: ~0.68;
    //  pair 256x256: ~0.95 — see profiles/README.md).
    const int sms = device_sm_count();
    struct Cand { int bn, ctas; double eff; };
    const Cand cands[4] = {{128, 1, 0.50}, {256, 1, 0.68}, {128, 2, 0.68}, {256, 2, 0.95}};
    double best = 1e30;
    for (const Cand& c : cands) {
      if (c.bn > 128 && N < 256) continue;
      if (c.ctas == 2 && M <= 128) continue;
      const long mt = (M + 128 * c.ctas - 1) / (128 * c.ctas);
      const long nt = (N + c.bn - 1) / c.bn;
      const long units = sms / c.ctas;
      const long rounds = (mt * nt + units - 1) / units;
      const double cost = rounds * (c.bn / 128.0) / c.eff + 0.15 * rounds;   // + per-tile epilogue/issue overhead
      if (cost < best - 1e-9) {
        best = cost;
        BN = c.bn;
        CTAS = c.ctas;
      }
    }
  }
  CUtensorMap ma, mb;
  int rc;
  if (!a_mn)
    rc = make_tmap_2d_bf16(&ma, A, M, K, lda, GEMM_BM, 64);
  else
    rc = make_tmap_2d_bf16(&ma, A, K, M, lda, GEMM_BK, 64);
  if (rc) return rc;
  if (!b_mn)
    rc = make_tmap_2d_bf16(&mb, B, N, K, ldb, BN / CTAS, 64);
  else
    rc = make_tmap_2d_bf16(&mb, B, K, N, ldb, GEMM_BK, 64);
  if (rc) return rc;
  // outputs: 32-row x 64-column boxes, one per epilogue warp and 64-column chunk
  CUtensorMap md, md2;
  rc = make_tmap_2d_bf16(&md, D, M, N, ldd, 32, 64);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&md2, D2 ? D2 : D, M, N, ldd, 32, 64);
  if (rc) return rc;

  GemmParams p;
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
  p.gather_ws = ga.ws;
  if (ga.ws > 0) {
    if (ga.ws > 8 || ldb != ga.cols || (ga.cols % 8) || (ga.flat_off % 8) || (ga.shard % 8)) return -2;
    p.g_rows = ga.rows;
    p.g_cols = ga.cols;
    p.g_chunk_rows = 32;   // small chunks: every CTA's gather warps get work, flags stay fine-grained
    p.g_flat_off = ga.flat_off;
    p.g_shard = ga.shard;
    for (int r = 0; r < 8; ++r)
      p.g_src[r] = r < ga.ws ? static_cast<const __nv_bfloat16*>(ga.shard_adj[r]) : nullptr;
    p.g_dst = static_cast<__nv_bfloat16*>(const_cast<void*>(B));
    p.g_flags = static_cast<uint32_t*>(ga.flags);
    p.g_epoch = static_cast<const uint32_t*>(ga.epoch);
    p.g_self = ga.self_gather;
    p.pf_rows = ga.pf_rows;
    p.pf_cols = ga.pf_cols;
    p.pf_flat_off = ga.pf_flat_off;
    p.pf_dst = static_cast<__nv_bfloat16*>(ga.pf_dst);
    p.pf_flags = static_cast<uint32_t*>(ga.pf_flags);
    if (ga.pf_rows > 0 && ((ga.pf_cols % 8) || (ga.pf_flat_off % 8) || (ga.pf_rows + 31) / 32 > 1024)) return -2;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int rc2;
  if (CTAS == 2)
    rc2 = (BN == 256) ? launch_gemm<256, 2>(ma, mb, md, md2, p, max_ctas, s) : launch_gemm<128, 2>(ma, mb, md, md2, p, max_ctas, s);
  else
    rc2 = (BN == 256) ? launch_gemm<256, 1>(ma, mb, md, md2, p, max_ctas, s) : launch_gemm<128, 1>(ma, mb, md, md2, p, max_ctas, s);
  if (rc2 == 0 && ga.ws > 0) {
    gather_epoch_bump_kernel<<<1, 1, 0, s>>>(static_cast<uint32_t*>(ga.epoch));
    rc2 = static_cast<int>(cudaGetLastError());
  }
  return rc2;
}

extern "C" int b200_set_gemm_pdl(int on) {
  g_gemm_pdl = on ? 1 : 0;
  return 0;
}

extern "C" int b200_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int lda,
                              int ldb, int ldd, int a_mn, int b_mn, int epi, const void* bias,
                              const void* aux, void* D2, float drop_p, unsigned int drop_seed,
                              int bn_hint, int max_ctas, void* stream) {
  return gemm_impl(A, B, D, M, N, K, lda, ldb, ldd, a_mn, b_mn, epi, bias, aux, D2, drop_p, drop_seed,
                   bn_hint, max_ctas, stream, GatherArgs());
}

// K14: same GEMM, but B (a weight matrix of `rows` x `cols`, partitioned over `ws` ranks by flat
// offset) is first pulled from the owners' shards into `B_staging` by extra warps of the SAME kernel,
// chunk by chunk, while the tensor cores already work on the chunks that have arrived.
extern "C" int b200_gemm_bf16_gather(const void* A, void* B_staging, void* D, int M, int N, int K,
                                     int lda, int ldb, int ldd, int a_mn, int b_mn, int epi,
                                     const void* bias, const void* aux, void* D2, float drop_p,
                                     unsigned int drop_seed, int bn_hint, int max_ctas, void* stream,
                                     const void* const* shard_adj, int ws,
                                     unsigned long long flat_off, unsigned long long shard_elems,
                                     int rows, int cols, void* chunk_flags, void* epoch_word,
                                     int self_gather, unsigned long long pf_flat_off, int pf_rows,
                                     int pf_cols, void* pf_staging, void* pf_flags) {
  GatherArgs ga;
  ga.self_gather = self_gather;
  ga.pf_flat_off = pf_flat_off;
  ga.pf_rows = pf_rows;
  ga.pf_cols = pf_cols;
  ga.pf_dst = pf_staging;
  ga.pf_flags = pf_flags;
  ga.ws = ws;
  ga.shard_adj = shard_adj;
  ga.flat_off = flat_off;
  ga.shard = shard_elems;
  ga.rows = rows;
  ga.cols = cols;
  ga.flags = chunk_flags;
  ga.epoch = epoch_word;
  if ((rows + 31) / 32 > 1024) return -3;   // chunk_flags holds 1024 words
  return gemm_impl(A, B_staging, D, M, N, K, lda, ldb, ldd, a_mn, b_mn, epi, bias, aux, D2, drop_p,
                   drop_seed, bn_hint, max_ctas, stream, ga);
}
