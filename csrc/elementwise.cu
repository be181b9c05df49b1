// Bandwidth-bound kernels of the TinyGPT step: LayerNorm fwd/bwd, embedding gather / scatter-add,
// fused softmax-cross-entropy forward+backward, column reductions (bias / LayerNorm parameter
// gradients), flat fused AdamW, gradient sum-of-squares, dropout re-application.
//
// Reference call sites (all reached through ATen there):
//   K1 embedding+dropout   benchmarking/train_harness.py:85-89
//   K2 LayerNorm           TH:113,117,56 (ln_1 evaluated 3x per block at TH:127; once here)
//   K9 cross_entropy       TH:99-103 (ignore_index=-1, mean over valid rows)
//   K12 AdamW              TH:329 / configs/deepspeed/zero2.json:27-44 (FusedAdam, clip, WarmupLR)
// All kernels are sized for HBM3e streaming: 16-byte vector accesses, one pass over the data
// wherever the row fits in registers, fp32 math, bf16 storage.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "b200_internal.h"
#include "sm100_ptx.cuh"

namespace b200 {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 t = unpack_bf16x2(w[j]);
    f[2 * j] = t.x;
    f[2 * j + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  u.x = pack_bf16x2(f[0], f[1]);
  u.y = pack_bf16x2(f[2], f[3]);
  u.z = pack_bf16x2(f[4], f[5]);
  u.w = pack_bf16x2(f[6], f[7]);
  return u;
}

// ----------------------------------------------------------------------------------------
// LayerNorm forward: one warp per row, row cached in registers (NV uint4 per lane).
// ----------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(256)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ gamma,
                     const __nv_bfloat16* __restrict__ beta, __nv_bfloat16* __restrict__ y,
                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int M, int d,
                     float eps) {
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= M) return;
  const int nvec = d >> 3;  // uint4 vectors per row
  const uint4* xr = reinterpret_cast<const uint4*>(x + static_cast<size_t>(row) * d);
  uint4 v[NV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + i * 32;
    if (c < nvec) {
      v[i] = xr[c];
      float f[8];
      unpack8(v[i], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) s += f[j];
    }
  }
  const float mean = warp_sum(s) / d;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + i * 32;
    if (c < nvec) {
      float f[8];
      unpack8(v[i], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float t = f[j] - mean;
        q += t * t;
      }
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / d + eps);
  if (lane == 0) {
    mean_out[row] = mean;
    rstd_out[row] = rstd;
  }
  uint4* yr = reinterpret_cast<uint4*>(y + static_cast<size_t>(row) * d);
  const uint4* g4 = reinterpret_cast<const uint4*>(gamma);
  const uint4* b4 = reinterpret_cast<const uint4*>(beta);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + i * 32;
    if (c < nvec) {
      float f[8], g[8], b[8];
      unpack8(v[i], f);
      unpack8(__ldg(g4 + c), g);
      unpack8(__ldg(b4 + c), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = (f[j] - mean) * rstd * g[j] + b[j];
      yr[c] = pack8(f);
    }
  }
}

// LayerNorm backward (input gradient): dx = rstd * (g*dy - mean(g*dy) - xhat*mean(g*dy*xhat)) [+ dres]
template <int NV>
__global__ void __launch_bounds__(256)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                     const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean_in,
                     const float* __restrict__ rstd_in, const __nv_bfloat16* __restrict__ dres,
                     __nv_bfloat16* __restrict__ dx, int M, int d) {
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= M) return;
  const int nvec = d >> 3;
  const size_t off = static_cast<size_t>(row) * d;
  const uint4* xr = reinterpret_cast<const uint4*>(x + off);
  const uint4* dyr = reinterpret_cast<const uint4*>(dy + off);
  const uint4* g4 = reinterpret_cast<const uint4*>(gamma);
  const float mean = mean_in[row], rstd = rstd_in[row];
  uint4 vx[NV], vd[NV];
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + i * 32;
    if (c < nvec) {
      vx[i] = xr[c];
      vd[i] = dyr[c];
      float fx[8], fd[8], g[8];
      unpack8(vx[i], fx);
      unpack8(vd[i], fd);
      unpack8(__ldg(g4 + c), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float gd = g[j] * fd[j];
        s1 += gd;
        s2 += gd * (fx[j] - mean) * rstd;
      }
    }
  }
  s1 = warp_sum(s1) / d;
  s2 = warp_sum(s2) / d;
  uint4* dxr = reinterpret_cast<uint4*>(dx + off);
  const uint4* rr = dres ? reinterpret_cast<const uint4*>(dres + off) : nullptr;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + i * 32;
    if (c < nvec) {
      float fx[8], fd[8], g[8], o[8];
      unpack8(vx[i], fx);
      unpack8(vd[i], fd);
      unpack8(__ldg(g4 + c), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xh = (fx[j] - mean) * rstd;
        o[j] = rstd * (g[j] * fd[j] - s1 - xh * s2);
      }
      if (rr) {
        float r[8];
        unpack8(rr[c], r);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] += r[j];
      }
      dxr[c] = pack8(o);
    }
  }
}

// ----------------------------------------------------------------------------------------
// LayerNorm backward, everything in ONE pass over dy and x:
//   dx      = rstd * (g*dy - mean(g*dy) - xhat*mean(g*dy*xhat)) [+ dres]
//   dx_drop = dropout_mask(seed, row*d + col) * dx / (1-p)          (optional second output: the gradient of the
//             fc2 output of the block below, whose dropout mask the GEMM epilogue drew with the same indexing)
//   partial column sums of dy*xhat (dgamma) and dy (dbeta): one fp32 [2][d] row per CTA, folded by
//   ln_param_fold_kernel (a few microseconds, off the critical path: only the optimizer needs its result)
// A CTA owns 8*RPW consecutive rows; every warp loads its RPW rows (x, dy, dres) up front, so a CTA has all of its
// DRAM requests in flight at once.  The column sums of the CTA are accumulated in shared memory - the warps take
// turns, plain loads and stores in a fixed order, no atomics - so the result is deterministic.
// ----------------------------------------------------------------------------------------
struct LnBwdFused {
  const __nv_bfloat16* dy;
  const __nv_bfloat16* x;
  const __nv_bfloat16* gamma;
  const float* mean;
  const float* rstd;
  const __nv_bfloat16* dres;   // nullable
  __nv_bfloat16* dx;
  __nv_bfloat16* dx_drop;      // nullable
  uint32_t drop_thr;
  float drop_scale;
  uint32_t drop_seed;
  const uint32_t* seed_ptr;
  float* partials;             // [ctas][2][d] in the shared-memory layout (see s_acc), nullable
  int M, d;
};

template <int NV, int RPW>
__global__ void __launch_bounds__(256)
layernorm_bwd_fused_kernel(const LnBwdFused p) {
  pdl_wait();
  // [2][2][d/8][4] floats: (dgamma | dbeta) x (elements 0-3 | 4-7 of each 8-column vector) - 16-byte lane stride,
  // so the float4 accesses below are free of bank conflicts
  extern __shared__ __align__(16) float s_acc[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int d = p.d, nvec = d >> 3;
  const bool params = p.partials != nullptr;
  if (params) {
    for (int i = threadIdx.x; i < 2 * d; i += 256) s_acc[i] = 0.f;
  }
  const uint4* g4 = reinterpret_cast<const uint4*>(p.gamma);
  uint32_t seed = p.drop_seed;
  if (p.dx_drop && p.seed_ptr) seed ^= __ldg(p.seed_ptr);
  constexpr bool PRELOAD_RES = NV <= 8;      // d = 4096: the residual gradient would not fit in registers as well
  uint4 vx[RPW][NV], vd[RPW][NV], vr[RPW][PRELOAD_RES ? NV : 1];
  float mean[RPW], rstd[RPW];
  int rows[RPW];
  // ---- every load of this warp's rows goes out before anything is consumed ----------------------------------
#pragma unroll
  for (int r = 0; r < RPW; ++r) {
    rows[r] = blockIdx.x * (8 * RPW) + r * 8 + warp;
    const bool valid = rows[r] < p.M;
    const size_t off = static_cast<size_t>(valid ? rows[r] : 0) * d;
    const uint4* xr = reinterpret_cast<const uint4*>(p.x + off);
    const uint4* dyr = reinterpret_cast<const uint4*>(p.dy + off);
    const uint4* rr = reinterpret_cast<const uint4*>(p.dres + off);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + i * 32;
      const bool ok = valid && c < nvec;
      vx[r][i] = ok ? xr[c] : make_uint4(0, 0, 0, 0);
      vd[r][i] = ok ? dyr[c] : make_uint4(0, 0, 0, 0);
      if (PRELOAD_RES) vr[r][i] = (ok && p.dres) ? rr[c] : make_uint4(0, 0, 0, 0);
    }
    mean[r] = valid ? p.mean[rows[r]] : 0.f;
    rstd[r] = valid ? p.rstd[rows[r]] : 0.f;
  }
#pragma unroll
  for (int r = 0; r < RPW; ++r) {
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + i * 32;
      if (c < nvec) {
        float fx[8], fd[8], g[8];
        unpack8(vx[r][i], fx);
        unpack8(vd[r][i], fd);
        unpack8(__ldg(g4 + c), g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float gd = g[j] * fd[j];
          s1 += gd;
          s2 += gd * (fx[j] - mean[r]) * rstd[r];
        }
      }
    }
    s1 = warp_sum(s1) / d;
    s2 = warp_sum(s2) / d;
    if (rows[r] < p.M) {
      const size_t off = static_cast<size_t>(rows[r]) * d;
      uint4* dxr = reinterpret_cast<uint4*>(p.dx + off);
      uint4* dropr = p.dx_drop ? reinterpret_cast<uint4*>(p.dx_drop + off) : nullptr;
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = lane + i * 32;
        if (c < nvec) {
          float fx[8], fd[8], g[8], o[8], res[8];
          unpack8(vx[r][i], fx);
          unpack8(vd[r][i], fd);
          if (PRELOAD_RES) unpack8(vr[r][i], res);
          else unpack8(p.dres ? reinterpret_cast<const uint4*>(p.dres + off)[c] : make_uint4(0, 0, 0, 0), res);
          unpack8(__ldg(g4 + c), g);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float xh = (fx[j] - mean[r]) * rstd[r];
            o[j] = rstd[r] * (g[j] * fd[j] - s1 - xh * s2) + res[j];
          }
          const uint4 packed = pack8(o);
          dxr[c] = packed;
          if (dropr) {
            float q[8];
            unpack8(packed, q);      // mask the bf16 value that was stored, like a separate pass over dx would
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint64_t idx = static_cast<uint64_t>(rows[r]) * d + c * 8 + j;
              q[j] = dropout_keep(seed, idx, p.drop_thr) ? q[j] * p.drop_scale : 0.f;
            }
            dropr[c] = pack8(q);
          }
        }
      }
    }
  }
  if (!params) return;
  // ---- column sums over the CTA's rows: one turn per warp ------------------------------------------------------
#pragma unroll 1
  for (int turn = 0; turn < 8; ++turn) {
    __syncthreads();
    if (turn == warp) {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = lane + i * 32;
        if (c < nvec) {
          float cg[8], cb[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) cg[j] = cb[j] = 0.f;
#pragma unroll
          for (int r = 0; r < RPW; ++r) {
            float fx[8], fd[8];
            unpack8(vx[r][i], fx);
            unpack8(vd[r][i], fd);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              cg[j] += fd[j] * ((fx[j] - mean[r]) * rstd[r]);
              cb[j] += fd[j];
            }
          }
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            float4* ag = reinterpret_cast<float4*>(s_acc + h * (d >> 1) + c * 4);
            float4* ab = reinterpret_cast<float4*>(s_acc + d + h * (d >> 1) + c * 4);
            float4 a = *ag, b = *ab;
            a.x += cg[4 * h + 0]; a.y += cg[4 * h + 1]; a.z += cg[4 * h + 2]; a.w += cg[4 * h + 3];
            b.x += cb[4 * h + 0]; b.y += cb[4 * h + 1]; b.z += cb[4 * h + 2]; b.w += cb[4 * h + 3];
            *ag = a;
            *ab = b;
          }
        }
      }
    }
  }
  __syncthreads();
  float4* mine = reinterpret_cast<float4*>(p.partials + static_cast<size_t>(blockIdx.x) * 2 * d);
  for (int i = threadIdx.x; i < (2 * d) >> 2; i += 256) mine[i] = reinterpret_cast<const float4*>(s_acc)[i];
}

// Folds the per-CTA partial column sums [n][2][d] of layernorm_bwd_fused_kernel into dgamma / dbeta.
// One CTA per 64 (stored) columns: 64 column lanes x 4 row groups, every thread keeps 8 loads in flight.
__global__ void __launch_bounds__(256)
ln_param_fold_kernel(const float* __restrict__ partials, int n, int d, __nv_bfloat16* __restrict__ dgamma,
                     __nv_bfloat16* __restrict__ dbeta, int accumulate) {
  pdl_wait();
  __shared__ float red[4][64];
  const int cl = threadIdx.x & 63, rg = threadIdx.x >> 6;
  const int i = blockIdx.x * 64 + cl;              // index into the stored [2][d] layout
  float s = 0.f;
  if (i < 2 * d) {
    int k = rg;
    for (; k + 28 < n; k += 32) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = __ldcg(partials + static_cast<size_t>(k + 4 * u) * 2 * d + i);
#pragma unroll
      for (int u = 0; u < 8; ++u) s += v[u];
    }
    for (; k < n; k += 4) s += __ldcg(partials + static_cast<size_t>(k) * 2 * d + i);
  }
  red[rg][cl] = s;
  __syncthreads();
  if (rg == 0 && i < 2 * d) {
    s = (red[0][cl] + red[1][cl]) + (red[2][cl] + red[3][cl]);
    // undo the shared-memory layout: i = which*d + half*(d/2) + c*4 + jj  ->  column c*8 + half*4 + jj
    const int which = i / d, r = i - which * d;
    const int half = r / (d >> 1), rr = r - half * (d >> 1);
    const int col = (rr >> 2) * 8 + half * 4 + (rr & 3);
    __nv_bfloat16* out = which ? dbeta : dgamma;
    if (accumulate) s += __bfloat162float(out[col]);
    out[col] = __float2bfloat16(s);
  }
}

// ----------------------------------------------------------------------------------------
// Column reductions over rows (bias gradients; LayerNorm gamma/beta gradients).
// grid (ceil(N/256), R): block = 32 column-vectors x 8 row lanes, each thread owns 8 columns;
// per-block partial sums -> scratch[R][K][N] (fp32); the last CTA of each column block to finish
// (completion counter at the head of the scratch buffer, self-resetting) sums the R partials and writes bf16,
// optionally accumulating into the destination.  The counter region must be zero before the
// FIRST call on a scratch buffer.
//   MODE 0: out0[n] = sum_m dy[m,n]
//   MODE 1: out0[n] = sum_m dy[m,n] * xhat[m,n]  (dgamma),  out1[n] = sum_m dy[m,n]  (dbeta)
// ----------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256)
colreduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                 const float* __restrict__ mean, const float* __restrict__ rstd,
                 float* __restrict__ scratch, unsigned int* __restrict__ counters, int M, int N, int ld,
                 __nv_bfloat16* __restrict__ out0, __nv_bfloat16* __restrict__ out1, int accumulate) {
  pdl_wait();
  constexpr int K = (MODE == 1) ? 2 : 1;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + tx) * 8;
  float acc[K][8];
#pragma unroll
  for (int k = 0; k < K; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[k][j] = 0.f;
  if (col < N) {
    for (int r = blockIdx.y * 8 + ty; r < M; r += gridDim.y * 8) {
      float fd[8];
      unpack8(*reinterpret_cast<const uint4*>(dy + static_cast<size_t>(r) * ld + col), fd);
      if (MODE == 1) {
        float fx[8];
        unpack8(*reinterpret_cast<const uint4*>(x + static_cast<size_t>(r) * ld + col), fx);
        const float mu = mean[r], rs = rstd[r];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc[0][j] += fd[j] * (fx[j] - mu) * rs;
          acc[K - 1][j] += fd[j];
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[0][j] += fd[j];
      }
    }
  }
  __shared__ float red[8][K][32][9];  // +1 padding against bank conflicts
  __shared__ bool is_last;
#pragma unroll
  for (int k = 0; k < K; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) red[ty][k][tx][j] = acc[k][j];
  __syncthreads();
  const int R = gridDim.y;
  for (int idx = threadIdx.x; idx < K * 256; idx += 256) {
    const int k = idx / 256, c = idx % 256;
    const int gcol = blockIdx.x * 256 + c;
    if (gcol < N) {
      float s = 0.f;
#pragma unroll
      for (int y = 0; y < 8; ++y) s += red[y][k][c >> 3][c & 7];
      scratch[(static_cast<size_t>(blockIdx.y) * K + k) * N + gcol] = s;
    }
  }
  // The last CTA of each column block to finish folds the R partials (threadfence-reduction).
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(&counters[blockIdx.x], 1u);
    is_last = (prev == static_cast<unsigned int>(R - 1));
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  for (int idx = threadIdx.x; idx < K * 256; idx += 256) {
    const int k = idx / 256, c = idx % 256;
    const int gcol = blockIdx.x * 256 + c;
    if (gcol < N) {
      float s = 0.f;
#pragma unroll 8
      for (int r = 0; r < R; ++r) s += __ldcg(&scratch[(static_cast<size_t>(r) * K + k) * N + gcol]);
      __nv_bfloat16* out = (k == 0) ? out0 : out1;
      if (accumulate) s += __bfloat162float(out[gcol]);
      out[gcol] = __float2bfloat16(s);
    }
  }
  if (threadIdx.x == 0) counters[blockIdx.x] = 0;  // leave the counters ready for the next call
}

// ----------------------------------------------------------------------------------------
// Embedding: out[m,:] = dropout(wte[tok[m],:] + wpe[m % T,:])
// ----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
embedding_fwd_kernel(const int* __restrict__ tok, const __nv_bfloat16* __restrict__ wte,
                     const __nv_bfloat16* __restrict__ wpe, __nv_bfloat16* __restrict__ out, int M,
                     int T, int d, uint32_t thr, float scale, uint32_t seed, const uint32_t* __restrict__ seed_ptr) {
  pdl_wait();
  if (thr && seed_ptr) seed ^= __ldg(seed_ptr);
  const int nvec = d >> 3;
  const size_t total = static_cast<size_t>(M) * nvec;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int m = static_cast<int>(i / nvec), c = static_cast<int>(i % nvec);
    const int t = m % T;
    const int id = tok[m];
    float a[8], b[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(wte + static_cast<size_t>(id) * d) + c), a);
    unpack8(__ldg(reinterpret_cast<const uint4*>(wpe + static_cast<size_t>(t) * d) + c), b);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float v = a[j] + b[j];
      if (thr) {
        const uint64_t idx = static_cast<uint64_t>(m) * d + c * 8 + j;
        v = dropout_keep(seed, idx, thr) ? v * scale : 0.f;
      }
      a[j] = v;
    }
    reinterpret_cast<uint4*>(out + static_cast<size_t>(m) * d)[c] = pack8(a);
  }
}

// Backward: dwte[tok[m],:] += sum over the positions holding that token of mask(dx[m,:]), dwpe[t,:] (+)= sum_b
// mask(dx[b,t,:]).  Tokens repeat (heavily at a 512-word vocabulary, and in real text), and dwte already holds the
// lm_head weight gradient of the tied matrix: adding position by position in bf16 would round after every add and
// depend on the order of the atomics.  Instead ONE CTA owns each distinct token - the CTA of the token's first
// position: it collects the later positions with the same token in increasing order (ballot compaction, 4096
// positions per round), sums their rows in fp32 and adds the sum to the dwte row once, without atomics.  The result
// is deterministic and rounded once per 4096 positions.  CTAs of repeated positions exit after the scan.
// One CTA per position m; the CTAs of batch element 0 also own the position-embedding row t = m.
constexpr int EB_THREADS = 128;
constexpr int EB_SEG = 4096;                       // positions per compaction round
constexpr int EB_WARPS = EB_THREADS / 32;
constexpr int EB_PER_WARP = EB_SEG / EB_WARPS;

__global__ void __launch_bounds__(EB_THREADS)
embedding_bwd_kernel(const int* __restrict__ tok, const __nv_bfloat16* __restrict__ dx,
                     __nv_bfloat16* __restrict__ dwte, __nv_bfloat16* __restrict__ dwpe, int B, int T,
                     int d, uint32_t thr, float scale, uint32_t seed, int accumulate_wpe,
                     const uint32_t* __restrict__ seed_ptr) {
  pdl_wait();
  if (thr && seed_ptr) seed ^= __ldg(seed_ptr);
  __shared__ int s_list[EB_WARPS][EB_PER_WARP];
  __shared__ int s_n[EB_WARPS];
  const int M = B * T, m = blockIdx.x, nvec = d >> 3;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  auto load_masked = [&](int row, int c, float (&g)[8]) {
    unpack8(reinterpret_cast<const uint4*>(dx + static_cast<size_t>(row) * d)[c], g);
    if (thr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint64_t idx = static_cast<uint64_t>(row) * d + c * 8 + j;
        g[j] = dropout_keep(seed, idx, thr) ? g[j] * scale : 0.f;
      }
    }
  };
  // ---- position embedding: row t = m, summed over the batch ---------------------------------
  if (m < T) {
    for (int c = threadIdx.x; c < nvec; c += EB_THREADS) {
      float pos[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int b = 0; b < B; ++b) {
        float g[8];
        load_masked(b * T + m, c, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) pos[j] += g[j];
      }
      uint4* pw = reinterpret_cast<uint4*>(dwpe + static_cast<size_t>(m) * d) + c;
      if (accumulate_wpe) {
        float old[8];
        unpack8(*pw, old);
#pragma unroll
        for (int j = 0; j < 8; ++j) pos[j] += old[j];
      }
      *pw = pack8(pos);
    }
  }
  // ---- token embedding: only the first position of a token goes on ---------------------------
  const int id = tok[m];
  int earlier = 0;
  for (int j = threadIdx.x; j < m; j += EB_THREADS) earlier |= (tok[j] == id);
  if (__syncthreads_or(earlier)) return;
  for (int s0 = (m / EB_SEG) * EB_SEG; s0 < M; s0 += EB_SEG) {
    int n = 0;                                  // warp-uniform: matches found by this warp so far
    for (int k = 0; k < EB_PER_WARP; k += 32) {
      const int j = s0 + warp * EB_PER_WARP + k + lane;
      const bool hit = j >= m && j < M && tok[j] == id;
      const unsigned mask = __ballot_sync(0xffffffffu, hit);
      if (hit) s_list[warp][n + __popc(mask & ((1u << lane) - 1u))] = j;
      n += __popc(mask);
    }
    if (lane == 0) s_n[warp] = n;
    __syncthreads();
    for (int c = threadIdx.x; c < nvec; c += EB_THREADS) {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int w = 0; w < EB_WARPS; ++w) {
        const int nw = s_n[w];
#pragma unroll 4
        for (int i = 0; i < nw; ++i) {
          float g[8];
          load_masked(s_list[w][i], c, g);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += g[j];
        }
      }
      uint4* pw = reinterpret_cast<uint4*>(dwte + static_cast<size_t>(id) * d) + c;
      float old[8];
      unpack8(*pw, old);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += old[j];
      *pw = pack8(acc);
    }
    __syncthreads();                            // the lists are rebuilt in the next round
  }
}

// ----------------------------------------------------------------------------------------
// Fused softmax cross-entropy forward + backward over materialised bf16 logits.
// One CTA per row; the row is read ONCE into registers, the gradient overwrites the logits.
//   loss_row[m] = lse - logit[target]   (0 when target == ignore_index)
//   dlogits     = (softmax - onehot) * grad_scale / n_valid
// ----------------------------------------------------------------------------------------
constexpr int CE_THREADS = 512;
constexpr int CE_MAXV = 8;  // uint4 per thread cached in registers: V <= 8*8*512 = 32768

__global__ void count_valid_kernel(const int* __restrict__ targets, int M, int ignore_index,
                                   int* __restrict__ n_valid) {
  pdl_wait();
  __shared__ int s;
  if (threadIdx.x == 0) s = 0;
  __syncthreads();
  int c = 0;
  for (int i = threadIdx.x; i < M; i += blockDim.x) c += (targets[i] != ignore_index);
  atomicAdd(&s, c);
  __syncthreads();
  if (threadIdx.x == 0) *n_valid = s;
}

__device__ __forceinline__ float block_reduce(float v, float* sh, bool is_max) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  v = is_max ? warp_max(v) : warp_sum(v);
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  const int nw = blockDim.x >> 5;
  float r = (lane < nw) ? sh[lane] : (is_max ? -INFINITY : 0.f);
  r = is_max ? warp_max(r) : warp_sum(r);
  return r;
}

__global__ void __launch_bounds__(CE_THREADS)
cross_entropy_kernel(__nv_bfloat16* __restrict__ logits, const int* __restrict__ targets,
                     float* __restrict__ loss_rows, int V, int ld, int ignore_index,
                     const int* __restrict__ n_valid, float grad_scale, int write_grad) {
  pdl_wait();
  __shared__ float sh[32];
  const int row = blockIdx.x;
  uint4* lr = reinterpret_cast<uint4*>(logits + static_cast<size_t>(row) * ld);
  const int nvec = V >> 3;
  const int tgt = targets[row];
  const bool cached = nvec <= CE_MAXV * CE_THREADS;
  uint4 v[CE_MAXV];
  float mx = -INFINITY;
  if (cached) {
#pragma unroll
    for (int i = 0; i < CE_MAXV; ++i) {
      const int c = threadIdx.x + i * CE_THREADS;
      if (c < nvec) {
        v[i] = lr[c];
        float f[8];
        unpack8(v[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) mx = fmaxf(mx, f[j]);
      }
    }
  } else {
    for (int c = threadIdx.x; c < nvec; c += CE_THREADS) {
      float f[8];
      unpack8(lr[c], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) mx = fmaxf(mx, f[j]);
    }
  }
  mx = block_reduce(mx, sh, true);
  float se = 0.f;
  constexpr float L2E = 1.4426950408889634f;
  if (cached) {
#pragma unroll
    for (int i = 0; i < CE_MAXV; ++i) {
      const int c = threadIdx.x + i * CE_THREADS;
      if (c < nvec) {
        float f[8];
        unpack8(v[i], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) se += exp2f((f[j] - mx) * L2E);
      }
    }
  } else {
    for (int c = threadIdx.x; c < nvec; c += CE_THREADS) {
      float f[8];
      unpack8(lr[c], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) se += exp2f((f[j] - mx) * L2E);
    }
  }
  se = block_reduce(se, sh, false);
  const float lse = mx + logf(se);
  const bool valid = (tgt != ignore_index);
  if (threadIdx.x == 0) {
    float l = 0.f;
    if (valid) l = lse - __bfloat162float(logits[static_cast<size_t>(row) * ld + tgt]);
    loss_rows[row] = l;
  }
  if (!write_grad) return;
  __syncthreads();  // the target logit was read above before anyone overwrites it
  const int nv = *n_valid;
  const float gs = (valid && nv > 0) ? grad_scale / nv : 0.f;
  const float inv_se = 1.f / se;
  auto emit = [&](int c, const uint4& u) {
    float f[8];
    unpack8(u, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float p = exp2f((f[j] - mx) * L2E) * inv_se;
      if (c * 8 + j == tgt) p -= 1.f;
      f[j] = p * gs;
    }
    lr[c] = pack8(f);
  };
  if (cached) {
#pragma unroll
    for (int i = 0; i < CE_MAXV; ++i) {
      const int c = threadIdx.x + i * CE_THREADS;
      if (c < nvec) emit(c, v[i]);
    }
  } else {
    for (int c = threadIdx.x; c < nvec; c += CE_THREADS) emit(c, lr[c]);
  }
}

__global__ void ce_finalize_kernel(const float* __restrict__ loss_rows, int M,
                                   const int* __restrict__ n_valid, float* __restrict__ loss_out,
                                   int accumulate, float scale) {
  pdl_wait();
  __shared__ float sh[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < M; i += blockDim.x) s += loss_rows[i];
  s = block_reduce(s, sh, false);
  if (threadIdx.x == 0) {
    const int nv = *n_valid;
    const float l = nv > 0 ? s / nv * scale : 0.f;
    *loss_out = accumulate ? (*loss_out + l) : l;
  }
}

// ----------------------------------------------------------------------------------------
// Flat fused AdamW.  Hyper-parameters are read from device memory so that the launch can sit in
// a CUDA graph while lr / bias corrections change every step:
//   hp[0]=lr hp[1]=beta1 hp[2]=beta2 hp[3]=eps hp[4]=weight_decay hp[5]=1/bias_corr1
//   hp[6]=1/sqrt(bias_corr2) hp[7]=grad_scale hp[8]=clip_norm (0 = off)
// gnorm_sq (nullable): squared global gradient norm (already multiplied by grad_scale^2 or not,
// see sumsq scale) used for clipping: coef = min(1, clip / (sqrt(gnorm_sq) + 1e-6)).
// ----------------------------------------------------------------------------------------
struct AdamCoef {
  float lr, b1, b2, eps, wd, ibc1, isbc2, gscale;
};
__device__ __forceinline__ AdamCoef load_adam_coef(const float* __restrict__ hp,
                                                   const float* __restrict__ gnorm_sq) {
  AdamCoef c;
  c.lr = hp[0];
  c.b1 = hp[1];
  c.b2 = hp[2];
  c.eps = hp[3];
  c.wd = hp[4];
  c.ibc1 = hp[5];
  c.isbc2 = hp[6];
  c.gscale = hp[7];
  const float clip = hp[8];
  if (gnorm_sq != nullptr && clip > 0.f) {
    const float norm = sqrtf(*gnorm_sq);
    if (norm > clip) c.gscale *= clip / (norm + 1e-6f);
  }
  return c;
}
__device__ __forceinline__ void adam_update4(float4& w, float4& m, float4& v, const float (&g)[4],
                                             const AdamCoef& c) {
  float* wp = &w.x;
  float* mp = &m.x;
  float* vp = &v.x;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float gj = g[j] * c.gscale;
    mp[j] = c.b1 * mp[j] + (1.f - c.b1) * gj;
    vp[j] = c.b2 * vp[j] + (1.f - c.b2) * gj * gj;
    // IEEE square root and division (round-to-nearest intrinsics are not affected by --use_fast_math), so the
    // update stays comparable with torch.optim.AdamW to fp32 rounding
    const float denom = __fsqrt_rn(vp[j]) * c.isbc2 + c.eps;
    wp[j] = wp[j] * (1.f - c.lr * c.wd) - __fdiv_rn(c.lr * c.ibc1 * mp[j], denom);
  }
}

__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
             const __nv_bfloat16* __restrict__ grad, __nv_bfloat16* __restrict__ w_out, size_t n,
             const float* __restrict__ hp, const float* __restrict__ gnorm_sq) {
  pdl_wait();
  const AdamCoef c = load_adam_coef(hp, gnorm_sq);
  const size_t nvec = n >> 3;  // 8 elements per iteration (one uint4 of bf16 grads)
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    float g[8];
    unpack8(reinterpret_cast<const uint4*>(grad)[i], g);
    float4* wp = reinterpret_cast<float4*>(master) + 2 * i;
    float4* mp = reinterpret_cast<float4*>(m) + 2 * i;
    float4* vp = reinterpret_cast<float4*>(v) + 2 * i;
    float4 w0 = wp[0], w1 = wp[1], m0 = mp[0], m1 = mp[1], v0 = vp[0], v1 = vp[1];
    const float g0[4] = {g[0], g[1], g[2], g[3]}, g1[4] = {g[4], g[5], g[6], g[7]};
    adam_update4(w0, m0, v0, g0, c);
    adam_update4(w1, m1, v1, g1, c);
    wp[0] = w0;
    wp[1] = w1;
    mp[0] = m0;
    mp[1] = m1;
    vp[0] = v0;
    vp[1] = v1;
    const float o[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
    reinterpret_cast<uint4*>(w_out)[i] = pack8(o);
  }
}

__global__ void __launch_bounds__(256)
sumsq_kernel(const __nv_bfloat16* __restrict__ g, size_t n, float scale, float* __restrict__ out) {
  pdl_wait();
  __shared__ float sh[32];
  float s = 0.f;
  const size_t nvec = n >> 3;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    float f[8];
    unpack8(reinterpret_cast<const uint4*>(g)[i], f);
#pragma unroll
    for (int j = 0; j < 8; ++j) s += f[j] * f[j];
  }
  s = block_reduce(s, sh, false);
  if (threadIdx.x == 0) atomicAdd(out, s * scale * scale);
}

__global__ void __launch_bounds__(256)
dropout_apply_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, int M,
                     int N, uint32_t thr, float scale, uint32_t seed,
                     const uint32_t* __restrict__ seed_ptr) {
  pdl_wait();
  if (seed_ptr) seed ^= __ldg(seed_ptr);
  const int nvec = N >> 3;
  const size_t total = static_cast<size_t>(M) * nvec;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t m = i / nvec;
    const int c = static_cast<int>(i % nvec);
    float f[8];
    unpack8(reinterpret_cast<const uint4*>(in)[i], f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint64_t idx = m * N + c * 8 + j;
      f[j] = dropout_keep(seed, idx, thr) ? f[j] * scale : 0.f;
    }
    reinterpret_cast<uint4*>(out)[i] = pack8(f);
  }
}

__global__ void fill_adam_hp_kernel(float* hp, float lr, float b1, float b2, float eps, float wd,
                                    float ibc1, float isbc2, float gscale, float clip) {
  hp[0] = lr; hp[1] = b1; hp[2] = b2; hp[3] = eps; hp[4] = wd;
  hp[5] = ibc1; hp[6] = isbc2; hp[7] = gscale; hp[8] = clip;
}

static const uint32_t* g_step_seed_ptr = nullptr;
const uint32_t* step_seed_ptr() { return g_step_seed_ptr; }

__global__ void bump_step_seed_kernel(uint32_t* w) { *w = *w * 1664525u + 1013904223u; }

static inline uint32_t drop_thr(float p) {
  if (!(p > 0.f) || p >= 1.f) return 0;
  return static_cast<uint32_t>(static_cast<double>(p) * 4294967296.0);
}
static inline int grid_for(size_t work_items, int threads, int per_sm) {
  size_t blocks = (work_items + threads - 1) / threads;
  const size_t cap = static_cast<size_t>(device_sm_count()) * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

constexpr int kColreduceCounters = 256;
// number of row splits for the column reductions: about one CTA per SM in total
static inline int colreduce_splits(int M, int N) {
  const int gx = (N + 255) / 256;
  int R = (device_sm_count() + gx - 1) / gx;
  const int max_r = (M + 31) / 32;
  if (R > max_r) R = max_r;
  if (R > 64) R = 64;
  if (R < 1) R = 1;
  return R;
}
static inline size_t colreduce_partial_floats(int M, int N, int outputs) {
  return (static_cast<size_t>(colreduce_splits(M, N)) * outputs * N + 63) / 64 * 64;
}

}  // namespace b200

using namespace b200;

#define LAUNCH_LN(KERNEL, ...)                                                    \
  do {                                                                            \
    const int nv = (d / 8 + 31) / 32;                                             \
    const int blocks = (M + 7) / 8;                                               \
    if (nv <= 1) launch_pdl(KERNEL<1>, dim3(blocks), dim3(256), 0, s, __VA_ARGS__);                 \
    else if (nv <= 2) launch_pdl(KERNEL<2>, dim3(blocks), dim3(256), 0, s, __VA_ARGS__);              \
    else if (nv <= 4) launch_pdl(KERNEL<4>, dim3(blocks), dim3(256), 0, s, __VA_ARGS__);              \
    else if (nv <= 8) launch_pdl(KERNEL<8>, dim3(blocks), dim3(256), 0, s, __VA_ARGS__);              \
    else if (nv <= 16) launch_pdl(KERNEL<16>, dim3(blocks), dim3(256), 0, s, __VA_ARGS__);            \
    else return -3;                                                               \
  } while (0)

extern "C" int b200_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y,
                                  float* mean, float* rstd, int M, int d, float eps, void* stream) {
  if (d % 8) return -2;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  LAUNCH_LN(layernorm_fwd_kernel, static_cast<const __nv_bfloat16*>(x),
            static_cast<const __nv_bfloat16*>(gamma), static_cast<const __nv_bfloat16*>(beta),
            static_cast<__nv_bfloat16*>(y), mean, rstd, M, d, eps);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int b200_layernorm_bwd(const void* dy, const void* x, const void* gamma,
                                  const float* mean, const float* rstd, const void* dres, void* dx,
                                  int M, int d, void* stream) {
  if (d % 8) return -2;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  LAUNCH_LN(layernorm_bwd_kernel, static_cast<const __nv_bfloat16*>(dy),
            static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(gamma), mean,
            rstd, static_cast<const __nv_bfloat16*>(dres), static_cast<__nv_bfloat16*>(dx), M, d);
  return static_cast<int>(cudaGetLastError());
}

// One row per warp: the kernel then needs < 128 registers per thread at d = 1024, so its CTAs fit on an SM next to
// the CTAs of the memory-bound side-stream kernels (AdamW) instead of waiting for a whole free register file.
static inline int lnb_rows_per_cta(int d) { (void)d; return 8; }
extern "C" size_t b200_layernorm_bwd_fused_scratch_bytes(int M, int d) {
  const int rpc = lnb_rows_per_cta(d);
  const size_t ctas = (M + rpc - 1) / rpc;
  return ctas * 2 * static_cast<size_t>(d) * sizeof(float);
}
extern "C" int b200_layernorm_bwd_fused(const void* dy, const void* x, const void* gamma, const float* mean,
                                        const float* rstd, const void* dres, void* dx, void* dx_drop,
                                        float drop_p, unsigned int drop_seed, void* partials,
                                        size_t partials_bytes, int M, int d, void* stream) {
  if (d % 8) return -2;
  const int rpc = lnb_rows_per_cta(d);
  const int ctas = (M + rpc - 1) / rpc;
  if (partials && partials_bytes < b200_layernorm_bwd_fused_scratch_bytes(M, d)) return -4;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  LnBwdFused p;
  p.dy = static_cast<const __nv_bfloat16*>(dy);
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.gamma = static_cast<const __nv_bfloat16*>(gamma);
  p.mean = mean;
  p.rstd = rstd;
  p.dres = static_cast<const __nv_bfloat16*>(dres);
  p.dx = static_cast<__nv_bfloat16*>(dx);
  const uint32_t thr = drop_thr(drop_p);
  p.dx_drop = thr ? static_cast<__nv_bfloat16*>(dx_drop) : nullptr;
  p.drop_thr = thr;
  p.drop_scale = thr ? 1.f / (1.f - drop_p) : 1.f;
  p.drop_seed = drop_seed;
  p.seed_ptr = step_seed_ptr();
  p.partials = static_cast<float*>(partials);
  p.M = M;
  p.d = d;
  const size_t smem = 2 * static_cast<size_t>(d) * sizeof(float);
  const int nv = (d / 8 + 31) / 32;
#define LNB_LAUNCH(NVV, RPW)                                                                                   \
  do {                                                                                                         \
    static bool attr_done[64] = {false};                                                                       \
    int dev = 0;                                                                                               \
    cudaGetDevice(&dev);                                                                                       \
    if (smem > 48 * 1024 && !attr_done[dev & 63]) {                                                            \
      cudaFuncSetAttribute(layernorm_bwd_fused_kernel<NVV, RPW>, cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                           131072);                                                                            \
      attr_done[dev & 63] = true;                                                                              \
    }                                                                                                          \
    launch_pdl(layernorm_bwd_fused_kernel<NVV, RPW>, dim3(ctas), dim3(256), smem, s, p);                       \
  } while (0)
  if (nv <= 1) LNB_LAUNCH(1, 1);
  else if (nv <= 2) LNB_LAUNCH(2, 1);
  else if (nv <= 4) LNB_LAUNCH(4, 1);
  else if (nv <= 8) LNB_LAUNCH(8, 1);
  else if (nv <= 16) LNB_LAUNCH(16, 1);
  else return -3;
#undef LNB_LAUNCH
  return static_cast<int>(cudaGetLastError());
}
/* dgamma / dbeta (+)= the partial column sums a b200_layernorm_bwd_fused call left in `partials` */
extern "C" int b200_layernorm_param_fold(const void* partials, int M, int d, void* dgamma, void* dbeta,
                                         int accumulate, void* stream) {
  const int rpc = lnb_rows_per_cta(d);
  const int n = (M + rpc - 1) / rpc;
  launch_pdl(ln_param_fold_kernel, dim3((2 * d + 63) / 64), dim3(256), 0, static_cast<cudaStream_t>(stream),
             static_cast<const float*>(partials), n, d, static_cast<__nv_bfloat16*>(dgamma),
             static_cast<__nv_bfloat16*>(dbeta), accumulate);
  return static_cast<int>(cudaGetLastError());
}

extern "C" size_t b200_colreduce_scratch_bytes(int M, int N, int outputs) {
  // 256 completion counters (one per 256-column block, fixed place at the head of the buffer so that
  // calls with different shapes can share one scratch buffer) + fp32 partials [splits][outputs][N]
  return kColreduceCounters * sizeof(unsigned int) +
         colreduce_partial_floats(M, N, outputs) * sizeof(float);
}

extern "C" int b200_layernorm_bwd_params(const void* dy, const void* x, const float* mean,
                                         const float* rstd, void* dgamma, void* dbeta, int M, int d,
                                         int accumulate, void* scratch, size_t scratch_bytes,
                                         void* stream) {
  if (d % 8) return -2;
  if (scratch_bytes < b200_colreduce_scratch_bytes(M, d, 2)) return -4;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int R = colreduce_splits(M, d);
  dim3 grid((d + 255) / 256, R);
  if ((d + 255) / 256 > kColreduceCounters) return -3;
  unsigned int* counters = static_cast<unsigned int*>(scratch);
  float* part = reinterpret_cast<float*>(counters + kColreduceCounters);
  launch_pdl(colreduce_kernel<1>, dim3(grid), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(dy),
                                           static_cast<const __nv_bfloat16*>(x), mean, rstd, part,
                                           counters, M, d, d, static_cast<__nv_bfloat16*>(dgamma),
                                           static_cast<__nv_bfloat16*>(dbeta), accumulate);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int b200_colsum(const void* dy, void* out, int M, int N, int ld, int accumulate,
                           void* scratch, size_t scratch_bytes, void* stream) {
  if ((N % 8) || (ld % 8)) return -2;
  if (scratch_bytes < b200_colreduce_scratch_bytes(M, N, 1)) return -4;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int R = colreduce_splits(M, N);
  dim3 grid((N + 255) / 256, R);
  if ((N + 255) / 256 > kColreduceCounters) return -3;
  unsigned int* counters = static_cast<unsigned int*>(scratch);
  float* part = reinterpret_cast<float*>(counters + kColreduceCounters);
  launch_pdl(colreduce_kernel<0>, dim3(grid), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(dy), nullptr, nullptr,
                                           nullptr, part, counters, M, N, ld,
                                           static_cast<__nv_bfloat16*>(out), nullptr, accumulate);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int b200_embedding_fwd(const int* tokens, const void* wte, const void* wpe, void* out,
                                  int B, int T, int d, float drop_p, unsigned int seed,
                                  void* stream) {
  if (d % 8) return -2;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t work = static_cast<size_t>(B) * T * (d / 8);
  const float scale = drop_p > 0.f && drop_p < 1.f ? 1.f / (1.f - drop_p) : 1.f;
  launch_pdl(embedding_fwd_kernel, dim3(grid_for(work, 256, 8)), dim3(256), 0, s, 
      tokens, static_cast<const __nv_bfloat16*>(wte), static_cast<const __nv_bfloat16*>(wpe),
      static_cast<__nv_bfloat16*>(out), B * T, T, d, drop_thr(drop_p), scale, seed, step_seed_ptr());
  return static_cast<int>(cudaGetLastError());
}

extern "C" int b200_embedding_bwd(const int* tokens, const void* dx, void* dwte, void* dwpe, int B,
                                  int T, int d, float drop_p, unsigned int seed, int accumulate_wpe,
                                  void* stream) {
  if (d % 8) return -2;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const float scale = drop_p > 0.f && drop_p < 1.f ? 1.f / (1.f - drop_p) : 1.f;
  launch_pdl(embedding_bwd_kernel, dim3(B * T), dim3(EB_THREADS), 0, s, 
      tokens, static_cast<const __nv_bfloat16*>(dx), static_cast<__nv_bfloat16*>(dwte),
      static_cast<__nv_bfloat16*>(dwpe), B, T, d, drop_thr(drop_p), scale, seed, accumulate_wpe,
      step_seed_ptr());
  return static_cast<int>(cudaGetLastError());
}

extern "C" int b200_cross_entropy(void* logits, const int* targets, float* loss_rows,
                                  int* n_valid, float* loss_out, int M, int V, int ld,
                                  int ignore_index, float grad_scale, int write_grad,
                                  int accumulate_loss, float loss_scale, void* stream) {
  if ((V % 8) || (ld % 8)) return -2;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  launch_pdl(count_valid_kernel, dim3(1), dim3(1024), 0, s, targets, M, ignore_index, n_valid);
  launch_pdl(cross_entropy_kernel, dim3(M), dim3(CE_THREADS), 0, s, static_cast<__nv_bfloat16*>(logits), targets,
                                                loss_rows, V, ld, ignore_index, n_valid, grad_scale,
                                                write_grad);
  launch_pdl(ce_finalize_kernel, dim3(1), dim3(1024), 0, s, loss_rows, M, n_valid, loss_out, accumulate_loss,
                                        loss_scale);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int b200_adamw_set_hp(float* hp, float lr, float beta1, float beta2, float eps, float wd,
                                 float inv_bias_corr1, float inv_sqrt_bias_corr2, float grad_scale,
                                 float clip_norm, void* stream) {
  fill_adam_hp_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(
      hp, lr, beta1, beta2, eps, wd, inv_bias_corr1, inv_sqrt_bias_corr2, grad_scale, clip_norm);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int b200_adamw(float* master, float* m, float* v, const void* grad, void* w_out,
                          size_t n, const float* hp, const float* gnorm_sq, void* stream) {
  if (n % 8) return -2;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  launch_pdl(adamw_kernel, dim3(grid_for(n / 8, 256, 8)), dim3(256), 0, s, master, m, v,
                                                       static_cast<const __nv_bfloat16*>(grad),
                                                       static_cast<__nv_bfloat16*>(w_out), n, hp,
                                                       gnorm_sq);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int b200_sumsq(const void* grad, size_t n, float scale, float* out, void* stream) {
  if (n % 8) return -2;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  launch_pdl(sumsq_kernel, dim3(grid_for(n / 8, 256, 4)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(grad), n,
                                                       scale, out);
  return static_cast<int>(cudaGetLastError());
}

extern "C" int b200_dropout_apply(const void* in, void* out, int M, int N, float drop_p,
                                  unsigned int seed, void* stream) {
  if (N % 8) return -2;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const uint32_t thr = drop_thr(drop_p);
  if (!thr) {
    if (in != out)
      cudaMemcpyAsync(out, in, static_cast<size_t>(M) * N * 2, cudaMemcpyDeviceToDevice, s);
    return static_cast<int>(cudaGetLastError());
  }
  const size_t work = static_cast<size_t>(M) * (N / 8);
  launch_pdl(dropout_apply_kernel, dim3(grid_for(work, 256, 8)), dim3(256), 0, s, 
      static_cast<const __nv_bfloat16*>(in), static_cast<__nv_bfloat16*>(out), M, N, thr,
      1.f / (1.f - drop_p), seed, step_seed_ptr());
  return static_cast<int>(cudaGetLastError());
}

// Registers a device word that every dropout site XORs into its seed, and advances it.  The engine
// captures b200_step_seed_advance at the head of its micro-step graph, so each replay draws new
// masks while forward and backward of the same micro-step still agree.
extern "C" int b200_step_seed_register(unsigned int* device_word) {
  g_step_seed_ptr = device_word;
  return 0;
}
extern "C" int b200_step_seed_advance(void* stream) {
  if (!g_step_seed_ptr) return -6;
  bump_step_seed_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(
      const_cast<uint32_t*>(g_step_seed_ptr));
  return static_cast<int>(cudaGetLastError());
}

// ----------------------------------------------------------------------------------------------
// b200_colsum_static: the bias-gradient column sum for callers that must not allocate.
//
// b200_colsum needs a scratch buffer (fp32 partial sums + completion counters) that the caller has to
// obtain from somewhere.  Some callers have nowhere to get it: a C plug-in running inside a host
// framework whose memory pool already owns almost all of HBM (a fresh cudaMalloc there either fails
// or fragments the pool), or a step loop that may not touch the CUDA allocator at all because
// cudaMalloc/cudaFree take driver-wide locks and synchronise the device.  For them the scratch lives
// in a file-scope device array that is part of the library image: every GPU that loads the library
// holds kStaticScratchBytes (512 KiB) of it until the process exits, and nothing is ever allocated or
// freed by this entry point (no cudaMalloc, no cudaFree, no library routine that allocates).
//
// In host code the name of a __device__ array is only a host-side stand-in for the symbol, not a
// device address, so it cannot be handed to a kernel; the address is looked up with
// cudaGetSymbolAddress once per device, with that device current, on the FIRST call made on that
// device.  Make that first call before any stream capture starts.  The per-device address cache is written
// without a lock: two host threads that make the first call on the same device at the same moment both store
// the same address, which is harmless.
//
// The array is shared by every call and keeps whatever the previous call left in it.  The completion
// counters at its head must be zero when the kernel starts, so they are cleared on the caller's
// stream before every launch; the partial sums need no clearing (every slot that is read is written
// first).  Because there is only one array per GPU, two calls whose kernels overlap on the same GPU
// (different streams, or different host threads) race on it: keep calls on one device stream-ordered.
// Shapes beyond the built-in buffer (more than 65536 columns, or more than kStaticScratchBytes of
// partials) are rejected with B200_ERR_STATIC_SCRATCH and a message on stderr; use b200_colsum with
// your own buffer for those.
// ----------------------------------------------------------------------------------------------
namespace b200 {
constexpr size_t kStaticScratchBytes = 512u << 10;  // 512 KiB per GPU: enough for any N <= 65536
static __device__ __align__(256) unsigned char g_colsum_static_scratch[kStaticScratchBytes];
constexpr int kMaxDevices = 64;
static void* g_static_scratch_addr[kMaxDevices] = {nullptr};
}  // namespace b200

extern "C" size_t b200_colsum_static_max_bytes(void) { return b200::kStaticScratchBytes; }

extern "C" int b200_colsum_static(const void* dy, void* out, int M, int N, int ld, int accumulate,
                                  void* stream) {
  if (M <= 0 || N <= 0) return 0;
  if ((N % 8) || (ld % 8)) return -2;
  const size_t need = b200_colreduce_scratch_bytes(M, N, 1);
  if ((N + 255) / 256 > kColreduceCounters || need > kStaticScratchBytes) {
    fprintf(stderr,
            "b200_colsum_static: M=%d N=%d is beyond the built-in scratch (at most %d columns and %zu "
            "bytes; this shape needs %zu bytes); call b200_colsum with a caller-provided buffer\n",
            M, N, kColreduceCounters * 256, static_cast<size_t>(kStaticScratchBytes), need);
    return B200_ERR_STATIC_SCRATCH;
  }
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return static_cast<int>(e);
  if (dev < 0 || dev >= kMaxDevices) return -3;
  if (g_static_scratch_addr[dev] == nullptr) {
    // once per device, with that device current: resolve the symbol to a real device address
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_colsum_static_scratch);
    if (e != cudaSuccess) return static_cast<int>(e);
    g_static_scratch_addr[dev] = ptr;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  unsigned int* counters = static_cast<unsigned int*>(g_static_scratch_addr[dev]);
  float* part = reinterpret_cast<float*>(counters + kColreduceCounters);
  // the previous call's leftovers must not be mistaken for arrivals of this one
  e = cudaMemsetAsync(counters, 0, kColreduceCounters * sizeof(unsigned int), s);
  if (e != cudaSuccess) return static_cast<int>(e);
  const int R = colreduce_splits(M, N);
  dim3 grid((N + 255) / 256, R);
  launch_pdl(colreduce_kernel<0>, dim3(grid), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(dy), nullptr, nullptr,
                                           nullptr, part, counters, M, N, ld,
                                           static_cast<__nv_bfloat16*>(out), nullptr, accumulate);
  return static_cast<int>(cudaGetLastError());
}
