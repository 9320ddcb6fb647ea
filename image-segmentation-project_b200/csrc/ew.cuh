// Shared plumbing of the HBM-bound layer kernels (section 4 of include/b2seg.h).
//
// Activations are NHWC bf16 seen as [P pixels][C channels] with a leading
// dimension ld; every thread moves 8 channels (one 128-bit word).  Per-channel
// reductions are two-phase: each block writes one row of fp32 partials, a
// finalize kernel sums the rows in fixed order (fp64) -> bit-reproducible.
#pragma once
#include <stdlib.h>
#include "ops.cuh"

struct EwArgs {
  void* p[B2SEG_EW_NPTR];
  long long i[B2SEG_EW_NINT];
  float f[4];
};

typedef int (*EwLaunch)(const EwArgs& a, dim3 grid, dim3 block, size_t smem, cudaStream_t st);

struct EwPlan {
  EwLaunch fn;
  dim3 grid, block;
  size_t smem;
  int launches;
  size_t smem_limit;  // 0 = the default 48 KB; planners that opt a kernel in to more set this
};

// planners, one per source file; each fills `plan` for desc->kind or returns an error
int ew_plan_bn(const b2seg_ew_desc* d, EwPlan* plan);
int ew_plan_stem(const b2seg_ew_desc* d, EwPlan* plan);
int ew_plan_attn(const b2seg_ew_desc* d, EwPlan* plan);
int ew_plan_se(const b2seg_ew_desc* d, EwPlan* plan);

constexpr int kEwThreads = 256;
// bn_bwd_reduce keeps four pixels' operands in registers (119 registers): two resident blocks per SM
constexpr int kBnBwdReduceBlocksPerSm = 2;

static inline int ew_pl(int C) {  // pixel lanes of a 256-thread block whose threads own 8 channels
  const int cv = C / 8;
  const int pl = kEwThreads / (cv > 0 ? cv : 1);
  return pl < 1 ? 1 : pl;
}
static inline int ew_rows(long long P, int C, int blocks_per_sm = 3) {
  const long long per = (long long)ew_pl(C) * 8;
  long long r = (P + per - 1) / per;
  if (r < 1) r = 1;
  // one block per partial row: at most 3 resident 256-thread blocks per SM (76-101 registers),
  // so more rows than 3 x SMs would run as a second, mostly empty wave
  long long cap = (long long)blocks_per_sm * b2_num_sms();
  if (cap > B2SEG_EW_MAX_PARTIAL_ROWS) cap = B2SEG_EW_MAX_PARTIAL_ROWS;
  if (r > cap) r = cap;
  return (int)r;
}
static inline int ew_blocks(long long work_items) {  // elementwise grid: 256 threads, grid-stride
  long long b = (work_items + kEwThreads - 1) / kEwThreads;
  const long long cap = (long long)b2_num_sms() * 12;  // whole waves at 2, 3, 4 or 6 blocks per SM
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}
// Grid of a grid-stride streaming kernel as ONE resident wave (occupancy API): with more blocks than fit,
// the blocks run in several waves of a few items per thread each and the batched loads never get going
// (bn_bwd_apply: 124 registers -> 2 blocks per SM, the old cap of 12 blocks per SM meant 6 waves).
// B2SEG_EW_WAVE=0 restores the capped grid.
#ifdef __CUDACC__
template <typename K>
static inline int ew_resident_blocks(K kernel, long long work_items) {
  int per_sm = 0;  // queried per op creation (both kernels share one function type: no static cache)
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kEwThreads, 0) != cudaSuccess || per_sm < 1) {
    cudaGetLastError();
    per_sm = 2;
  }
  const char* e = getenv("B2SEG_EW_WAVE");
  if (e && e[0] == '0') return ew_blocks(work_items);
  long long b = (work_items + kEwThreads - 1) / kEwThreads;
  const long long cap = (long long)per_sm * b2_num_sms();
  if (b > cap) b = cap;
  return (int)(b < 1 ? 1 : b);
}
#endif
// Number of row slices (gridDim.y) of a column-reduce finalize kernel: each (32 x 32)-thread
// block then walks <= ~32 rows per thread (independent loads, one L2 round trip); block-partial
// buffers have <= 592 rows, so these kernels are single-phase -- no ticket, no second fold.
static inline int ew_fin_slices(long long rows, int colblocks) {
  long long s = (rows + 1023) / 1024;
  if (s > 64) s = 64;
  while (s > 1 && s * colblocks > 1024) --s;
  if (s < 1) s = 1;
  return (int)s;
}
constexpr int kFinMaxColBlocks = 128;  // 4096 columns
constexpr int kFinMaxSlices = 64;

// sub-warp group size (power of two <= 32) for kernels that reduce over the channels of a pixel
static inline int ew_group(int C) {
  const int cv = C / 8;
  int g = 1;
  while (g < cv && g < 32) g <<= 1;
  return g;
}

#ifdef __CUDACC__
struct F8 {
  float v[8];
};
__device__ __forceinline__ F8 ld_bf8(const bf16* p) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  F8 r;
  float2 t;
  t = bf2_to_f2(u.x); r.v[0] = t.x; r.v[1] = t.y;
  t = bf2_to_f2(u.y); r.v[2] = t.x; r.v[3] = t.y;
  t = bf2_to_f2(u.z); r.v[4] = t.x; r.v[5] = t.y;
  t = bf2_to_f2(u.w); r.v[6] = t.x; r.v[7] = t.y;
  return r;
}
__device__ __forceinline__ F8 unpack_bf8(const uint4& u) {
  F8 r;
  float2 t;
  t = bf2_to_f2(u.x); r.v[0] = t.x; r.v[1] = t.y;
  t = bf2_to_f2(u.y); r.v[2] = t.x; r.v[3] = t.y;
  t = bf2_to_f2(u.z); r.v[4] = t.x; r.v[5] = t.y;
  t = bf2_to_f2(u.w); r.v[6] = t.x; r.v[7] = t.y;
  return r;
}
__device__ __forceinline__ void st_bf8(bf16* p, const F8& r) {
  uint4 u;
  u.x = f2_to_bf2(r.v[0], r.v[1]);
  u.y = f2_to_bf2(r.v[2], r.v[3]);
  u.z = f2_to_bf2(r.v[4], r.v[5]);
  u.w = f2_to_bf2(r.v[6], r.v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ F8 ld_f8(const float* p) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  F8 r;
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}

// idx -> (idx / d, idx % d).  Work-item counts of the layer kernels fit 31 bits for every shape
// the plan builds, and a 64-bit division costs ~100 instructions against ~20 for 32 bits: next to
// one 16-byte load per item the index arithmetic otherwise dominates the instruction stream.
__device__ __forceinline__ void ew_divmod(long long idx, int d, long long& q, int& r) {
  if (idx < (1ll << 31)) {
    const unsigned u = (unsigned)idx, ud = (unsigned)d;
    const unsigned uq = u / ud;
    q = uq;
    r = (int)(u - uq * ud);
  } else {
    q = idx / d;
    r = (int)(idx - q * d);
  }
}

// Scratch of the column-reduce finalize kernels.  One set per device context: the kernels that
// use it are enqueued on ONE stream by a plan (they never overlap each other).
static __device__ double g_fin_scratch[kFinMaxColBlocks * kFinMaxSlices * 2 * 32];
static __device__ unsigned g_fin_counters[kFinMaxColBlocks];

// Deterministic sum over the rows of a block-partial buffer.  grid = (ceil(ncols/32), S),
// block = (32 columns, 32 row lanes).  Accumulator k of column c in row r lives at
// partial[r*rstride + k*koff + c].  Slice s sums rows s*32+ty, s*32+ty+32*S, ... in fp64; the
// last slice block to arrive (atomic ticket) adds the S slice sums in slice order and calls
// epi(c, sums) for its 32 columns -- the summation order is fixed, so results are bit-reproducible.
template <int NACC, class Epi>
__device__ __forceinline__ void ew_rows_reduce(const float* partial, int rows, long long rstride,
                                               long long koff, int ncols, Epi epi) {
  __shared__ double sh[NACC][32][33];
  __shared__ int is_last;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int c = blockIdx.x * 32 + tx;
  const int S = gridDim.y, s = blockIdx.y;
  double acc[NACC];
#pragma unroll
  for (int k = 0; k < NACC; ++k) acc[k] = 0.0;
  if (c < ncols && partial != nullptr) {
    // the loads of a batch of rows are issued together (one L2 round trip per 8 rows, not per row);
    // the fp64 additions keep the row order, so the result does not depend on the batching
    const int step = 32 * S;
    for (int r = s * 32 + ty; r < rows; r += 8 * step) {
      float v[8][NACC];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int rr = r + u * step;
#pragma unroll
        for (int k = 0; k < NACC; ++k)
          v[u][k] = rr < rows ? __ldg(partial + (size_t)rr * rstride + k * koff + c) : 0.0f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u)
#pragma unroll
        for (int k = 0; k < NACC; ++k) acc[k] += (double)v[u][k];
    }
  }
#pragma unroll
  for (int k = 0; k < NACC; ++k) sh[k][ty][tx] = acc[k];
  __syncthreads();
  if (ty < NACC) {
    double v = 0.0;
    for (int l = 0; l < 32; ++l) v += sh[ty][l][tx];
    sh[ty][0][tx] = v;
    if (S > 1) g_fin_scratch[(((size_t)blockIdx.x * S + s) * NACC + ty) * 32 + tx] = v;
  }
  if (S > 1) {
    __threadfence();
    __syncthreads();
    if (tx == 0 && ty == 0) is_last = (atomicAdd(&g_fin_counters[blockIdx.x], 1u) == (unsigned)(S - 1));
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    if (ty < NACC) {
      double v = 0.0;
      for (int q = 0; q < S; ++q)
        v += *((volatile double*)&g_fin_scratch[(((size_t)blockIdx.x * S + q) * NACC + ty) * 32 + tx]);
      sh[ty][0][tx] = v;
    }
    if (tx == 0 && ty == 0) g_fin_counters[blockIdx.x] = 0;
  }
  __syncthreads();
  if (ty == 0 && c < ncols) {
    double v[NACC];
#pragma unroll
    for (int k = 0; k < NACC; ++k) v[k] = sh[k][0][tx];
    epi(c, v);
  }
}

// Block-level per-channel reduction.  Thread t owns channel vector t % cv and pixel lane
// t / cv; `body(p, c0, acc)` adds pixel p's contribution for channels [c0, c0+8) into
// acc[NACC][8].  Row `blockIdx.x` of `partial` ([gridDim.x][NACC][C]) receives the block sums.
// Batched form: `fetch(p, c0)` only loads (returns the raw operands of one pixel), `consume(raw, c0, acc)`
// does the arithmetic.  NB pixels are fetched before the first is consumed, so a thread keeps NB x (32-64) bytes
// in flight: with one pixel per iteration the BatchNorm-backward reduce had ~24 KB in flight per SM and ran at
// 3.6-4.1 TB/s -- Little's law, not HBM.
// fixed != 0: `partial` is ONE int64 row [NACC][C] instead; the block adds its sums as fixed-point integers
// (x 2^40, clamped to +-2^62) with 64-bit atomics -- exact and order-independent (B2SEG_BWD_FIX_SCALE).
template <int NACC, int NB, class Fetch, class Consume>
__device__ __forceinline__ void ew_channel_reduce_batched(long long P0, long long P1, int C, float* partial,
                                                          float* red /* smem NACC*pl*C */, Fetch fetch,
                                                          Consume consume, bool fixed = false) {
  const int cv = C >> 3;
  const int pl = blockDim.x / cv > 0 ? blockDim.x / cv : 1;
  const int lc = threadIdx.x % cv, lp = threadIdx.x / cv;
  float acc[NACC][8];
#pragma unroll
  for (int k = 0; k < NACC; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[k][j] = 0.0f;
  if (lp < pl) {
    const long long step = (long long)gridDim.x * pl;
    long long p = P0 + (long long)blockIdx.x * pl + lp;
    for (; p + (NB - 1) * step < P1; p += NB * step) {
      decltype(fetch(p, 0)) raw[NB];
#pragma unroll
      for (int b = 0; b < NB; ++b) raw[b] = fetch(p + b * step, lc * 8);
#pragma unroll
      for (int b = 0; b < NB; ++b) consume(raw[b], lc * 8, acc);
    }
    for (; p < P1; p += step) consume(fetch(p, lc * 8), lc * 8, acc);
#pragma unroll
    for (int k = 0; k < NACC; ++k)
#pragma unroll
      for (int j = 0; j < 8; ++j) red[((size_t)k * pl + lp) * C + lc * 8 + j] = acc[k][j];
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < NACC * C; idx += blockDim.x) {
    const int k = idx / C, c = idx % C;
    float s = 0.0f;
    for (int l = 0; l < pl; ++l) s += red[((size_t)k * pl + l) * C + c];
    if (fixed) {
      const double v = fmin(fmax((double)s * B2SEG_BWD_FIX_SCALE, -4.611686018427388e18), 4.611686018427388e18);
      atomicAdd(reinterpret_cast<unsigned long long*>(partial) + (size_t)k * C + c,
                (unsigned long long)__double2ll_rn(v));
    } else {
      partial[((size_t)blockIdx.x * NACC + k) * C + c] = s;
    }
  }
}

template <int NACC, class Body>
__device__ __forceinline__ void ew_channel_reduce(long long P0, long long P1, int C, float* partial,
                                                  float* red /* smem NACC*pl*C */, Body body) {
  const int cv = C >> 3;
  const int pl = blockDim.x / cv > 0 ? blockDim.x / cv : 1;
  const int lc = threadIdx.x % cv, lp = threadIdx.x / cv;
  float acc[NACC][8];
#pragma unroll
  for (int k = 0; k < NACC; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[k][j] = 0.0f;
  if (lp < pl) {
#pragma unroll 2
    for (long long p = P0 + (long long)blockIdx.x * pl + lp; p < P1; p += (long long)gridDim.x * pl)
      body(p, lc * 8, acc);
#pragma unroll
    for (int k = 0; k < NACC; ++k)
#pragma unroll
      for (int j = 0; j < 8; ++j) red[((size_t)k * pl + lp) * C + lc * 8 + j] = acc[k][j];
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < NACC * C; idx += blockDim.x) {
    const int k = idx / C, c = idx % C;
    float s = 0.0f;
    for (int l = 0; l < pl; ++l) s += red[((size_t)k * pl + l) * C + c];
    partial[((size_t)blockIdx.x * NACC + k) * C + c] = s;
  }
}
#endif
