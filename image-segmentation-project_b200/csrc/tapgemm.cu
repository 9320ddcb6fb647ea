// Tap-gather implicit GEMM for convolutions on tcgen05 tensor cores.
//
//   out[pix, col] = sum_{tap} sum_{c} src_tap[pix + off_tap, c] * W[tap][col][c]
//
// M = 128 output pixels (a wt x ht x nt brick of the NHWC pixel grid),
// N = BLOCK_N output columns, K = taps x channels in chunks of 64.
// The A operand never exists as an im2col matrix: for every (tap, channel
// chunk) one rank-4 TMA box load with a shifted coordinate lands a 128 x 64
// bf16 K-major tile in shared memory (128B swizzle); out-of-bounds pixels are
// zero-filled by TMA, which is exactly the conv zero padding.  Weights arrive
// as [tap][col][c] boxes.  One elected thread issues tcgen05.mma with the fp32
// accumulator in TMEM; four epilogue warps read it back with tcgen05.ld, add
// bias, emit per-channel BatchNorm partial sums (sum, sum of squares) and store
// bf16 through a strided / pixel-shuffle address map.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM alloc + MMA
// issuer, warps 2..5 = epilogue (TMEM lane quadrant = warp_id % 4).
#include <new>
#include <stdlib.h>
#include <string.h>

#include "ops.cuh"

namespace {

// Division by a run-time constant as multiply-high + shift (dividends here are tile indices < 2^24).
struct FastDiv {
  uint32_t mul, shift, d;
};
static inline FastDiv make_fastdiv(uint32_t d) {
  FastDiv f;
  f.d = d;
  uint32_t l = 0;
  while ((1u << l) < d) ++l;
  f.shift = l;
  f.mul = (uint32_t)(((uint64_t)1 << 32) * (((uint64_t)1 << l) - d) / d + 1);
  return f;
}
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t fdiv(uint32_t n, const FastDiv& f) {
  return (__umulhi(n, f.mul) + n) >> f.shift;
}
#endif

struct TapSub {
  int ntaps;
  int8_t src[B2SEG_MAX_TAPS], dh[B2SEG_MAX_TAPS], dw[B2SEG_MAX_TAPS], wt[B2SEG_MAX_TAPS];
  long long out_off;
};

struct alignas(64) TapGemmParams {
  CUtensorMap tmA[4];
  CUtensorMap tmB;
  TapSub sub[B2SEG_MAX_SUB];
  int kchunks;
  int wt, ht, nt, tilesW, tilesH, mtiles;
  int on, oh, ow;
  bf16* out;
  long long osn, osh, osw, oqh, oqw;
  int cq_shift;  // 30 = no pixel-shuffle groups
  int ngemm;
  const float* bias;
  float* stats;
  int accumulate;
  int nsub, stat_rows;
  // TMA epilogue (nsub == 1): every epilogue warp stages its 32 pixels x 32 columns in shared memory and
  // stores / reduce-adds them as a pixel box; with a pixel shuffle there is one map per output phase
  CUtensorMap tmO[4];
  int tma_out, epi_off;
  int box_ch;            // channels per store: 32, or 16 when a 32-column chunk spans two shuffle phases
  int box_h, box_n;      // the warp's 32 pixels = wt x box_h x box_n
  FastDiv dv_mtiles, dv_w, dv_hw;  // / mtiles, / tilesW, / (tilesW * tilesH)
  long long* trace;      // -DTG_TRACE_ENABLE + B2SEG_TG_TRACE=file: CTA 0's per-tile clock64 timeline (debug)
};

#ifndef TG_TRACE_ENABLE
#define TG_TRACE(k, tileidx) do {} while (0)
#else
#define TG_TRACE(k, tileidx)                                                                 \
  do {                                                                                       \
    if (p.trace && blockIdx.x == 0 && lane == 0 && (tileidx) < 64) p.trace[(tileidx) * 8 + (k)] = clock64(); \
  } while (0)
#endif

constexpr int kThreadsTG = 192;
constexpr int A_BYTES = 128 * 128;  // 128 pixels x 64 bf16

template <int BLOCK_N, int NSTAGES>
struct TGSmem {
  static constexpr int B_BYTES = BLOCK_N * 128;
  static constexpr int STAGE = A_BYTES + B_BYTES;
  static constexpr int BAR_OFF = NSTAGES * STAGE;
  // full[NS], empty[NS], tmem_full[2], tmem_empty[2], tmem slot
  // [4 warps][2][BLOCK_N] f32 per-tile sums + [2][BLOCK_N] f32 running totals
  static constexpr int STAT_OFF = BAR_OFF + (2 * NSTAGES + 4) * 8 + 16;
  // TMA-epilogue staging: 2 KB slots per epilogue warp; two slots where two CTAs still fit an SM with them
  static constexpr int EPI_SLOTS = BLOCK_N == 32 ? 2 : 1;
  static constexpr int EPI_OFF = (STAT_OFF + 5 * 2 * BLOCK_N * 4 + 1023) / 1024 * 1024;
  static constexpr int TOTAL = EPI_OFF + 4 * EPI_SLOTS * 2048 + 1024 /*align slack*/;
  static_assert(2 * (TOTAL + 1024) <= 233472, "two CTAs per SM");
};

// A CTA keeps ONE column tile (blockIdx.x % ntiles; neighbouring CTAs share the A bricks through
// L2) and walks the (pixel tile, sub-problem) pairs  g, g + groups, ...  of its group g.
struct TGItem {
  int mtile, sub;
};
__device__ __forceinline__ TGItem tg_item(int j, int mtiles) {
  TGItem it;
  it.mtile = j % mtiles;
  it.sub = j / mtiles;
  return it;
}

// Persistent kernel: every CTA walks the work items blockIdx.x, blockIdx.x + gridDim.x, ...
// The shared-memory ring never drains between items, and the fp32 accumulator is double
// buffered in TMEM (2 x BLOCK_N columns), so the TMA producer and the MMA issuer run ahead
// into the next item while the four epilogue warps drain the previous one.
template <int BLOCK_N, int NSTAGES>
__global__ void __launch_bounds__(kThreadsTG, 2)
    tapgemm_kernel(const __grid_constant__ TapGemmParams p) {
  pdl_trigger();
  using L = TGSmem<BLOCK_N, NSTAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* empty_bar = full_bar + NSTAGES;
  uint64_t* tmem_full = empty_bar + NSTAGES;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;        // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = warp_id_uniform(), lane = threadIdx.x & 31;
  const int ntiles = p.ngemm / BLOCK_N;
  const int ntile = blockIdx.x % ntiles, grp = blockIdx.x / ntiles, groups = gridDim.x / ntiles;
  const int njobs = p.mtiles * p.nsub;
  const int col0 = ntile * BLOCK_N;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < NSTAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 4);  // one arrive per epilogue warp
    }
    fence_barrier_init();
    tma_prefetch_desc(&p.tmB);
    tma_prefetch_desc(&p.tmA[0]);
    if (p.tma_out) tma_prefetch_desc(&p.tmO[0]);
  }
  if (warp == 1) tmem_alloc<2 * BLOCK_N>(tmem_slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // barriers, TMEM and descriptors are set up; now wait for the producer grid

  if (warp == 0) {
    // ===================== TMA producer (converged warp, elected issue) =====================
    {
      uint32_t s = 0, ph = 1;  // ring slot and the parity a free slot shows
      uint32_t ptile = 0;
      for (int j = grp; j < njobs; j += groups, ++ptile) {
        const TGItem w = tg_item(j, p.mtiles);
        const TapSub& sub = p.sub[w.sub];
        const int tw = w.mtile % p.tilesW, th = (w.mtile / p.tilesW) % p.tilesH,
                  tn = w.mtile / (p.tilesW * p.tilesH);
        const int w0 = tw * p.wt, h0 = th * p.ht, n0 = tn * p.nt;
        for (int t = 0; t < sub.ntaps; ++t) {
          const CUtensorMap* tmA = &p.tmA[sub.src[t]];
          const int cw = w0 + sub.dw[t], ch = h0 + sub.dh[t], wtap = sub.wt[t];
          for (int kc = 0; kc < p.kchunks; ++kc) {
            mbar_wait(&empty_bar[s], ph);
            if (elect_one_sync()) {
              mbar_expect_tx(&full_bar[s], L::STAGE);
              uint8_t* sa = smem + s * L::STAGE;
              tma_load_4d(sa, tmA, &full_bar[s], kc * 64, cw, ch, n0);
              tma_load_3d(sa + A_BYTES, &p.tmB, &full_bar[s], kc * 64, col0, wtap);
            }
            __syncwarp();
            if (t == 0 && kc == 0) TG_TRACE(0, ptile);
            if (++s == NSTAGES) { s = 0; ph ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (converged warp, elected issue) =====================
    {
      constexpr uint32_t idesc = umma_idesc_bf16(128, BLOCK_N, 0, 0);
      // descriptors of ring slot 0; slot s adds s * STAGE / 16 to the 14-bit address field
      const uint64_t adesc0 = umma_smem_desc(smem_u32(smem), 16, 1024);
      const uint64_t bdesc0 = umma_smem_desc(smem_u32(smem) + A_BYTES, 16, 1024);
      uint32_t s = 0, ph = 0, tile = 0;
      for (int j = grp; j < njobs; j += groups, ++tile) {
        const TGItem w = tg_item(j, p.mtiles);
        const int iters = p.sub[w.sub].ntaps * p.kchunks;
        const uint32_t acc = tile & 1u, aph = (tile >> 1) & 1u;
        mbar_wait(&tmem_empty[acc], aph ^ 1u);  // epilogue has drained this accumulator
        tcgen05_fence_after();
        TG_TRACE(6, tile);
        const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
        for (int i = 0; i < iters; ++i) {
          mbar_wait(&full_bar[s], ph);
          tcgen05_fence_after();
          if (i == 0) TG_TRACE(1, tile);
          const uint64_t so = (uint64_t)(s * (uint32_t)(L::STAGE >> 4));
          const uint64_t adesc = adesc0 + so, bdesc = bdesc0 + so;
          if (elect_one_sync()) {
            umma_bf16_ss(tmem_d, adesc, bdesc, idesc, i != 0 ? 1u : 0u);
#pragma unroll
            for (int k = 1; k < 4; ++k)  // 4 x (K = 16 bf16 = 32 B) per 64-channel chunk
              umma_bf16_ss(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, 1u);
            tcgen05_commit(&empty_bar[s]);  // smem slot free once these MMAs retire
            if (i == iters - 1) tcgen05_commit(&tmem_full[acc]);
          }
          __syncwarp();
          if (++s == NSTAGES) { s = 0; ph ^= 1u; }
        }
        TG_TRACE(2, tile);
      }
    }
  } else {
    // ===================== epilogue =====================
    const int wq = warp & 3;  // TMEM lane quadrant this warp may access
    const int r = wq * 32 + lane;
    const int hw = p.ht * p.wt;
    const int ln = r / hw, lh = (r % hw) / p.wt, lw = r % p.wt;
    const int cq_mask = (1 << p.cq_shift) - 1;
    float* stat_sm = reinterpret_cast<float*>(smem + L::STAT_OFF) + wq * 2 * BLOCK_N;
    // running (sum, sumsq) of this CTA over all its tiles: entry e is owned by one thread
    float* run_sm = reinterpret_cast<float*>(smem + L::STAT_OFF) + 8 * BLOCK_N;
    // BatchNorm partial sums: for BLOCK_N == 32 they stay in registers across all tiles of the CTA and the
    // cross-lane reduction runs once after the loop (conv3x3.cu); the N = 128 tiles (deep-K layers, where
    // the epilogue is not the bottleneck) keep the per-tile fold through shared memory.  The per-tile fold
    // cost the thin 1x1 layers ~1500 of their ~2300 epilogue cycles per tile (profiles/r02_tapgemm_timeline.txt).
    constexpr bool REGSTATS = BLOCK_N == 32;  // 64 accumulators per thread at N = 64 spill (ptxas: 444 B)
    constexpr int RCH = REGSTATS ? BLOCK_N / 32 : 1;
    float rs1[RCH][32], rs2[RCH][32];
    if (REGSTATS) {
#pragma unroll
      for (int c = 0; c < RCH; ++c)
#pragma unroll
        for (int jx = 0; jx < 32; ++jx) rs1[c][jx] = rs2[c][jx] = 0.0f;
    }
    if (p.stats)
      for (int e = threadIdx.x - 64; e < 2 * BLOCK_N; e += 128) run_sm[e] = 0.0f;
    uint32_t tile = 0, nstore = 0;
    uint8_t* epi_sm = smem + L::EPI_OFF + wq * (L::EPI_SLOTS * 2048);  // this warp's 2 KB staging slots
    const int r0 = wq * 32, ln0 = r0 / hw, lh0 = (r0 % hw) / p.wt;  // first pixel of this warp's box
    for (int j = grp; j < njobs; j += groups, ++tile) {
      TGItem w;
      w.sub = (int)fdiv((uint32_t)j, p.dv_mtiles);
      w.mtile = j - w.sub * p.mtiles;
      const TapSub& sub = p.sub[w.sub];
      const int tn = (int)fdiv((uint32_t)w.mtile, p.dv_hw);
      const int rem = w.mtile - tn * (p.tilesW * p.tilesH);
      const int th = (int)fdiv((uint32_t)rem, p.dv_w), tw = rem - th * p.tilesW;
      const int pn = tn * p.nt + ln, phh = th * p.ht + lh, pw = tw * p.wt + lw;
      const bool row_ok = pn < p.on && phh < p.oh && pw < p.ow;
      const long long row_off =
          sub.out_off + (long long)pn * p.osn + (long long)phh * p.osh + (long long)pw * p.osw;
      const uint32_t acc = tile & 1u, aph = (tile >> 1) & 1u;
      // accumulate mode: fetch the old output values BEFORE waiting for the accumulator, so
      // their (uncoalesced, one pixel row per lane) latency hides behind the MMAs
      uint4 old[BLOCK_N / 32][4];
      if (p.accumulate && row_ok && !p.tma_out) {
#pragma unroll
        for (int c = 0; c < BLOCK_N / 32; ++c)
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int col = col0 + c * 32 + g * 8;
            const int q = col >> p.cq_shift;
            old[c][g] = *reinterpret_cast<const uint4*>(
                p.out + row_off + (long long)(q >> 1) * p.oqh + (long long)(q & 1) * p.oqw + (col & cq_mask));
          }
      }
      mbar_wait(&tmem_full[acc], aph);
      tcgen05_fence_after();
      if (warp == 2) TG_TRACE(3, tile);
      const uint32_t tmem_d = tmem_base + acc * BLOCK_N + ((uint32_t)(wq * 32) << 16);
#pragma unroll
      for (int c = 0; c < BLOCK_N / 32; ++c) {
        float v[32];
        tmem_ld_32x32b_x32(tmem_d + (uint32_t)(c * 32), v);
        if (warp == 2 && c == 0) TG_TRACE(4, tile);
        if (c == BLOCK_N / 32 - 1) {  // accumulator fully read: hand it back to the MMA issuer
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        }
        const int colc = col0 + c * 32;
        if (p.bias) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += __ldg(p.bias + ((colc + j) & cq_mask));
        }
        if (p.tma_out) {
          // stage 32 pixels x 32 columns and let TMA write (or reduce-add) the pixel box: the warp's
          // scattered 16-byte global accesses cost 32 L1 wavefronts per instruction (conv3x3.cu)
          uint8_t* sb = epi_sm + ((nstore & (uint32_t)(L::EPI_SLOTS - 1)) << 11);
          if (lane == 0) bulk_wait_group_read<L::EPI_SLOTS - 1>();  // the store that last used this slot has read it
          __syncwarp();
          uint4 pk[4];
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            pk[g].x = f2_to_bf2(v[g * 8 + 0], v[g * 8 + 1]);
            pk[g].y = f2_to_bf2(v[g * 8 + 2], v[g * 8 + 3]);
            pk[g].z = f2_to_bf2(v[g * 8 + 4], v[g * 8 + 5]);
            pk[g].w = f2_to_bf2(v[g * 8 + 6], v[g * 8 + 7]);
          }
          if (p.box_ch == 32) {  // 64 B rows, 64 B swizzle
            const int sw = (lane >> 1) & 3;
#pragma unroll
            for (int g = 0; g < 4; ++g) sts_u4(sb + lane * 64 + ((g ^ sw) << 4), pk[g]);
          } else {               // two 16-channel boxes of 32 B rows, no swizzle
#pragma unroll
            for (int g = 0; g < 4; ++g) sts_u4(sb + (g >> 1) * 1024 + lane * 32 + ((g & 1) << 4), pk[g]);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            const int cw = tw * p.wt, chh = th * p.ht + lh0, cn = tn * p.nt + ln0;
            for (int hh = 0; hh * p.box_ch < 32; ++hh) {
              const int col = colc + hh * p.box_ch;
              const int q = col >> p.cq_shift;
              const void* tm = &p.tmO[q & 3];
              const uint8_t* src = sb + hh * 1024;
              if (p.accumulate) tma_reduce_add_4d(tm, src, col & cq_mask, cw, chh, cn);
              else tma_store_4d(tm, src, col & cq_mask, cw, chh, cn);
            }
            bulk_commit_group();
          }
          ++nstore;
        } else if (row_ok) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int col = colc + g * 8;
            const int q = col >> p.cq_shift;
            bf16* dst = p.out + row_off + (long long)(q >> 1) * p.oqh + (long long)(q & 1) * p.oqw +
                        (col & cq_mask);
            float o[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o[j] = v[g * 8 + j];
            if (p.accumulate) {
              const uint4 od = old[c][g];
              const float2 a0 = bf2_to_f2(od.x), a1 = bf2_to_f2(od.y), a2 = bf2_to_f2(od.z),
                           a3 = bf2_to_f2(od.w);
              o[0] += a0.x; o[1] += a0.y; o[2] += a1.x; o[3] += a1.y;
              o[4] += a2.x; o[5] += a2.y; o[6] += a3.x; o[7] += a3.y;
            }
            uint4 pk;
            pk.x = f2_to_bf2(o[0], o[1]);
            pk.y = f2_to_bf2(o[2], o[3]);
            pk.z = f2_to_bf2(o[4], o[5]);
            pk.w = f2_to_bf2(o[6], o[7]);
            *reinterpret_cast<uint4*>(dst) = pk;
          }
        }
        if (REGSTATS) {
          if (p.stats && row_ok) {
#pragma unroll
            for (int jx = 0; jx < 32; ++jx) {
              rs1[c < RCH ? c : 0][jx] += v[jx];
              rs2[c < RCH ? c : 0][jx] = fmaf(v[jx], v[jx], rs2[c < RCH ? c : 0][jx]);
            }
          }
        } else if (p.stats) {
          float s1[32], s2[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float x = row_ok ? v[j] : 0.0f;
            s1[j] = x;
            s2[j] = x * x;
          }
          const float a = warp_transpose_sum32(s1, lane);
          const float b = warp_transpose_sum32(s2, lane);
          stat_sm[c * 32 + lane] = a;
          stat_sm[BLOCK_N + c * 32 + lane] = b;
        }
      }
      if (!REGSTATS && p.stats) {  // fold the four warps' sums into the CTA's running totals
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const float* sm0 = reinterpret_cast<const float*>(smem + L::STAT_OFF);
        for (int e = threadIdx.x - 64; e < 2 * BLOCK_N; e += 128)
          run_sm[e] += (sm0[e] + sm0[2 * BLOCK_N + e]) + (sm0[4 * BLOCK_N + e] + sm0[6 * BLOCK_N + e]);
        asm volatile("bar.sync 1, 128;" ::: "memory");  // stat_sm is rewritten by the next item
      }
      if (warp == 2) TG_TRACE(5, tile);
    }
    if (p.tma_out && lane == 0) bulk_wait_group<0>();  // this warp's stores have completed
    if (REGSTATS && p.stats) {  // one transposing reduction per warp and chunk, then the four warps' rows
#pragma unroll
      for (int c = 0; c < RCH; ++c) {
        const float a = warp_transpose_sum32(rs1[c], lane);
        const float b = warp_transpose_sum32(rs2[c], lane);
        stat_sm[c * 32 + lane] = a;
        stat_sm[BLOCK_N + c * 32 + lane] = b;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      const float* sm0 = reinterpret_cast<const float*>(smem + L::STAT_OFF);
      for (int e = threadIdx.x - 64; e < 2 * BLOCK_N; e += 128)
        run_sm[e] = (sm0[e] + sm0[2 * BLOCK_N + e]) + (sm0[4 * BLOCK_N + e] + sm0[6 * BLOCK_N + e]);
    }
    if (p.stats) {  // one (sum, sumsq) row per CTA group; rows no group owns are zero-filled
      for (int e = threadIdx.x - 64; e < 2 * BLOCK_N; e += 128) {
        const int k = e / BLOCK_N, cc = e % BLOCK_N;
        if (p.stat_rows == 0) {  // fixed-point accumulators (b2seg_tapgemm_desc::stats_fixed)
          atomicAdd(reinterpret_cast<unsigned long long*>(p.stats) + (size_t)k * p.ngemm + col0 + cc,
                    (unsigned long long)__float2ll_rn(run_sm[e] * (k ? B2SEG_STATS_FIX_SQ : B2SEG_STATS_FIX_SUM)));
          continue;
        }
        p.stats[((size_t)grp * 2 + k) * p.ngemm + col0 + cc] = run_sm[e];
        for (int rr = grp + groups; rr < p.stat_rows; rr += groups)
          p.stats[((size_t)rr * 2 + k) * p.ngemm + col0 + cc] = 0.0f;
      }
    }
    tcgen05_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc<2 * BLOCK_N>(tmem_base);
  }
}

struct TapGemmOp : b2seg_op {
  TapGemmParams p;
  dim3 grid;
  int block_n, stages;
  size_t smem;
  const char* name() const override { return "tapgemm"; }
  int run(cudaStream_t st) override {
#define TG_LAUNCH(BN, NS)                                                            \
  B2_CHECK_CUDA(b2_launch(tapgemm_kernel<BN, NS>, grid, dim3(kThreadsTG), (size_t)TGSmem<BN, NS>::TOTAL, st, p))
    if (block_n == 128) TG_LAUNCH(128, 3);
    else if (block_n == 64) TG_LAUNCH(64, 4);
    else TG_LAUNCH(32, 4);
#undef TG_LAUNCH
    B2_CHECK_CUDA(cudaGetLastError());
#ifdef TG_TRACE_ENABLE
    if (p.trace) {
      B2_CHECK_CUDA(cudaStreamSynchronize(st));
      long long h[64 * 8];
      B2_CHECK_CUDA(cudaMemcpy(h, p.trace, sizeof(h), cudaMemcpyDeviceToHost));
      if (FILE* f = fopen(getenv("B2SEG_TG_TRACE"), "a")) {
        fprintf(f, "# tapgemm BN=%d kchunks=%d mtiles=%d grid=%d nsub=%d tma_out=%d  columns: tile prodA mma_full mma_issued epi_full epi_ld epi_done mma_tmem_empty\n",
                block_n, p.kchunks, p.mtiles, grid.x, p.nsub, p.tma_out);
        const long long t0 = h[0];
        for (int i = 0; i < 64 && h[i * 8 + 1]; ++i) {
          fprintf(f, "%2d", i);
          for (int k = 0; k < 7; ++k) fprintf(f, " %8lld", h[i * 8 + k] ? h[i * 8 + k] - t0 : -1LL);
          fprintf(f, "\n");
        }
        fclose(f);
      }
      B2_CHECK_CUDA(cudaMemset(p.trace, 0, sizeof(h)));
    }
#endif
    return B2SEG_OK;
  }
};

template <int BN, int NS>
int set_smem_attr() {
  static bool done = false;
  if (!done) {
    B2_CHECK_CUDA(cudaFuncSetAttribute(tapgemm_kernel<BN, NS>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       TGSmem<BN, NS>::TOTAL));
    done = true;
  }
  return B2SEG_OK;
}

int ilog2_exact(int x) {
  int s = 0;
  while ((1 << s) < x) ++s;
  return (1 << s) == x ? s : -1;
}

}  // namespace

int b2_view_tmap(CUtensorMap* out, const b2seg_tensor4* v, int wt, int ht, int nt) {
  const uint64_t dims[4] = {(uint64_t)v->c, (uint64_t)v->w, (uint64_t)v->h, (uint64_t)v->n};
  const uint64_t strides[3] = {(uint64_t)v->sw * 2, (uint64_t)v->sh * 2, (uint64_t)v->sn * 2};
  const uint32_t box[4] = {64, (uint32_t)wt, (uint32_t)ht, (uint32_t)nt};
  return b2_encode_tmap_bf16(out, v->ptr, 4, dims, strides, box, 1);
}

extern "C" int b2seg_tapgemm_stat_rows(int on, int oh, int ow) {
  int wt, ht, nt;
  b2_pick_pixel_tile(on, oh, ow, &wt, &ht, &nt);
  const int tiles = ((ow + wt - 1) / wt) * ((oh + ht - 1) / ht) * ((on + nt - 1) / nt);
  const int cap = 2 * b2_num_sms();  // one row per CTA group of the persistent conv kernels
  return tiles < cap ? tiles : cap;
}

extern "C" int b2seg_tapgemm_create(const b2seg_tapgemm_desc* d, b2seg_op** out) {
  B2_REQUIRE(d && out, "tapgemm_create: null argument");
  B2_REQUIRE(d->nsrc >= 1 && d->nsrc <= 4, "tapgemm_create: nsrc %d", d->nsrc);
  B2_REQUIRE(d->nsub >= 1 && d->nsub <= B2SEG_MAX_SUB, "tapgemm_create: nsub %d", d->nsub);
  B2_REQUIRE(d->weights && d->out, "tapgemm_create: null weights/out");
  B2_REQUIRE(d->cin >= 8 && d->cin % 8 == 0, "tapgemm_create: cin %d must be a multiple of 8",
             d->cin);
  B2_REQUIRE(d->ngemm >= 32 && d->ngemm % 32 == 0,
             "tapgemm_create: ngemm %d must be a multiple of 32", d->ngemm);
  B2_REQUIRE(d->on > 0 && d->oh > 0 && d->ow > 0, "tapgemm_create: empty pixel grid");
  B2_REQUIRE(d->stats == nullptr || d->nsub == 1, "tapgemm_create: stats need nsub == 1");
  int cq_shift = 30;
  if (d->cq > 0) {
    cq_shift = ilog2_exact(d->cq);
    B2_REQUIRE(cq_shift >= 3, "tapgemm_create: cq %d must be a power of two >= 8", d->cq);
  }
  if (!getenv("B2SEG_NO_HALO")) {  // 3x3 stride-1: halo-reuse kernel (conv3x3.cu)
    b2seg_op* h = nullptr;
    int hrc = b2_halo3x3_try_create(d, &h);
    if (hrc != B2SEG_OK) return hrc;
    if (h) { *out = h; return B2SEG_OK; }
  }
  auto* op = new (std::nothrow) TapGemmOp();
  B2_REQUIRE(op, "tapgemm_create: out of host memory");
  TapGemmParams& p = op->p;
  memset(&p, 0, sizeof(p));
  b2_pick_pixel_tile(d->on, d->oh, d->ow, &p.wt, &p.ht, &p.nt);
  p.tilesW = (d->ow + p.wt - 1) / p.wt;
  p.tilesH = (d->oh + p.ht - 1) / p.ht;
  const int tilesN = (d->on + p.nt - 1) / p.nt;
  p.mtiles = p.tilesW * p.tilesH * tilesN;
  for (int i = 0; i < d->nsrc; ++i) {
    const b2seg_tensor4& v = d->src[i];
    if (v.c != d->cin || !v.ptr) {
      b2seg_set_error("tapgemm_create: src[%d] has %d channels, expected %d", i, v.c, d->cin);
      delete op;
      return B2SEG_EINVAL;
    }
    int rc = b2_view_tmap(&p.tmA[i], &v, p.wt, p.ht, p.nt);
    if (rc != B2SEG_OK) { delete op; return rc; }
  }
  for (int i = d->nsrc; i < 4; ++i) p.tmA[i] = p.tmA[0];

  const int nm = d->ngemm;
  int bn;
  // N = 128 tiles once they fill most of the SMs (B2SEG_TG_BN128_MIN = minimum tile count)
  long long bn128_min = b2_num_sms();
  if (const char* e = getenv("B2SEG_TG_BN128_MIN")) bn128_min = atoll(e);
  if (nm % 128 == 0 && (long long)p.mtiles * (nm / 128) >= bn128_min) bn = 128;
  else if (nm % 64 == 0) bn = 64;
  else bn = 32;
  op->block_n = bn;
  {
    const uint64_t dims[3] = {(uint64_t)d->cin, (uint64_t)d->ngemm, (uint64_t)d->wtaps};
    const uint64_t strides[2] = {(uint64_t)d->cin * 2, (uint64_t)d->cin * d->ngemm * 2};
    const uint32_t box[3] = {64, (uint32_t)bn, 1};
    int rc = b2_encode_tmap_bf16(&p.tmB, d->weights, 3, dims, strides, box, 1);
    if (rc != B2SEG_OK) { delete op; return rc; }
  }
  for (int s = 0; s < d->nsub; ++s) {
    const auto& ds = d->sub[s];
    if (ds.ntaps < 1 || ds.ntaps > B2SEG_MAX_TAPS) {
      b2seg_set_error("tapgemm_create: sub %d ntaps %d", s, ds.ntaps);
      delete op;
      return B2SEG_EINVAL;
    }
    p.sub[s].ntaps = ds.ntaps;
    p.sub[s].out_off = ds.out_offset;
    for (int t = 0; t < ds.ntaps; ++t) {
      if (ds.taps[t].src < 0 || ds.taps[t].src >= d->nsrc || ds.taps[t].wtap < 0 ||
          ds.taps[t].wtap >= d->wtaps) {
        b2seg_set_error("tapgemm_create: sub %d tap %d out of range", s, t);
        delete op;
        return B2SEG_EINVAL;
      }
      p.sub[s].src[t] = ds.taps[t].src;
      p.sub[s].dh[t] = ds.taps[t].dh;
      p.sub[s].dw[t] = ds.taps[t].dw;
      p.sub[s].wt[t] = ds.taps[t].wtap;
    }
  }
  p.kchunks = (d->cin + 63) / 64;
  p.on = d->on; p.oh = d->oh; p.ow = d->ow;
  p.out = (bf16*)d->out;
  p.osn = d->osn; p.osh = d->osh; p.osw = d->osw;
  p.oqh = d->oqh; p.oqw = d->oqw;
  p.cq_shift = cq_shift;
  p.ngemm = d->ngemm;
  p.bias = d->bias;
  p.stats = d->stats;
  p.accumulate = d->accumulate;
  p.nsub = d->nsub;
  p.stat_rows = d->stats_fixed ? 0 : b2seg_tapgemm_stat_rows(d->on, d->oh, d->ow);
  p.dv_mtiles = make_fastdiv((uint32_t)p.mtiles);
  p.dv_w = make_fastdiv((uint32_t)p.tilesW);
  p.dv_hw = make_fastdiv((uint32_t)(p.tilesW * p.tilesH));
  {  // TMA epilogue: one output map (no shuffle) or one per shuffle phase; not for phase-split outputs
    const char* e = getenv("B2SEG_TG_TMA_OUT");
    const bool want = !(e && e[0] == '0');
    const int cq = cq_shift < 30 ? (1 << cq_shift) : d->ngemm;
    p.box_ch = cq >= 32 ? 32 : 16;
    p.box_h = p.ht < 32 / p.wt ? p.ht : 32 / p.wt;
    p.box_n = 32 / (p.wt * p.box_h);
    p.tma_out = want && d->nsub == 1 && d->sub[0].out_offset == 0 && cq >= 16 && cq % p.box_ch == 0 &&
                (cq_shift == 30 || d->ngemm <= 4 * cq);
    if (p.tma_out) {
      const int nq = cq_shift < 30 ? d->ngemm / cq : 1;
      for (int q = 0; q < 4; ++q) {
        const int qq = q < nq ? q : 0;
        const bf16* base = (const bf16*)d->out + (long long)(qq >> 1) * d->oqh + (long long)(qq & 1) * d->oqw;
        const uint64_t dims[4] = {(uint64_t)cq, (uint64_t)d->ow, (uint64_t)d->oh, (uint64_t)d->on};
        const uint64_t strides[3] = {(uint64_t)d->osw * 2, (uint64_t)d->osh * 2, (uint64_t)d->osn * 2};
        const uint32_t box[4] = {(uint32_t)p.box_ch, (uint32_t)p.wt, (uint32_t)p.box_h, (uint32_t)p.box_n};
        int rc = b2_encode_tmap_bf16(&p.tmO[q], base, 4, dims, strides, box, p.box_ch == 32 ? 2 : 0);
        if (rc != B2SEG_OK) { p.tma_out = 0; break; }  // e.g. a stride that TMA cannot express: direct stores
      }
    }
  }
  {  // persistent grid: two CTAs per SM (96 KB of shared memory and <= 256 TMEM columns each),
     // a multiple of the column-tile count so that a CTA keeps one column tile
    const int ntiles = d->ngemm / bn;
    int groups = 2 * b2_num_sms() / ntiles;
    if (groups > p.mtiles * d->nsub) groups = p.mtiles * d->nsub;
    if (groups < 1) groups = 1;
    op->grid = dim3(groups * ntiles, 1, 1);
  }
#ifdef TG_TRACE_ENABLE
  if (const char* tr = getenv("B2SEG_TG_TRACE")) {
    if (tr[0] && cudaMalloc(&p.trace, 64 * 8 * 8) == cudaSuccess) cudaMemset(p.trace, 0, 64 * 8 * 8);
    else p.trace = nullptr;
  }
#endif
  int rc;
  if (bn == 128) rc = set_smem_attr<128, 3>();
  else if (bn == 64) rc = set_smem_attr<64, 4>();
  else rc = set_smem_attr<32, 4>();
  if (rc != B2SEG_OK) { delete op; return rc; }
  *out = op;
  return B2SEG_OK;
}

// ---- generic op / sequence plumbing ----------------------------------------------
extern "C" int b2seg_op_run(b2seg_op* op, void* stream) {
  B2_REQUIRE(op, "op_run: null op");
  return op->run((cudaStream_t)stream);
}
extern "C" int b2seg_op_destroy(b2seg_op* op) {
  delete op;
  return B2SEG_OK;
}
extern "C" int b2seg_op_launches(const b2seg_op* op) { return op ? op->launches() : 0; }
extern "C" int b2seg_seq_create(b2seg_seq** out) {
  B2_REQUIRE(out, "seq_create: null");
  *out = new (std::nothrow) b2seg_seq();
  B2_REQUIRE(*out, "seq_create: out of host memory");
  return B2SEG_OK;
}
extern "C" int b2seg_seq_append_lane(b2seg_seq* seq, b2seg_op* op, int lane) {
  B2_REQUIRE(seq && op, "seq_append: null");
  B2_REQUIRE(lane == 0 || lane == 1, "seq_append: lane %d (0 = chain, 1 = side branch)", lane);
  B2_REQUIRE(!seq->exec, "seq_append: sequence already captured");
  seq->ops.push_back(op);
  seq->lanes.push_back(lane);
  return B2SEG_OK;
}
extern "C" int b2seg_seq_append(b2seg_seq* seq, b2seg_op* op) {
  return b2seg_seq_append_lane(seq, op, 0);
}
extern "C" int b2seg_seq_join(b2seg_seq* seq) {
  B2_REQUIRE(seq, "seq_join: null");
  B2_REQUIRE(!seq->exec, "seq_join: sequence already captured");
  seq->join_at.push_back(seq->ops.size());
  return B2SEG_OK;
}
namespace {
bool graphs_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B2SEG_GRAPH");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
}
// Capture the op list into an executable graph.  Any failure (e.g. a driver that cannot capture
// programmatic-dependent launches) leaves the sequence on the direct-launch path.
bool side_lane_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B2SEG_SIDE");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
}
bool seq_capture_once(b2seg_seq* seq, bool use_lanes) {
  if (cudaStreamBeginCapture(seq->capture_stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  int rc = B2SEG_OK;
  bool side_used = false, forked = false;  // forked: side branch already waits on the latest chain op
  bool side_pending = false;  // side-branch work not yet joined
  for (size_t i = 0; i < seq->ops.size() && rc == B2SEG_OK; ++i) {
    if (side_pending)
      for (size_t j : seq->join_at)
        if (j == i) {
          if (cudaEventRecord(seq->join_event, seq->side_stream) != cudaSuccess ||
              cudaStreamWaitEvent(seq->capture_stream, seq->join_event, 0) != cudaSuccess)
            rc = B2SEG_ECUDA;
          side_pending = forked = false;
          break;
        }
    if (rc != B2SEG_OK) break;
    if (use_lanes && seq->lanes[i] == 1) {
      side_pending = true;
      if (!forked) {
        if (cudaEventRecord(seq->fork_event, seq->capture_stream) != cudaSuccess ||
            cudaStreamWaitEvent(seq->side_stream, seq->fork_event, 0) != cudaSuccess) {
          rc = B2SEG_ECUDA;
          break;
        }
        forked = side_used = true;
      }
      rc = seq->ops[i]->run(seq->side_stream);
    } else {
      rc = seq->ops[i]->run(seq->capture_stream);
      forked = false;
    }
  }
  if (side_used && side_pending && rc == B2SEG_OK) {
    if (cudaEventRecord(seq->join_event, seq->side_stream) != cudaSuccess ||
        cudaStreamWaitEvent(seq->capture_stream, seq->join_event, 0) != cudaSuccess)
      rc = B2SEG_ECUDA;
  }
  cudaGraph_t g = nullptr;
  cudaError_t e = cudaStreamEndCapture(seq->capture_stream, &g);
  if (rc != B2SEG_OK || e != cudaSuccess || !g) {
    if (g) cudaGraphDestroy(g);
    cudaGetLastError();
    return false;
  }
  cudaGraphExec_t ex = nullptr;
  if (cudaGraphInstantiate(&ex, g, 0) != cudaSuccess) {
    cudaGraphDestroy(g);
    cudaGetLastError();
    return false;
  }
  seq->graph = g;
  seq->exec = ex;
  return true;
}
// Capture the op list into an executable graph.  Any failure (e.g. a driver that cannot capture
// programmatic-dependent launches) leaves the sequence on the direct-launch path.
void seq_try_capture(b2seg_seq* seq) {
  seq->capture_failed = true;  // cleared on success
  if (cudaStreamCreateWithFlags(&seq->capture_stream, cudaStreamNonBlocking) != cudaSuccess) return;
  bool lanes = false;
  if (side_lane_enabled())
    for (int l : seq->lanes) lanes = lanes || l == 1;
  if (lanes) {
    lanes = cudaStreamCreateWithFlags(&seq->side_stream, cudaStreamNonBlocking) == cudaSuccess &&
            cudaEventCreateWithFlags(&seq->fork_event, cudaEventDisableTiming) == cudaSuccess &&
            cudaEventCreateWithFlags(&seq->join_event, cudaEventDisableTiming) == cudaSuccess;
    if (!lanes) cudaGetLastError();
  }
  if (lanes && seq_capture_once(seq, true)) {
    seq->capture_failed = false;
    return;
  }
  if (seq_capture_once(seq, false)) seq->capture_failed = false;
}
}  // namespace

extern "C" int b2seg_seq_run(b2seg_seq* seq, void* stream) {
  B2_REQUIRE(seq, "seq_run: null");
  if (graphs_enabled() && seq->ops.size() >= 4) {
    if (!seq->exec && !seq->capture_failed) seq_try_capture(seq);
    if (seq->exec) {
      B2_CHECK_CUDA(cudaGraphLaunch(seq->exec, (cudaStream_t)stream));
      return B2SEG_OK;
    }
  }
  for (b2seg_op* op : seq->ops) {
    int rc = op->run((cudaStream_t)stream);
    if (rc != B2SEG_OK) return rc;
  }
  return B2SEG_OK;
}
extern "C" int b2seg_seq_is_graph(const b2seg_seq* seq) { return seq && seq->exec ? 1 : 0; }
extern "C" int b2seg_seq_run_timed(b2seg_seq* seq, void* stream, float* ms, int n_ops) {
  B2_REQUIRE(seq && ms, "seq_run_timed: null");
  B2_REQUIRE(n_ops == (int)seq->ops.size(), "seq_run_timed: n_ops %d != %d", n_ops,
             (int)seq->ops.size());
  cudaStream_t st = (cudaStream_t)stream;
  std::vector<cudaEvent_t> ev(seq->ops.size() + 1);
  for (auto& e : ev) B2_CHECK_CUDA(cudaEventCreate(&e));
  int rc = B2SEG_OK;
  B2_CHECK_CUDA(cudaEventRecord(ev[0], st));
  for (size_t i = 0; i < seq->ops.size() && rc == B2SEG_OK; ++i) {
    rc = seq->ops[i]->run(st);
    cudaEventRecord(ev[i + 1], st);
  }
  cudaError_t err = cudaStreamSynchronize(st);
  if (rc == B2SEG_OK && err == cudaSuccess)
    for (size_t i = 0; i < seq->ops.size(); ++i) cudaEventElapsedTime(&ms[i], ev[i], ev[i + 1]);
  for (auto& e : ev) cudaEventDestroy(e);
  B2_CHECK_CUDA(err);
  return rc;
}
extern "C" int b2seg_seq_launches(const b2seg_seq* seq) {
  int n = 0;
  if (seq)
    for (const b2seg_op* op : seq->ops) n += op->launches();
  return n;
}
extern "C" int b2seg_seq_destroy(b2seg_seq* seq) {
  if (seq) {
    if (seq->exec) cudaGraphExecDestroy(seq->exec);
    if (seq->graph) cudaGraphDestroy(seq->graph);
    if (seq->capture_stream) cudaStreamDestroy(seq->capture_stream);
    if (seq->side_stream) cudaStreamDestroy(seq->side_stream);
    if (seq->fork_event) cudaEventDestroy(seq->fork_event);
    if (seq->join_event) cudaEventDestroy(seq->join_event);
  }
  delete seq;
  return B2SEG_OK;
}
