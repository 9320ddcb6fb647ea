// 3x3 stride-1 convolution (forward and data gradient) as a halo-reuse implicit GEMM on tcgen05.
//
// The generic tap-gather kernel (tapgemm.cu) lands one 128-pixel A brick per filter tap: nine
// TMA loads of (almost) the same pixels per 64-channel chunk.  Here ONE rank-4 TMA box brings the
// brick with its 1-pixel halo -- (ht+2) x (wt+2) pixels x 64 channels, zero-filled outside the
// image, which is the conv padding -- and all nine taps are tcgen05.mma instructions whose A
// descriptors point INTO that halo tile: tap (dh,dw) starts (dh+1)*(wt+2) + (dw+1) rows of 128 B
// further, and the 8-row group stride (SBO) is the halo pitch (wt+2)*128 B with wt = 8.  The
// tensor core applies the 128B swizzle on absolute shared-memory address bits (measured:
// profiles/r01_umma_swizzle_probe.txt), so unaligned starts and a 1280 B group stride read
// exactly what TMA wrote.  A traffic drops from 9 x 16 KB to 22.5 KB per chunk.
//
// Weights: when the [9 taps][BLOCK_N][Cin] slice a CTA needs fits beside the A ring it is loaded
// ONCE and stays resident while the persistent CTA walks its pixel tiles; otherwise 16 KB tap
// slabs stream through a second ring.  Accumulators are double buffered in TMEM; eight epilogue
// warps (two per TMEM lane quadrant) drain one tile while the next one is being multiplied.
//
// Warp roles (352 threads): warp 0 = TMA producer, warps 1 and 2 = MMA issuers (even / odd pixel tiles,
// one TMEM accumulator each; warp 1 owns the TMEM allocation), warps 3..10 = epilogue.
#include <new>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "ops.cuh"

namespace {

constexpr int kThreadsH = 352;  // producer, two MMA issuers, eight epilogue warps
constexpr int kEpiWarp0 = 3;
constexpr int kMaxAStages = 8;
constexpr int kMaxBSlots = 18;
constexpr int kAcc = 2;  // TMEM accumulator stages (4 measured no faster; 2 leaves TMEM for a PDL-overlapped successor)

struct alignas(64) HaloParams {
  CUtensorMap tmA;  // box (64 ch, wt+2, ht+2, 1)
  CUtensorMap tmB;  // box (64 ch, BLOCK_N, 1 tap)
  int8_t wtap[9];    // weight tap index of raster tap r = (dh+1)*3 + (dw+1)
  int kchunks;
  int ksteps_last;   // K = 16 steps of the last 64-channel chunk (cin % 64 != 0: TMA zero-fills the rest)
  int wt, ht, tilesW, tilesH, mtiles, ntiles;
  int on, oh, ow;
  bf16* out;
  long long osn, osh, osw;
  int ngemm;
  const float* bias;
  float* stats;
  int accumulate;
  int resident, a_stages, b_slots, stat_rows;
  int probe;  // B2SEG_PROBE bits (perf experiments only): 1 no output stores, 2 no stats, 4 no MMAs, 16 no A traffic after the first tile, 32 no TMEM reads
  int a_stage_bytes, b_off, bar_off, stat_off;
  CUtensorMap tmO;   // output view, box (32 ch, 8, 4, 1), 64 B swizzle: one epilogue warp's 32 pixels x 32 channels
  int issuers;       // 2: MMA warps 1 and 2 alternate pixel tiles; 1: warp 1 issues every tile
  int tma_out;       // epilogue stages through shared memory and stores / reduce-adds with TMA
  int epi_off;
  long long* trace;  // B2SEG_HALO_TRACE=1: CTA 0 records clock64 per tile and role (debug only)
};

#ifndef HALO_TRACE_ENABLE  // the trace points cost ~25 % even when off at run time: compile them in only on demand
#define HALO_TRACE(k, tileidx) do {} while (0)
#else
#define HALO_TRACE(k, tileidx)                                                              \
  do {                                                                                     \
    if (p.trace && blockIdx.x == 0 && lane == 0 && (tileidx) < 64) p.trace[(tileidx) * 8 + (k)] = clock64(); \
  } while (0)
#endif

constexpr int kHaloPitch = 10;  // halo tile pitch in pixels: wt + 2 with wt = 8 (enforced by the creator)

// Taps [T0, T1) of one 64-channel chunk of a pixel tile (raster order) x KS K-steps into one accumulator,
// issued by the elected thread.  Every operand offset is an immediate added to the two chunk descriptors,
// which keeps the block at ~3 SASS instructions per MMA (2 x UIADD3.64 + UTCHMMA).  Keep ONE copy of this
// block per code path: when the same offsets appeared in several unrolled copies ptxas hoisted all 72
// descriptors into vector registers (~110 instructions ahead of the first MMA of every tile, R2UR pairs
// per MMA) and the tensor pipe idled behind the issuing thread.
template <int BLOCK_N, int KS, int T0, int T1>
__device__ __forceinline__ void halo_issue_taps(uint32_t tmem_d, uint64_t ad, uint64_t bd, uint32_t idesc,
                                                uint32_t acc0) {
#pragma unroll
  for (int t = T0; t < T1; ++t) {
#pragma unroll
    for (int k = 0; k < KS; ++k)
      umma_bf16_ss(tmem_d, ad + (uint64_t)(((t / 3) * kHaloPitch + (t % 3)) * 8 + 2 * k),
                   bd + (uint64_t)(t * ((BLOCK_N * 128) >> 4) + 2 * k), idesc, (t != T0 || k != 0) ? 1u : acc0);
  }
}
template <int BLOCK_N, int T0, int T1>
__device__ __forceinline__ void halo_issue_taps_ks(int ks, uint32_t tmem_d, uint64_t ad, uint64_t bd,
                                                   uint32_t idesc, uint32_t acc0) {
  if (ks == 4) halo_issue_taps<BLOCK_N, 4, T0, T1>(tmem_d, ad, bd, idesc, acc0);
  else if (ks == 2) halo_issue_taps<BLOCK_N, 2, T0, T1>(tmem_d, ad, bd, idesc, acc0);
  else if (ks == 1) halo_issue_taps<BLOCK_N, 1, T0, T1>(tmem_d, ad, bd, idesc, acc0);
  else halo_issue_taps<BLOCK_N, 3, T0, T1>(tmem_d, ad, bd, idesc, acc0);
}

template <int BLOCK_N, bool STATS>
__global__ void __launch_bounds__(kThreadsH, 1)
    halo3x3_kernel(const __grid_constant__ HaloParams p) {
  pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  constexpr int B_SLAB = BLOCK_N * 128;
  uint64_t* a_full = reinterpret_cast<uint64_t*>(smem + p.bar_off);
  uint64_t* a_empty = a_full + kMaxAStages;
  uint64_t* b_full = a_empty + kMaxAStages;
  uint64_t* b_empty = b_full + kMaxBSlots;
  uint64_t* tmem_full = b_empty + kMaxBSlots;  // [kAcc]
  uint64_t* tmem_empty = tmem_full + kAcc;     // [kAcc]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + kAcc);

  const int warp = warp_id_uniform(), lane = threadIdx.x & 31;
  // this CTA keeps one column tile (its weights may be resident) and strides over pixel tiles
  const int ntile = blockIdx.x % p.ntiles;
  const int mt0 = blockIdx.x / p.ntiles, mstep = gridDim.x / p.ntiles;
  const int col0 = ntile * BLOCK_N;
  constexpr int pitch = kHaloPitch;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < p.a_stages; ++i) {
      mbar_init(&a_full[i], 1);
      mbar_init(&a_empty[i], 1);
    }
    const int nb = p.resident ? p.kchunks : p.b_slots;  // resident weights: one barrier per 64-channel chunk
    for (int i = 0; i < nb; ++i) {
      mbar_init(&b_full[i], 1);
      mbar_init(&b_empty[i], 1);
    }
    for (int i = 0; i < kAcc; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 8);  // one arrive per epilogue warp
    }
    fence_barrier_init();
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
    if (p.tma_out) tma_prefetch_desc(&p.tmO);
  }
  if (warp == 1) tmem_alloc<kAcc * BLOCK_N>(tmem_slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Resident weights do not depend on the predecessor grid (the packed operands were written by a
  // kernel that completed before any kernel of this chain started): request them BEFORE the
  // dependency wait, one mbarrier per 64-channel chunk (nine tap slabs each).
  if (warp == 0 && p.resident) {
    for (int kc = 0; kc < p.kchunks; ++kc) {
      if (elect_one_sync()) {
        mbar_expect_tx(&b_full[kc], 9 * B_SLAB);
        for (int t = 0; t < 9; ++t)
          tma_load_3d(smem + p.b_off + (kc * 9 + t) * B_SLAB, &p.tmB, &b_full[kc], kc * 64, col0, p.wtap[t]);
      }
      __syncwarp();
    }
  }
  pdl_wait();  // barriers, TMEM and descriptors are set up; now wait for the producer grid

  // pixel-tile walk without divisions: (tw, th, n) advance by the decomposition of mstep
  const int tilesHW = p.tilesW * p.tilesH;
  const int d_n = mstep / tilesHW, d_th = (mstep % tilesHW) / p.tilesW, d_tw = mstep % p.tilesW;
  int tw = mt0 % p.tilesW, th = (mt0 / p.tilesW) % p.tilesH, tn = mt0 / tilesHW;
  auto next_tile = [&]() {
    tw += d_tw;
    if (tw >= p.tilesW) { tw -= p.tilesW; ++th; }
    th += d_th;
    if (th >= p.tilesH) { th -= p.tilesH; ++tn; }
    tn += d_n;
  };

  // Streaming weights: every CTA would walk the (chunk, tap row) slabs in the same order and ask the
  // same few L2 lines for the same 48 KB at the same moment; each CTA group starts at its own position
  // of the cycle instead (the accumulation order of a tile depends on its CTA only: still reproducible).
  const int rot = p.resident ? 0 : mt0;
  const int kc0 = rot % p.kchunks, g0 = (rot / p.kchunks) % 3;

  if (warp == 0) {
    // ===================== TMA producer (converged warp, elected issue) =====================
    {
      uint32_t s = 0, ph = 1, bs = 0, bph = 1;  // ring slots and the parity a free slot shows
      bool first = true;
      uint32_t ptile = 0;
      const uint32_t halo_bytes = (uint32_t)(pitch * (p.ht + 2) * 128);
      for (int mt = mt0; mt < p.mtiles; mt += mstep, first = false, next_tile(), ++ptile) {
        const int n0 = tn;
        const int w0 = tw * p.wt - 1, h0 = th * p.ht - 1;
        for (int kq = 0; kq < p.kchunks; ++kq) {
          const int kc = kq + kc0 < p.kchunks ? kq + kc0 : kq + kc0 - p.kchunks;
          mbar_wait(&a_empty[s], ph);
          if (elect_one_sync()) {
            if ((p.probe & 16) && !first) {
              mbar_arrive(&a_full[s]);  // probe: no A traffic after the first tile
            } else {
              mbar_expect_tx(&a_full[s], halo_bytes);
              tma_load_4d(smem + s * p.a_stage_bytes, &p.tmA, &a_full[s], kc * 64, w0, h0, n0);
            }
          }
          __syncwarp();
          if (kq == 0) HALO_TRACE(0, ptile);
          if (++s == (uint32_t)p.a_stages) { s = 0; ph ^= 1u; }
          if (!p.resident) {
            for (int gq = 0; gq < 3; ++gq) {  // one ring slot = one tap row (three 64-channel slabs)
              const int g = gq + g0 < 3 ? gq + g0 : gq + g0 - 3;
              mbar_wait(&b_empty[bs], bph);
              if (elect_one_sync()) {
                mbar_expect_tx(&b_full[bs], 3 * B_SLAB);
                for (int j = 0; j < 3; ++j)
                  tma_load_3d(smem + p.b_off + (bs * 3 + j) * B_SLAB, &p.tmB, &b_full[bs], kc * 64, col0,
                              p.wtap[g * 3 + j]);
              }
              __syncwarp();
              if (++bs == (uint32_t)p.b_slots) { bs = 0; bph ^= 1u; }
            }
          }
        }
      }
    }
  } else if (warp <= 2) {
    // ===================== MMA issuers (converged warp, elected issue) =====================
    // The issuing thread is instruction bound, not tensor bound, unless its loop is a few
    // instructions per MMA (profiles/r02_ncu_conv_source_notes.txt: 22 SASS instructions per
    // UTCHMMA made a 128x64x16 MMA cost 85 cycles instead of 48).  Taps are therefore kept in raster
    // order so that every operand offset inside a chunk is an immediate added to two running
    // descriptors: tap (i, j) reads the halo tile (i * 10 + j) rows further and weight slab 3i + j.
    {
      constexpr uint32_t idesc = umma_idesc_bf16(128, BLOCK_N, 0, 0);
      constexpr uint32_t BSLAB16 = B_SLAB >> 4;
      const uint32_t a_stage16 = (uint32_t)p.a_stage_bytes >> 4;
      const uint64_t adesc0 = umma_smem_desc(smem_u32(smem), 16, (uint32_t)kHaloPitch * 128);
      const uint64_t bdesc0 = umma_smem_desc(smem_u32(smem + p.b_off), 16, 1024);
      // tcgen05.mma issue is nearly synchronous with execution (tools/umma_rate.cu: a 36-MMA burst is
      // issued in 1825 cycles and complete after 1952), so the tensor pipe idles for every cycle
      // the issuing warp spends elsewhere: the whole role is straight-line converged code.
      // ONE elected thread runs the whole role.  tcgen05.mma issue is nearly synchronous with execution
      // (tools/umma_rate.cu: a 36-MMA burst is issued in 1825 cycles and complete after 1952), and every
      // separate `if (elect_one_sync()) {...} __syncwarp()` region costs ~180 cycles of tensor idle time
      // (same probe: groups of 4 MMAs take 380 cycles, of 36 take 1905), so the tile loop, its barrier
      // waits (3 cycles each once the phase is complete) and the commits all live inside one region.
      // TWO issuers alternate pixel tiles (warp 1 the even ones into accumulator 0, warp 2 the odd ones
      // into accumulator 1): between two tiles an issuer spends ~600 cycles on its barrier waits and
      // commits (same probe: 2 waits + 36 MMAs + 2 commits = 2316 cycles against 1814 for the MMAs alone),
      // and that now overlaps the other issuer's MMAs instead of idling the tensor pipe.
      static_assert(kAcc == 2, "one accumulator per MMA issuer");
      // An issuer must never be two phases ahead of a ring barrier (a parity wait cannot tell phase n
      // from phase n - 2), so the second issuer only runs with resident weights and a_stages > kchunks
      // (the creator sets p.issuers); the streaming path keeps one issuer for both accumulators.
      if (elect_one_sync() && warp - 1 < p.issuers) {
        const uint32_t me = (uint32_t)(warp - 1), nis = (uint32_t)p.issuers;
        // ring positions advance in the global (tile, chunk) order the producer fills them in
        uint32_t s = 0, ph = 0, bs = 0, bph = 0;
        auto skip_a = [&](uint32_t n) { s += n; while (s >= (uint32_t)p.a_stages) { s -= p.a_stages; ph ^= 1u; } };
        auto skip_b = [&](uint32_t n) { bs += n; while (bs >= (uint32_t)p.b_slots) { bs -= p.b_slots; bph ^= 1u; } };
        if (me) skip_a(p.kchunks);
        uint32_t tile = me;
        for (int mt = mt0 + (int)me * mstep; mt < p.mtiles; mt += (int)nis * mstep, tile += nis) {
          const uint32_t acc = tile & 1u, aph = (tile >> 1) & 1u;
          const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
          mbar_wait(&tmem_empty[acc], aph ^ 1u);
          HALO_TRACE(6, tile);
          for (int kq = 0; kq < p.kchunks; ++kq) {
            const int kc = kq + kc0 < p.kchunks ? kq + kc0 : kq + kc0 - p.kchunks;
            const bool last = kq == p.kchunks - 1;
            const int ks = kc == p.kchunks - 1 ? p.ksteps_last : 4;
            mbar_wait(&a_full[s], ph);
            if (kq == 0) HALO_TRACE(1, tile);
            const uint64_t ad_s = adesc0 + (uint64_t)(s * a_stage16);
            if (p.resident) {
              if (tile < 2) mbar_wait(&b_full[kc], 0);  // this chunk's nine weight slabs have landed
              tcgen05_fence_after();
              const uint64_t bd_k = bdesc0 + (uint64_t)((uint32_t)kc * 9u * BSLAB16);
              if (!(p.probe & 4)) halo_issue_taps_ks<BLOCK_N, 0, 9>(ks, tmem_d, ad_s, bd_k, idesc, kq != 0 ? 1u : 0u);
            } else {
              // one ring slot of weights = one tap row (three taps); rows in this CTA's rotated order
#pragma unroll 1
              for (int gq = 0; gq < 3; ++gq) {
                const int g = gq + g0 < 3 ? gq + g0 : gq + g0 - 3;
                mbar_wait(&b_full[bs], bph);
                tcgen05_fence_after();
                const uint64_t ad = ad_s + (uint64_t)(g * kHaloPitch * 8);
                const uint64_t bd = bdesc0 + (uint64_t)(bs * 3u * BSLAB16);
                halo_issue_taps_ks<BLOCK_N, 0, 3>(ks, tmem_d, ad, bd, idesc, (kq | gq) != 0 ? 1u : 0u);
                tcgen05_commit(&b_empty[bs]);
                skip_b(1);
              }
            }
            tcgen05_commit(&a_empty[s]);
            if (last) tcgen05_commit(&tmem_full[acc]);
            skip_a(1);
          }
          if (nis == 2) skip_a(p.kchunks);  // the other issuer's tile
          HALO_TRACE(2, tile);
        }
      }
      __syncwarp();
    }
  } else {
    // ===================== epilogue (8 warps) =====================
    const int wq = warp & 3;             // TMEM lane quadrant this warp may access
    const int half = (warp - kEpiWarp0) >> 2;    // which 32-column chunks this warp drains
    const int r = wq * 32 + lane;
    const int lh = r / p.wt, lw = r % p.wt;
    constexpr int NCH = BLOCK_N / 32;
    constexpr int MYCH = (NCH + 1) / 2;  // chunks per warp (chunk c = half + 2*ci)
    // BatchNorm partial sums stay in registers across ALL tiles of this CTA (row r of every tile,
    // this warp's columns); the cross-lane reduction runs once, after the tile loop.
    constexpr int SCH = STATS ? MYCH : 1;
    float s1[SCH][32], s2[SCH][32];
    if (STATS) {
#pragma unroll
      for (int ci = 0; ci < SCH; ++ci)
#pragma unroll
        for (int j = 0; j < 32; ++j) s1[ci][j] = s2[ci][j] = 0.0f;
    }
    uint32_t tile = 0, nstore = 0;
    uint8_t* epi_sm = smem + p.epi_off + (warp - kEpiWarp0) * 4096;  // two 2 KB staging slots per epilogue warp
    for (int mt = mt0; mt < p.mtiles; mt += mstep, ++tile, next_tile()) {
      const int pn = tn;
      const int phh = th * p.ht + lh, pw = tw * p.wt + lw;
      const bool row_ok = phh < p.oh && pw < p.ow;
      const long long row_off = (long long)pn * p.osn + (long long)phh * p.osh + (long long)pw * p.osw;
      const uint32_t acc = tile % kAcc, aph = (tile / kAcc) & 1u;
      // accumulate mode: fetch the old output values BEFORE waiting for the accumulator, so
      // their (uncoalesced, one pixel row per lane) latency hides behind the MMAs
      uint4 old[MYCH][4];
      if (!STATS && p.accumulate && row_ok && !p.tma_out) {
#pragma unroll
        for (int ci = 0; ci < MYCH; ++ci) {
          const int c = half + 2 * ci;
          if (c < NCH) {
#pragma unroll
            for (int g = 0; g < 4; ++g)
              old[ci][g] = *reinterpret_cast<const uint4*>(p.out + row_off + col0 + c * 32 + g * 8);
          }
        }
      }
      mbar_wait(&tmem_full[acc], aph);
      tcgen05_fence_after();
      if (warp == kEpiWarp0) HALO_TRACE(3, tile);
      const uint32_t tmem_d = tmem_base + acc * BLOCK_N + ((uint32_t)(wq * 32) << 16);
      if (half >= NCH) {  // BLOCK_N == 32: the second warp of the quadrant has no columns
        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      }
#pragma unroll
      for (int ci = 0; ci < MYCH; ++ci) {
        const int c = half + 2 * ci;
        if (c < NCH) {
          float v[32];
          if (p.probe & 32) {  // probe: no TMEM read
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = 0.0f;
          } else {
            tmem_ld_32x32b_x32(tmem_d + (uint32_t)(c * 32), v);
          }
          if (warp == kEpiWarp0 && ci == 0) HALO_TRACE(4, tile);
          if (c + 2 >= NCH) {  // this warp's last chunk is in registers: release the accumulator
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[acc]);
          }
          const int colc = col0 + c * 32;
          if (p.bias) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] += __ldg(p.bias + colc + j);
          }
          if (p.tma_out) {
            // Stage this warp's 32 pixels x 32 channels in shared memory (64 B rows, 64 B swizzle: the
            // 16-byte stores of a quarter warp land in eight different bank groups) and let TMA write --
            // or reduce-add, in accumulate mode -- the 4 x 8 pixel box.  A warp's scattered 16-byte
            // global stores cost 32 L1 wavefronts per instruction, and those wavefronts share the
            // shared-memory pipe with the tensor core's operand reads (an N = 64 MMA needs all of it):
            // 1024 (2048 with the old-value loads) wavefronts per tile added straight to the MMA time.
            uint8_t* sb = epi_sm + ((nstore & 1u) << 11);
            if (lane == 0) bulk_wait_group_read<1>();  // the store that last used this slot has read it
            __syncwarp();
            const int sw = (lane >> 1) & 3;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              uint4 pk;
              pk.x = f2_to_bf2(v[g * 8 + 0], v[g * 8 + 1]);
              pk.y = f2_to_bf2(v[g * 8 + 2], v[g * 8 + 3]);
              pk.z = f2_to_bf2(v[g * 8 + 4], v[g * 8 + 5]);
              pk.w = f2_to_bf2(v[g * 8 + 6], v[g * 8 + 7]);
              sts_u4(sb + lane * 64 + ((g ^ sw) << 4), pk);
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && !(p.probe & 1)) {
              if (!STATS && p.accumulate) tma_reduce_add_4d(&p.tmO, sb, colc, tw * 8, th * 16 + wq * 4, pn);
              else tma_store_4d(&p.tmO, sb, colc, tw * 8, th * 16 + wq * 4, pn);
              bulk_commit_group();
            }
            ++nstore;
          } else if (row_ok && !(p.probe & 1)) {
            bf16* dst = p.out + row_off + colc;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              float o[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) o[j] = v[g * 8 + j];
              if (!STATS && p.accumulate) {
                const uint4 od = old[ci][g];
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
              *reinterpret_cast<uint4*>(dst + g * 8) = pk;
            }
          }
          if (row_ok) {
            if (STATS && !(p.probe & 2)) {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                s1[ci][j] += v[j];
                s2[ci][j] = fmaf(v[j], v[j], s2[ci][j]);
              }
            }
          }
        }
      }
      if (warp == kEpiWarp0) HALO_TRACE(5, tile);
    }
    if (p.tma_out && lane == 0) bulk_wait_group<0>();  // all of this warp's stores have completed
    if (STATS) {  // one (sum, sumsq) row per CTA group; rows no group owns are zero-filled
      float* stat_sm = reinterpret_cast<float*>(smem + p.stat_off) + wq * 2 * BLOCK_N;
#pragma unroll
      for (int ci = 0; ci < MYCH; ++ci) {
        const int c = half + 2 * ci;
        if (c < NCH) {
          const float a = warp_transpose_sum32(s1[ci], lane);
          const float b = warp_transpose_sum32(s2[ci], lane);
          stat_sm[c * 32 + lane] = a;
          stat_sm[BLOCK_N + c * 32 + lane] = b;
        }
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const float* sm0 = reinterpret_cast<const float*>(smem + p.stat_off);
      for (int e = threadIdx.x - 32 * kEpiWarp0; e < 2 * BLOCK_N; e += 256) {
        const float x = (sm0[e] + sm0[2 * BLOCK_N + e]) + (sm0[4 * BLOCK_N + e] + sm0[6 * BLOCK_N + e]);
        const int k = e / BLOCK_N, cc = e % BLOCK_N;
        if (p.stat_rows == 0) {  // fixed-point accumulators (b2seg_tapgemm_desc::stats_fixed)
          atomicAdd(reinterpret_cast<unsigned long long*>(p.stats) + (size_t)k * p.ngemm + col0 + cc,
                    (unsigned long long)__float2ll_rn(x * (k ? B2SEG_STATS_FIX_SQ : B2SEG_STATS_FIX_SUM)));
          continue;
        }
        p.stats[((size_t)mt0 * 2 + k) * p.ngemm + col0 + cc] = x;
        for (int rr = mt0 + mstep; rr < p.stat_rows; rr += mstep)
          p.stats[((size_t)rr * 2 + k) * p.ngemm + col0 + cc] = 0.0f;
      }
    }
    tcgen05_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc<kAcc * BLOCK_N>(tmem_base);
  }
}

struct HaloOp : b2seg_op {
  HaloParams p;
  dim3 grid;
  int block_n;
  size_t smem;
  const char* name() const override { return "halo3x3"; }
  int run(cudaStream_t st) override {
#define HALO_LAUNCH(BN)                                                        \
  do {                                                                        \
    if (p.stats) B2_CHECK_CUDA(b2_launch(halo3x3_kernel<BN, true>, grid, dim3(kThreadsH), smem, st, p));  \
    else B2_CHECK_CUDA(b2_launch(halo3x3_kernel<BN, false>, grid, dim3(kThreadsH), smem, st, p));         \
  } while (0)
    if (block_n == 128) HALO_LAUNCH(128);
    else if (block_n == 64) HALO_LAUNCH(64);
    else HALO_LAUNCH(32);
#undef HALO_LAUNCH
    B2_CHECK_CUDA(cudaGetLastError());
    if (p.trace) {  // debug: dump CTA 0's per-tile timeline after every run
      B2_CHECK_CUDA(cudaStreamSynchronize(st));
      std::vector<long long> h(64 * 8);
      B2_CHECK_CUDA(cudaMemcpy(h.data(), p.trace, h.size() * 8, cudaMemcpyDeviceToHost));
      if (FILE* f = fopen(getenv("B2SEG_HALO_TRACE"), "a")) {
        fprintf(f, "# halo3x3 BN=%d kchunks=%d mtiles=%d grid=%d resident=%d a_stages=%d  columns: tile prodA mma_afull mma_issued epi_full epi_ld epi_done mma_tmem_empty mma_half (cycles since first event)\n",
                block_n, p.kchunks, p.mtiles, grid.x, p.resident, p.a_stages);
        long long t0 = h[0];
        for (int i = 0; i < 64 && h[i * 8 + 1]; ++i) {
          fprintf(f, "%2d", i);
          const int order[8] = {0, 1, 2, 3, 4, 5, 6, 7};
          for (int k : order) fprintf(f, " %8lld", h[i * 8 + k] ? h[i * 8 + k] - t0 : -1LL);
          fprintf(f, "\n");
        }
        fclose(f);
      }
      B2_CHECK_CUDA(cudaMemset(p.trace, 0, 64 * 8 * 8));
    }
    return B2SEG_OK;
  }
};

constexpr int kSmemMax = 232448;  // 227 KB opt-in limit per CTA on sm_100

template <int BN>
int set_halo_attr() {
  static bool done = false;
  if (!done) {
    B2_CHECK_CUDA(cudaFuncSetAttribute(halo3x3_kernel<BN, true>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemMax));
    B2_CHECK_CUDA(cudaFuncSetAttribute(halo3x3_kernel<BN, false>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemMax));
    done = true;
  }
  return B2SEG_OK;
}

}  // namespace

// Returns B2SEG_OK with *out == nullptr when the descriptor is not a 3x3 stride-1 problem this
// kernel covers (the caller then builds the generic tap-gather op).
int b2_halo3x3_try_create(const b2seg_tapgemm_desc* d, b2seg_op** out) {
  *out = nullptr;
  if (d->nsrc != 1 || d->nsub != 1 || d->sub[0].ntaps != 9 || d->cq > 0 || d->wtaps != 9) return B2SEG_OK;
  const auto& sub = d->sub[0];
  if (sub.out_offset != 0) return B2SEG_OK;
  if (d->stats && d->accumulate) return B2SEG_OK;  // not a combination the halo epilogue implements
  unsigned seen = 0;
  for (int t = 0; t < 9; ++t) {
    const auto& tp = sub.taps[t];
    if (tp.src != 0 || tp.dh < -1 || tp.dh > 1 || tp.dw < -1 || tp.dw > 1) return B2SEG_OK;
    seen |= 1u << ((tp.dh + 1) * 3 + tp.dw + 1);
  }
  if (seen != 0x1ffu) return B2SEG_OK;
  const b2seg_tensor4& x = d->src[0];
  if (x.n != d->on || x.h != d->oh || x.w != d->ow) return B2SEG_OK;
  int wt, ht, nt;
  b2_pick_pixel_tile(d->on, d->oh, d->ow, &wt, &ht, &nt);
  if (nt != 1 || wt != 8) return B2SEG_OK;  // images smaller than one 8 x 16 brick: generic path

  const int nm = d->ngemm;
  const int kchunks = (d->cin + 63) / 64;
  const int a_stage = ((wt + 2) * (ht + 2) * 128 + 1023) / 1024 * 1024;
  const int bar_bytes = (2 * kMaxAStages + 2 * kMaxBSlots + 2 * kAcc) * 8 + 16;
  // choose BLOCK_N, the weight mode and the epilogue (TMA staging needs 4 KB per active epilogue warp)
  int bn = 0, resident = 0, a_stages = 0, b_slots = 0, tma_out = 0;
  const int cands[3] = {128, 64, 32};
  auto fixed = [&](int n) { return bar_bytes + 4 * 2 * n * 4 + 1024; };
  auto epi_bytes = [&](int n) { return (n >= 64 ? 8 : 4) * 4096 + 1024; };  // + alignment pad
  bool want_tma = true;
  if (const char* e = getenv("B2SEG_HALO_TMA_OUT")) want_tma = e[0] != '0';
  // 1) resident weights with the widest tile that leaves >= 3 A stages
  for (int c = 0; c < 3 && !bn; ++c) {
    const int n = cands[c];
    if (nm % n) continue;
    const int slabs = 9 * kchunks;
    if (slabs > kMaxBSlots) continue;
    const int left = kSmemMax - fixed(n) - slabs * n * 128;
    const int min_stages = 3;
    if (left >= min_stages * a_stage) {
      bn = n; resident = 1; b_slots = slabs;
      tma_out = want_tma && left - epi_bytes(n) >= min_stages * a_stage;
      a_stages = (left - (tma_out ? epi_bytes(n) : 0)) / a_stage;
      if (a_stages > kMaxAStages) a_stages = kMaxAStages;
    }
  }
  // 2) otherwise stream tap slabs through a second ring
  for (int c = 0; c < 3 && !bn; ++c) {
    const int n = cands[c];
    if (nm % n) continue;
    bn = n; resident = 0;
    a_stages = 3;
    tma_out = want_tma;
    b_slots = (kSmemMax - fixed(n) - a_stages * a_stage - (tma_out ? epi_bytes(n) : 0)) / (3 * n * 128);  // ring slot = one tap row (3 slabs)
    if (b_slots < 2 && tma_out) {
      tma_out = 0;
      b_slots = (kSmemMax - fixed(n) - a_stages * a_stage) / (3 * n * 128);
    }
    if (b_slots > 6) b_slots = 6;
    if (b_slots < 2) bn = 0;
  }
  if (!bn) return B2SEG_OK;

  auto* op = new (std::nothrow) HaloOp();
  B2_REQUIRE(op, "halo3x3: out of host memory");
  HaloParams& p = op->p;
  memset(&p, 0, sizeof(p));
  {
    const uint64_t dims[4] = {(uint64_t)x.c, (uint64_t)x.w, (uint64_t)x.h, (uint64_t)x.n};
    const uint64_t strides[3] = {(uint64_t)x.sw * 2, (uint64_t)x.sh * 2, (uint64_t)x.sn * 2};
    const uint32_t box[4] = {64, (uint32_t)(wt + 2), (uint32_t)(ht + 2), 1};
    int rc = b2_encode_tmap_bf16(&p.tmA, x.ptr, 4, dims, strides, box, 1);
    if (rc != B2SEG_OK) { delete op; return rc; }
  }
  {
    const uint64_t dims[3] = {(uint64_t)d->cin, (uint64_t)d->ngemm, (uint64_t)d->wtaps};
    const uint64_t strides[2] = {(uint64_t)d->cin * 2, (uint64_t)d->cin * d->ngemm * 2};
    const uint32_t box[3] = {64, (uint32_t)bn, 1};
    int rc = b2_encode_tmap_bf16(&p.tmB, d->weights, 3, dims, strides, box, 1);
    if (rc != B2SEG_OK) { delete op; return rc; }
  }
  for (int t = 0; t < 9; ++t)  // raster order: slot (dh+1)*3 + (dw+1)
    p.wtap[(sub.taps[t].dh + 1) * 3 + sub.taps[t].dw + 1] = (int8_t)sub.taps[t].wtap;
  p.kchunks = kchunks;
  p.ksteps_last = d->cin % 64 ? (d->cin % 64 + 15) / 16 : 4;
  p.wt = wt; p.ht = ht;
  p.tilesW = (d->ow + wt - 1) / wt;
  p.tilesH = (d->oh + ht - 1) / ht;
  p.mtiles = p.tilesW * p.tilesH * d->on;
  p.ntiles = nm / bn;
  p.on = d->on; p.oh = d->oh; p.ow = d->ow;
  p.out = (bf16*)d->out;
  p.osn = d->osn; p.osh = d->osh; p.osw = d->osw;
  p.ngemm = nm;
  p.bias = d->bias;
  p.stats = d->stats;
  p.accumulate = d->accumulate;
  p.stat_rows = d->stats_fixed ? 0 : b2seg_tapgemm_stat_rows(d->on, d->oh, d->ow);
  p.resident = resident; p.a_stages = a_stages; p.b_slots = b_slots;
  {
    const char* e = getenv("B2SEG_PROBE");
    p.probe = e ? atoi(e) : 0;
    if (const char* tr = getenv("B2SEG_HALO_TRACE")) {
      if (tr[0] && cudaMalloc(&p.trace, 64 * 8 * 8) == cudaSuccess) cudaMemset(p.trace, 0, 64 * 8 * 8);
      else p.trace = nullptr;
    }
    const char* st = getenv("B2SEG_ASTAGES");
    if (st && atoi(st) >= 1 && atoi(st) <= p.a_stages) p.a_stages = atoi(st);
  }
  p.a_stage_bytes = a_stage;
  p.b_off = a_stages * a_stage;
  p.bar_off = p.b_off + (resident ? b_slots : 3 * b_slots) * bn * 128;
  p.stat_off = p.bar_off + bar_bytes;
  p.epi_off = (p.stat_off + 4 * 2 * bn * 4 + 1023) / 1024 * 1024;
  p.tma_out = tma_out;
  p.issuers = (resident && a_stages > kchunks) ? 2 : 1;
  if (const char* e = getenv("B2SEG_HALO_ISSUERS")) if (e[0] == '1') p.issuers = 1;
  op->smem = (size_t)p.epi_off + (tma_out ? epi_bytes(bn) : 0) + 1024;
  if (tma_out) {
    const uint64_t dims[4] = {(uint64_t)nm, (uint64_t)d->ow, (uint64_t)d->oh, (uint64_t)d->on};
    const uint64_t strides[3] = {(uint64_t)d->osw * 2, (uint64_t)d->osh * 2, (uint64_t)d->osn * 2};
    const uint32_t box[4] = {32, 8, 4, 1};
    int rc = b2_encode_tmap_bf16(&p.tmO, d->out, 4, dims, strides, box, 2);
    if (rc != B2SEG_OK) { delete op; return rc; }
  }
  op->block_n = bn;
  // one persistent CTA per SM; a CTA owns one column tile, so the grid is a multiple of ntiles
  int groups = b2_num_sms() / p.ntiles;
  if (groups > p.mtiles) groups = p.mtiles;
  if (groups < 1) groups = 1;
  op->grid = dim3(groups * p.ntiles, 1, 1);
  int rc = bn == 128 ? set_halo_attr<128>() : bn == 64 ? set_halo_attr<64>() : set_halo_attr<32>();
  if (rc != B2SEG_OK) { delete op; return rc; }
  *out = op;
  return B2SEG_OK;
}
