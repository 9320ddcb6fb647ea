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
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM alloc + MMA issuer,
// warps 2..9 = epilogue.
#include <new>
#include <stdlib.h>
#include <string.h>

#include "ops.cuh"

namespace {

constexpr int kThreadsH = 320;
constexpr int kMaxAStages = 8;
constexpr int kMaxBSlots = 18;
constexpr int kAcc = 2;  // TMEM accumulator stages (4 measured no faster; 2 leaves TMEM for a PDL-overlapped successor)

struct alignas(64) HaloParams {
  CUtensorMap tmA;  // box (64 ch, wt+2, ht+2, 1)
  CUtensorMap tmB;  // box (64 ch, BLOCK_N, 1 tap)
  int8_t dh[9], dw[9], wtap[9];
  int kchunks;
  int wt, ht, tilesW, tilesH, mtiles, ntiles;
  int on, oh, ow;
  bf16* out;
  long long osn, osh, osw;
  int ngemm;
  const float* bias;
  float* stats;
  int accumulate;
  int resident, a_stages, b_slots, stat_rows;
  int probe;  // B2SEG_PROBE bits (perf experiments only): 1 no output stores, 2 no stats, 4 no MMAs, 8 issue every chunk's MMAs 3x
  int a_stage_bytes, b_off, bar_off, stat_off;
};

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
  const int pitch = p.wt + 2;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < p.a_stages; ++i) {
      mbar_init(&a_full[i], 1);
      mbar_init(&a_empty[i], 1);
    }
    for (int i = 0; i < p.b_slots; ++i) {
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
  }
  if (warp == 1) tmem_alloc<kAcc * BLOCK_N>(tmem_slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // barriers, TMEM and descriptors are set up; now wait for the producer grid

  if (warp == 0) {
    // ===================== TMA producer (converged warp, elected issue) =====================
    {
      uint32_t s = 0, ph = 1, bs = 0, bph = 1;  // ring slots and the parity a free slot shows
      bool first = true;
      const uint32_t halo_bytes = (uint32_t)(pitch * (p.ht + 2) * 128);
      for (int mt = mt0; mt < p.mtiles; mt += mstep, first = false) {
        const int tw = mt % p.tilesW, th = (mt / p.tilesW) % p.tilesH, n0 = mt / (p.tilesW * p.tilesH);
        const int w0 = tw * p.wt - 1, h0 = th * p.ht - 1;
        for (int kc = 0; kc < p.kchunks; ++kc) {
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
          if (++s == (uint32_t)p.a_stages) { s = 0; ph ^= 1u; }
          if (p.resident) {
            if (first) {
              for (int t = 0; t < 9; ++t) {
                const int slab = kc * 9 + t;
                if (elect_one_sync()) {
                  mbar_expect_tx(&b_full[slab], B_SLAB);
                  tma_load_3d(smem + p.b_off + slab * B_SLAB, &p.tmB, &b_full[slab], kc * 64, col0,
                              p.wtap[t]);
                }
                __syncwarp();
              }
            }
          } else {
            for (int t = 0; t < 9; ++t) {
              mbar_wait(&b_empty[bs], bph);
              if (elect_one_sync()) {
                mbar_expect_tx(&b_full[bs], B_SLAB);
                tma_load_3d(smem + p.b_off + bs * B_SLAB, &p.tmB, &b_full[bs], kc * 64, col0, p.wtap[t]);
              }
              __syncwarp();
              if (++bs == (uint32_t)p.b_slots) { bs = 0; bph ^= 1u; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (converged warp, elected issue) =====================
    {
      constexpr uint32_t idesc = umma_idesc_bf16(128, BLOCK_N, 0, 0);
      constexpr uint32_t BSLAB16 = B_SLAB >> 4;
      uint32_t toff16[9];  // tap (dh,dw) -> rows into the halo tile, in 16-byte descriptor units
#pragma unroll
      for (int t = 0; t < 9; ++t) toff16[t] = (uint32_t)(((p.dh[t] + 1) * pitch + (p.dw[t] + 1)) * 8);
      const uint32_t a_stage16 = (uint32_t)p.a_stage_bytes >> 4;
      const uint64_t adesc0 = umma_smem_desc(smem_u32(smem), 16, (uint32_t)pitch * 128);
      const uint64_t bdesc0 = umma_smem_desc(smem_u32(smem + p.b_off), 16, 1024);
      uint32_t s = 0, ph = 0, bs = 0, bph = 0, tile = 0;
      for (int mt = mt0; mt < p.mtiles; mt += mstep, ++tile) {
        const uint32_t acc = tile % kAcc, aph = (tile / kAcc) & 1u;
        mbar_wait(&tmem_empty[acc], aph ^ 1u);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
        for (int kc = 0; kc < p.kchunks; ++kc) {
          mbar_wait(&a_full[s], ph);
          tcgen05_fence_after();
          const uint64_t ad_s = adesc0 + (uint64_t)(s * a_stage16);
          const bool last = kc == p.kchunks - 1;
          if (p.resident && tile != 0) {
            // steady state: weights are resident, the whole chunk (9 taps x 4 MMAs) is one issue block
            const uint64_t bd_k = bdesc0 + (uint64_t)((uint32_t)kc * 9u * BSLAB16);
            if (elect_one_sync()) {
              for (int rep = (p.probe & 8) ? 2 : 0; rep > 0; --rep) {  // probe: extra MMA passes
#pragma unroll
                for (int t = 0; t < 9; ++t) {
                  const uint64_t ad = ad_s + (uint64_t)toff16[t];
                  const uint64_t bd = bd_k + (uint64_t)(t * BSLAB16);
#pragma unroll
                  for (int k = 0; k < 4; ++k)
                    umma_bf16_ss(tmem_d, ad + (uint64_t)(2 * k), bd + (uint64_t)(2 * k), idesc, 1u);
                }
              }
#pragma unroll
              for (int t = 0; t < 9; ++t) {
                if (p.probe & 4) break;
                const uint64_t ad = ad_s + (uint64_t)toff16[t];
                const uint64_t bd = bd_k + (uint64_t)(t * BSLAB16);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  umma_bf16_ss(tmem_d, ad + (uint64_t)(2 * k), bd + (uint64_t)(2 * k), idesc,
                               (t | k) != 0 ? 1u : (kc != 0 ? 1u : 0u));
              }
              tcgen05_commit(&a_empty[s]);
              if (last) tcgen05_commit(&tmem_full[acc]);
            }
            __syncwarp();
          } else {
#pragma unroll
            for (int t = 0; t < 9; ++t) {
              uint32_t slot;
              if (p.resident) {
                slot = kc * 9 + t;
                mbar_wait(&b_full[slot], 0);
              } else {
                slot = bs;
                mbar_wait(&b_full[bs], bph);
              }
              tcgen05_fence_after();
              const uint64_t ad = ad_s + (uint64_t)toff16[t];
              const uint64_t bd = bdesc0 + (uint64_t)(slot * BSLAB16);
              if (elect_one_sync()) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  umma_bf16_ss(tmem_d, ad + (uint64_t)(2 * k), bd + (uint64_t)(2 * k), idesc,
                               (kc | t | k) != 0 ? 1u : 0u);
                if (!p.resident) tcgen05_commit(&b_empty[bs]);
                if (t == 8) {
                  tcgen05_commit(&a_empty[s]);
                  if (last) tcgen05_commit(&tmem_full[acc]);
                }
              }
              __syncwarp();
              if (!p.resident && ++bs == (uint32_t)p.b_slots) { bs = 0; bph ^= 1u; }
            }
          }
          if (++s == (uint32_t)p.a_stages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else {
    // ===================== epilogue (8 warps) =====================
    const int wq = warp & 3;             // TMEM lane quadrant this warp may access
    const int half = (warp - 2) >> 2;    // which 32-column chunks this warp drains
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
    uint32_t tile = 0;
    for (int mt = mt0; mt < p.mtiles; mt += mstep, ++tile) {
      const int tw = mt % p.tilesW, th = (mt / p.tilesW) % p.tilesH, pn = mt / (p.tilesW * p.tilesH);
      const int phh = th * p.ht + lh, pw = tw * p.wt + lw;
      const bool row_ok = phh < p.oh && pw < p.ow;
      const long long row_off = (long long)pn * p.osn + (long long)phh * p.osh + (long long)pw * p.osw;
      const uint32_t acc = tile % kAcc, aph = (tile / kAcc) & 1u;
      // accumulate mode: fetch the old output values BEFORE waiting for the accumulator, so
      // their (uncoalesced, one pixel row per lane) latency hides behind the MMAs
      uint4 old[MYCH][4];
      if (!STATS && p.accumulate && row_ok) {
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
      const uint32_t tmem_d = tmem_base + acc * BLOCK_N + ((uint32_t)(wq * 32) << 16);
      if (half >= NCH) {  // BLOCK_N == 32: the second warp of the quadrant has no columns
        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      }
#pragma unroll
      for (int ci = 0; ci < MYCH; ++ci) {
        const int c = half + 2 * ci;
        if (c < NCH) {
          float v[32];
          tmem_ld_32x32b_x32(tmem_d + (uint32_t)(c * 32), v);
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
          if (row_ok && !(p.probe & 1)) {
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
    }
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
      for (int e = threadIdx.x - 64; e < 2 * BLOCK_N; e += 256) {
        const float x = (sm0[e] + sm0[2 * BLOCK_N + e]) + (sm0[4 * BLOCK_N + e] + sm0[6 * BLOCK_N + e]);
        const int k = e / BLOCK_N, cc = e % BLOCK_N;
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
  // choose BLOCK_N and the weight mode
  int bn = 0, resident = 0, a_stages = 0, b_slots = 0;
  const int cands[3] = {128, 64, 32};
  auto fixed = [&](int n) { return bar_bytes + 4 * 2 * n * 4 + 1024; };
  // 1) resident weights with the widest tile that leaves >= 3 A stages
  for (int c = 0; c < 3 && !bn; ++c) {
    const int n = cands[c];
    if (nm % n) continue;
    const int slabs = 9 * kchunks;
    if (slabs > kMaxBSlots) continue;
    const int left = kSmemMax - fixed(n) - slabs * n * 128;
    if (left >= 3 * a_stage) {
      bn = n; resident = 1; b_slots = slabs;
      a_stages = left / a_stage;
      if (a_stages > kMaxAStages) a_stages = kMaxAStages;
    }
  }
  // 2) otherwise stream tap slabs through a second ring
  for (int c = 0; c < 3 && !bn; ++c) {
    const int n = cands[c];
    if (nm % n) continue;
    bn = n; resident = 0;
    a_stages = 3;
    b_slots = (kSmemMax - fixed(n) - a_stages * a_stage) / (n * 128);
    if (b_slots > kMaxBSlots) b_slots = kMaxBSlots;
    if (b_slots < 4) bn = 0;
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
  for (int t = 0; t < 9; ++t) {
    p.dh[t] = sub.taps[t].dh;
    p.dw[t] = sub.taps[t].dw;
    p.wtap[t] = sub.taps[t].wtap;
  }
  p.kchunks = kchunks;
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
  p.stat_rows = b2seg_tapgemm_stat_rows(d->on, d->oh, d->ow);
  p.resident = resident; p.a_stages = a_stages; p.b_slots = b_slots;
  {
    const char* e = getenv("B2SEG_PROBE");
    p.probe = e ? atoi(e) : 0;
    const char* st = getenv("B2SEG_ASTAGES");
    if (st && atoi(st) >= 1 && atoi(st) <= p.a_stages) p.a_stages = atoi(st);
  }
  p.a_stage_bytes = a_stage;
  p.b_off = a_stages * a_stage;
  p.bar_off = p.b_off + b_slots * bn * 128;
  p.stat_off = p.bar_off + bar_bytes;
  op->smem = (size_t)p.stat_off + 4 * 2 * bn * 4 + 1024;
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
