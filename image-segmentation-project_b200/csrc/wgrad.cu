// Weight-gradient GEMM on tcgen05 tensor cores.
//
//   G[tap][i][j] = sum_{pix} A_tap[pix + off_tap, i] * B[pix, j]
//
// The reduction (K) dimension is the pixel grid, consumed in the same 128-pixel
// bricks the forward kernel uses.  TMA lands every brick as [128 pixels][64 ch]
// rows of 128 B (128B swizzle); read with pixels as K this is exactly the
// MN-major canonical UMMA layout, so both operands go to tcgen05.mma untouched
// (a_major = b_major = MN).  M = 128 = two 64-channel chunks of the tap side
// (chunk = (tap, channel block)), N = BLOCK_N channels of the other side.
// Split-K over pixel bricks; each CTA writes an fp32 partial tile, and a second
// kernel reduces the partials in fixed order (deterministic) and scatters into
// the parameter-gradient layout.
#include <new>
#include <string.h>

#include "ops.cuh"

namespace {

struct WTap {
  int8_t src, dh, dw, pad;
  long long out_base;
};

struct alignas(64) WgradParams {
  CUtensorMap tmA[4];
  CUtensorMap tmB;
  WTap taps[B2SEG_MAX_TAPS];
  int ntaps, cchunks, nchunks, npairs;  // chunks = ntaps * cchunks
  int a_channels, b_channels;
  int wt, ht, nt, tilesW, tilesH, ktiles;
  int nsplit, ncols_pad;
  float* ws;
  float* grad;
  long long row_stride, col_stride;
  int accumulate;
  int halo;          // 1: partials come from wgrad_halo_kernel (pair = cchunk * 5 + slot / 2)
  int8_t slot[12];   // halo: raster position (dh+1)*3 + (dw+1) of tap t
};

constexpr int kThreadsWG = 192;
constexpr int CHUNK_BYTES = 128 * 128;  // [128 pixels][64 ch] bf16

template <int BLOCK_N, int NSTAGES>
struct WGSmem {
  static constexpr int NB = BLOCK_N / 64;
  static constexpr int STAGE = (2 + NB) * CHUNK_BYTES;
  static constexpr int BAR_OFF = NSTAGES * STAGE;
  static constexpr int TOTAL = BAR_OFF + (2 * NSTAGES + 1) * 8 + 16 + 1024;
};

template <int BLOCK_N, int NSTAGES>
__global__ void __launch_bounds__(kThreadsWG, 1)
    wgrad_kernel(const __grid_constant__ WgradParams p) {
  pdl_trigger();
  using L = WGSmem<BLOCK_N, NSTAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* empty_bar = full_bar + NSTAGES;
  uint64_t* tmem_full = empty_bar + NSTAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = warp_id_uniform(), lane = threadIdx.x & 31;
  const int pair = blockIdx.x, ntile = blockIdx.y, split = blockIdx.z;
  const int my_ktiles = (p.ktiles - split + p.nsplit - 1) / p.nsplit;  // split, split+nsplit, ...

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < NSTAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
    tma_prefetch_desc(&p.tmB);
  }
  if (warp == 1) tmem_alloc<BLOCK_N>(tmem_slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // barriers, TMEM and descriptors are set up; now wait for the producer grid

  if (warp == 0) {
    {  // converged producer warp, elected issue
      // the two 64-channel chunks forming this CTA's M = 128
      int ch[2] = {2 * pair, min(2 * pair + 1, p.nchunks - 1)};
      const CUtensorMap* tm[2];
      int dh[2], dw[2], c0[2];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int tap = ch[i] / p.cchunks;
        tm[i] = &p.tmA[p.taps[tap].src];
        dh[i] = p.taps[tap].dh;
        dw[i] = p.taps[tap].dw;
        c0[i] = (ch[i] % p.cchunks) * 64;
      }
      uint32_t s = 0, ph = 1;  // ring slot and the parity a free slot shows
      int kt = split;
      for (int it = 0; it < my_ktiles; ++it, kt += p.nsplit) {
        const int tw = kt % p.tilesW, th = (kt / p.tilesW) % p.tilesH,
                  tn = kt / (p.tilesW * p.tilesH);
        const int w0 = tw * p.wt, h0 = th * p.ht, n0 = tn * p.nt;
        mbar_wait(&empty_bar[s], ph);
        if (elect_one_sync()) {
          mbar_expect_tx(&full_bar[s], L::STAGE);
          uint8_t* st = smem + s * L::STAGE;
          tma_load_4d(st, tm[0], &full_bar[s], c0[0], w0 + dw[0], h0 + dh[0], n0);
          tma_load_4d(st + CHUNK_BYTES, tm[1], &full_bar[s], c0[1], w0 + dw[1], h0 + dh[1], n0);
#pragma unroll
          for (int b = 0; b < L::NB; ++b)
            tma_load_4d(st + (2 + b) * CHUNK_BYTES, &p.tmB, &full_bar[s],
                        ntile * BLOCK_N + b * 64, w0, h0, n0);
        }
        __syncwarp();
        if (++s == NSTAGES) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    {  // converged MMA warp, elected issue
      constexpr uint32_t idesc = umma_idesc_bf16(128, BLOCK_N, 1, 1);
      // descriptors of ring slot 0, K step 0; slot s adds s*STAGE/16, K step k adds 2048/16
      const uint64_t adesc0 = umma_smem_desc(smem_u32(smem), CHUNK_BYTES, 1024);
      const uint64_t bdesc0 = umma_smem_desc(smem_u32(smem) + 2 * CHUNK_BYTES, CHUNK_BYTES, 1024);
      uint32_t s = 0, ph = 0;
      for (int it = 0; it < my_ktiles; ++it) {
        mbar_wait(&full_bar[s], ph);
        tcgen05_fence_after();
        const uint64_t so = (uint64_t)(s * (uint32_t)(L::STAGE >> 4));
        const uint64_t ad = adesc0 + so, bd = bdesc0 + so;
        if (elect_one_sync()) {
          umma_bf16_ss(tmem_base, ad, bd, idesc, it != 0 ? 1u : 0u);
#pragma unroll
          for (int k = 1; k < 8; ++k)  // 16 pixels (rows of 128 B) per MMA
            umma_bf16_ss(tmem_base, ad + (uint64_t)(k * 128), bd + (uint64_t)(k * 128), idesc, 1u);
          tcgen05_commit(&empty_bar[s]);
          if (it == my_ktiles - 1) tcgen05_commit(tmem_full);
        }
        __syncwarp();
        if (++s == NSTAGES) { s = 0; ph ^= 1u; }
      }
    }
  } else {
    const int wq = warp & 3;
    const int r = wq * 32 + lane;
    float* dst = p.ws + ((size_t)((size_t)split * p.npairs + pair) * 128 + r) * p.ncols_pad +
                 (size_t)ntile * BLOCK_N;
    if (my_ktiles > 0) {
      mbar_wait(tmem_full, 0);
      tcgen05_fence_after();
    }
#pragma unroll 1
    for (int c = 0; c < BLOCK_N / 32; ++c) {
      float v[32];
      if (my_ktiles > 0) {
        tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(wq * 32) << 16) + (uint32_t)(c * 32), v);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.0f;
      }
#pragma unroll
      for (int g = 0; g < 8; ++g)
        *reinterpret_cast<float4*>(dst + c * 32 + g * 4) =
            make_float4(v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
    }
    tcgen05_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc<BLOCK_N>(tmem_base);
  }
}

// ---- 3x3 stride-1 weight gradient with halo reuse ---------------------------------
// The generic kernel above loads one shifted copy of the input brick per tap: 48 KB of L2->SM
// traffic per 2.1 MFLOP, which caps it near 350 TFLOP/s.  Here a CTA owns ALL nine taps of one
// 64-channel input chunk x one 64-channel dY tile: per 8(w) x 16(h) pixel brick it loads the
// input brick once WITH its one-pixel halo (10 x 18 pixels, OOB zero fill = the padding) plus the
// dY brick, 39 KB for 9.4 MFLOP.  The tap (dh, dw) operand is the same tile read from row
// (dh+1)*10 + (dw+1) with an 8-pixel K-group stride of 10 rows (SBO = 1280 B); the hardware
// applies the 128B swizzle to absolute address bits, so unaligned starts read what TMA wrote
// (tools/umma_probe.cu, MN-major cases).  Two raster-adjacent taps form one M = 128 operand
// (LBO = their distance, 128 B or 1024 B); the ninth tap is paired with a dummy whose rows are
// never read back.  Five fp32 accumulators (320 TMEM columns) stay resident over the CTA's bricks.
constexpr int kHaloRows = 10 * 18;
constexpr int kHaloABytes = (kHaloRows * 128 + 1023) / 1024 * 1024;  // 23552
constexpr int kHaloStage = kHaloABytes + CHUNK_BYTES;
constexpr int kHaloStages = 5;
constexpr int kHaloBarOff = kHaloStages * kHaloStage;
constexpr int kHaloSmem = kHaloBarOff + (2 * kHaloStages + 1) * 8 + 16 + 1024;

__global__ void __launch_bounds__(kThreadsWG, 1)
    wgrad_halo_kernel(const __grid_constant__ WgradParams p) {
  pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kHaloBarOff);
  uint64_t* empty_bar = full_bar + kHaloStages;
  uint64_t* tmem_full = empty_bar + kHaloStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = warp_id_uniform(), lane = threadIdx.x & 31;
  const int cchunk = blockIdx.x, ntile = blockIdx.y, split = blockIdx.z;
  const int my_ktiles = (p.ktiles - split + p.nsplit - 1) / p.nsplit;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < kHaloStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
    tma_prefetch_desc(&p.tmA[0]);
    tma_prefetch_desc(&p.tmB);
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    uint32_t s = 0, ph = 1;
    int kt = split;
    for (int it = 0; it < my_ktiles; ++it, kt += p.nsplit) {
      const int tw = kt % p.tilesW, th = (kt / p.tilesW) % p.tilesH, tn = kt / (p.tilesW * p.tilesH);
      const int w0 = tw * 8, h0 = th * 16;
      mbar_wait(&empty_bar[s], ph);
      if (elect_one_sync()) {
        mbar_expect_tx(&full_bar[s], kHaloRows * 128 + CHUNK_BYTES);
        uint8_t* st = smem + s * kHaloStage;
        tma_load_4d(st, &p.tmA[0], &full_bar[s], cchunk * 64, w0 - 1, h0 - 1, tn);
        tma_load_4d(st + kHaloABytes, &p.tmB, &full_bar[s], ntile * 64, w0, h0, tn);
      }
      __syncwarp();
      if (++s == kHaloStages) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 1, 1);
    // pair q = raster taps (2q, 2q+1); tap r starts (r/3)*10 + r%3 rows into the halo tile
    uint64_t adesc0[5];
#pragma unroll
    for (int q = 0; q < 5; ++q) {
      const int r0 = 2 * q, r1 = 2 * q + 1;
      const int o0 = (r0 / 3) * 10 + r0 % 3, o1 = (r1 / 3) * 10 + r1 % 3;
      adesc0[q] = umma_smem_desc(smem_u32(smem) + (uint32_t)o0 * 128u, q < 4 ? (uint32_t)(o1 - o0) * 128u : 128u, 1280);
    }
    const uint64_t bdesc0 = umma_smem_desc(smem_u32(smem) + kHaloABytes, CHUNK_BYTES, 1024);
    uint32_t s = 0, ph = 0;
    for (int it = 0; it < my_ktiles; ++it) {
      mbar_wait(&full_bar[s], ph);
      tcgen05_fence_after();
      const uint64_t so = (uint64_t)(s * (uint32_t)(kHaloStage >> 4));
      if (elect_one_sync()) {
#pragma unroll
        for (int q = 0; q < 5; ++q) {
          const uint64_t ad = adesc0[q] + so, bd = bdesc0 + so;
#pragma unroll
          for (int k = 0; k < 8; ++k)  // 16 pixels = two brick rows per MMA
            umma_bf16_ss(tmem_base + (uint32_t)(q * 64), ad + (uint64_t)(k * 160), bd + (uint64_t)(k * 128),
                         idesc, (it != 0 || k != 0) ? 1u : 0u);
        }
        tcgen05_commit(&empty_bar[s]);
        if (it == my_ktiles - 1) tcgen05_commit(tmem_full);
      }
      __syncwarp();
      if (++s == kHaloStages) { s = 0; ph ^= 1u; }
    }
  } else {
    const int wq = warp & 3;
    const int r = wq * 32 + lane;
    if (my_ktiles > 0) {
      mbar_wait(tmem_full, 0);
      tcgen05_fence_after();
    }
#pragma unroll 1
    for (int q = 0; q < 5; ++q) {
      if (q == 4 && r >= 64) break;  // dummy half of the last pair
      float* dst = p.ws + ((size_t)((size_t)split * p.npairs + cchunk * 5 + q) * 128 + r) * p.ncols_pad +
                   (size_t)ntile * 64;
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        float v[32];
        if (my_ktiles > 0) {
          tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(wq * 32) << 16) + (uint32_t)(q * 64 + c * 32), v);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.0f;
        }
#pragma unroll
        for (int g = 0; g < 8; ++g)
          *reinterpret_cast<float4*>(dst + c * 32 + g * 4) =
              make_float4(v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
      }
    }
    tcgen05_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// Sum split-K partials in fixed order and scatter into the gradient layout.
// grid (cchunks, ceil(b_channels/kRedJ), 64/kRedI); a block owns kRedI channels of the tap side x kRedJ
// columns x all taps.  The reads are the whole cost (up to 148 splits of an L2-resident workspace) and
// they are latency bound, so they are spread as wide as the block allows: a thread owns one float4 of
// one (tap, channel) row and one of kRedG split groups, keeps all its loads in flight, and leaves a
// partial sum in shared memory; the partials are combined as ((g0+g1)+(g2+g3))+((g4+g5)+(g6+g7)) --
// fixed order, bit-reproducible -- and written along (channel, tap), contiguous in OIHW / IOHW.
// (The first version walked the splits serially per thread: 22 us per launch at 341 GB/s, 1.25 ms/step.)
constexpr int kRedI = 2, kRedJ = 32, kRedG = 8, kRedPitch = 36;
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const __grid_constant__ WgradParams p) {
  pdl_prologue();
  __shared__ __align__(16) float sm[kRedG][kRedI * B2SEG_MAX_TAPS][kRedPitch];
  const int cchunk = blockIdx.x, j0 = blockIdx.y * kRedJ, i0 = blockIdx.z * kRedI;
  const size_t split_stride = (size_t)p.npairs * 128 * p.ncols_pad;
  const int nrows = kRedI * p.ntaps;
  const int j4 = threadIdx.x & 7, slot = threadIdx.x >> 3;
  const int j = j0 + j4 * 4;
  for (int u = slot; u < nrows * kRedG; u += 32) {
    const int g = u % kRedG, r = u / kRedG;
    const int i = r / p.ntaps, tap = r % p.ntaps;
    int pair, row;
    if (p.halo) {
      const int q = p.slot[tap];
      pair = cchunk * 5 + (q >> 1);
      row = (q & 1) * 64 + i0 + i;
    } else {
      const int chunk = tap * p.cchunks + cchunk;
      pair = chunk >> 1;
      row = (chunk & 1) * 64 + i0 + i;
    }
    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (j < p.b_channels) {
      const float4* src = reinterpret_cast<const float4*>(p.ws + ((size_t)pair * 128 + row) * p.ncols_pad + j);
      const size_t st4 = split_stride / 4;
      int sidx = g;
      for (; sidx + 3 * kRedG < p.nsplit; sidx += 4 * kRedG) {  // four loads in flight per step
        const float4 a = __ldg(src + (size_t)sidx * st4), b = __ldg(src + (size_t)(sidx + kRedG) * st4),
                     c = __ldg(src + (size_t)(sidx + 2 * kRedG) * st4), d = __ldg(src + (size_t)(sidx + 3 * kRedG) * st4);
        acc.x = (((acc.x + a.x) + b.x) + c.x) + d.x;
        acc.y = (((acc.y + a.y) + b.y) + c.y) + d.y;
        acc.z = (((acc.z + a.z) + b.z) + c.z) + d.z;
        acc.w = (((acc.w + a.w) + b.w) + c.w) + d.w;
      }
      for (; sidx < p.nsplit; sidx += kRedG) {
        const float4 a = __ldg(src + (size_t)sidx * st4);
        acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
      }
    }
    *reinterpret_cast<float4*>(&sm[g][r][j4 * 4]) = acc;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < kRedJ * nrows; e += blockDim.x) {
    const int jj = e / nrows, r = e % nrows;
    const int i = r / p.ntaps, tap = r % p.ntaps;
    const int ch = cchunk * 64 + i0 + i, jo = j0 + jj;
    if (ch >= p.a_channels || jo >= p.b_channels) continue;
    float* dst = p.grad + p.taps[tap].out_base + (long long)ch * p.row_stride + (long long)jo * p.col_stride;
    const float v = ((sm[0][r][jj] + sm[1][r][jj]) + (sm[2][r][jj] + sm[3][r][jj])) +
                    ((sm[4][r][jj] + sm[5][r][jj]) + (sm[6][r][jj] + sm[7][r][jj]));
    *dst = p.accumulate ? *dst + v : v;
  }
}

struct WgradOp : b2seg_op {
  WgradParams p;
  dim3 grid;
  int block_n;
  dim3 reduce_grid;
  const char* name() const override { return "wgrad"; }
  int launches() const override { return 2; }
  int run(cudaStream_t st) override {
    if (p.halo)
      B2_CHECK_CUDA(b2_launch(wgrad_halo_kernel, grid, dim3(kThreadsWG), (size_t)kHaloSmem, st, p));
    else if (block_n == 128)
      B2_CHECK_CUDA(b2_launch(wgrad_kernel<128, 3>, grid, dim3(kThreadsWG), (size_t)WGSmem<128, 3>::TOTAL, st, p));
    else
      B2_CHECK_CUDA(b2_launch(wgrad_kernel<64, 4>, grid, dim3(kThreadsWG), (size_t)WGSmem<64, 4>::TOTAL, st, p));
    B2_CHECK_CUDA(cudaGetLastError());
    B2_CHECK_CUDA(b2_launch(wgrad_reduce_kernel, reduce_grid, dim3(256), 0, st, p));
    B2_CHECK_CUDA(cudaGetLastError());
    return B2SEG_OK;
  }
};

struct WgradGeom {
  int wt, ht, nt, tilesW, tilesH, ktiles, cchunks, nchunks, npairs, block_n, ntiles, nsplit,
      ncols_pad;
  size_t ws_bytes;
};

bool wgrad_halo_enabled() {
  const char* e = getenv("B2SEG_WGRAD_HALO");
  return !(e && e[0] == '0');
}

// geometry of the halo-reuse kernel; false when the pixel grid is smaller than one 8 x 16 brick
// Measured at 64 images (profiles/r01_op_times_b64*.txt, generic -> halo): 96->32 @128x128 235 -> 171 us,
// 32->32 @128x128 158 -> 114, 128->64 @64x64 74 -> 71; but 64->64 @64x64 58 -> 65 (its 148-way split-K
// reduction becomes the long pole), and every dY wider than 64 channels loses (256->256 @16x16
// 39 -> 54): this kernel's N = 64 MMAs run at 48 cycles each, shared-memory-operand bound, where
// the generic kernel's N = 128 MMAs do twice the work in 64 (tools/umma_rate.cu; MN-major operands
// cost nothing extra).
bool wgrad_halo_geom(int a_ch, int b_ch, int pn, int ph, int pw, WgradGeom* out) {
  WgradGeom g;
  b2_pick_pixel_tile(pn, ph, pw, &g.wt, &g.ht, &g.nt);
  if (g.nt != 1 || g.wt != 8 || g.ht != 16) return false;
  const char* force = getenv("B2SEG_WGRAD_HALO");
  const bool forced = force && force[0] == '2';  // 2 = wherever the geometry allows (tests, probes)
  if (!forced && (b_ch > 64 || (a_ch <= 64 && b_ch == 64))) return false;
  g.tilesW = (pw + 7) / 8;
  g.tilesH = (ph + 15) / 16;
  g.ktiles = g.tilesW * g.tilesH * pn;
  g.cchunks = (a_ch + 63) / 64;
  g.nchunks = 9 * g.cchunks;
  g.npairs = 5 * g.cchunks;
  g.block_n = 64;
  g.ntiles = (b_ch + 63) / 64;
  g.ncols_pad = g.ntiles * 64;
  const int ctas = g.cchunks * g.ntiles;
  int ns = b2_num_sms() / ctas;
  if (ns > (g.ktiles + 3) / 4) ns = (g.ktiles + 3) / 4;
  if (ns < 1) ns = 1;
  if (ns > 65535) ns = 65535;
  g.nsplit = ns;
  g.ws_bytes = (size_t)ns * g.npairs * 128 * g.ncols_pad * sizeof(float);
  *out = g;
  return true;
}

WgradGeom wgrad_geom(int a_ch, int b_ch, int ntaps, int pn, int ph, int pw) {
  WgradGeom g;
  b2_pick_pixel_tile(pn, ph, pw, &g.wt, &g.ht, &g.nt);
  g.tilesW = (pw + g.wt - 1) / g.wt;
  g.tilesH = (ph + g.ht - 1) / g.ht;
  g.ktiles = g.tilesW * g.tilesH * ((pn + g.nt - 1) / g.nt);
  g.cchunks = (a_ch + 63) / 64;
  g.nchunks = ntaps * g.cchunks;
  g.npairs = (g.nchunks + 1) / 2;
  g.block_n = b_ch > 64 ? 128 : 64;
  g.ntiles = (b_ch + g.block_n - 1) / g.block_n;
  g.ncols_pad = g.ntiles * g.block_n;
  // one CTA per SM (192 KB of shared memory): split K so the grid is a single full wave
  const int ctas = g.npairs * g.ntiles;
  int ns = b2_num_sms() / ctas;
  // keep at least 4 pixel bricks per CTA so the pipeline has something to hide
  if (ns > (g.ktiles + 3) / 4) ns = (g.ktiles + 3) / 4;
  if (ns < 1) ns = 1;
  if (ns > 65535) ns = 65535;
  g.nsplit = ns;
  g.ws_bytes = (size_t)ns * g.npairs * 128 * g.ncols_pad * sizeof(float);
  return g;
}

template <int BN, int NS>
int set_wg_attr() {
  static bool done = false;
  if (!done) {
    B2_CHECK_CUDA(cudaFuncSetAttribute(wgrad_kernel<BN, NS>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       WGSmem<BN, NS>::TOTAL));
    done = true;
  }
  return B2SEG_OK;
}

}  // namespace

extern "C" size_t b2seg_wgrad_workspace_bytes(int a_channels, int b_channels, int ntaps, int pn,
                                              int ph, int pw) {
  if (a_channels <= 0 || b_channels <= 0 || ntaps <= 0 || pn <= 0 || ph <= 0 || pw <= 0) return 0;
  size_t need = wgrad_geom(a_channels, b_channels, ntaps, pn, ph, pw).ws_bytes;
  WgradGeom h;  // a 9-tap problem may take the halo path, whose partials are laid out differently
  if (ntaps == 9 && wgrad_halo_enabled() && wgrad_halo_geom(a_channels, b_channels, pn, ph, pw, &h) &&
      h.ws_bytes > need)
    need = h.ws_bytes;
  return need;
}

extern "C" int b2seg_wgrad_create(const b2seg_wgrad_desc* d, b2seg_op** out) {
  B2_REQUIRE(d && out, "wgrad_create: null argument");
  B2_REQUIRE(d->na_src >= 1 && d->na_src <= 4, "wgrad_create: na_src %d", d->na_src);
  B2_REQUIRE(d->ntaps >= 1 && d->ntaps <= B2SEG_MAX_TAPS, "wgrad_create: ntaps %d", d->ntaps);
  B2_REQUIRE(d->grad && d->workspace && d->b_src.ptr, "wgrad_create: null pointer");
  B2_REQUIRE(d->pn > 0 && d->ph > 0 && d->pw > 0, "wgrad_create: empty pixel grid");
  const int a_ch = d->a_src[0].c, b_ch = d->b_src.c;
  B2_REQUIRE(a_ch >= 8 && a_ch % 8 == 0 && b_ch >= 8 && b_ch % 8 == 0,
             "wgrad_create: channels (%d, %d) must be multiples of 8", a_ch, b_ch);
  WgradGeom g = wgrad_geom(a_ch, b_ch, d->ntaps, d->pn, d->ph, d->pw);
  // 3x3 stride 1 (one source of the pixel grid's own size, all nine offsets): halo-reuse kernel
  bool halo = false;
  int8_t slot[12] = {0};
  if (d->ntaps == 9 && d->na_src == 1 && wgrad_halo_enabled() && d->a_src[0].n == d->pn &&
      d->a_src[0].h == d->ph && d->a_src[0].w == d->pw) {
    unsigned seen = 0;
    bool ok = true;
    for (int t = 0; t < 9; ++t) {
      const auto& tp = d->taps[t];
      if (tp.src != 0 || tp.dh < -1 || tp.dh > 1 || tp.dw < -1 || tp.dw > 1) { ok = false; break; }
      slot[t] = (int8_t)((tp.dh + 1) * 3 + tp.dw + 1);
      seen |= 1u << slot[t];
    }
    WgradGeom h;
    if (ok && seen == 0x1ffu && wgrad_halo_geom(a_ch, b_ch, d->pn, d->ph, d->pw, &h)) {
      g = h;
      halo = true;
    }
  }
  B2_REQUIRE(d->workspace_bytes >= g.ws_bytes, "wgrad_create: workspace %zu < %zu bytes",
             d->workspace_bytes, g.ws_bytes);
  auto* op = new (std::nothrow) WgradOp();
  B2_REQUIRE(op, "wgrad_create: out of host memory");
  WgradParams& p = op->p;
  memset(&p, 0, sizeof(p));
  for (int i = 0; i < d->na_src; ++i) {
    if (d->a_src[i].c != a_ch || !d->a_src[i].ptr) {
      b2seg_set_error("wgrad_create: a_src[%d] channel mismatch", i);
      delete op;
      return B2SEG_EINVAL;
    }
    int rc;
    if (halo) {  // the brick with its one-pixel halo: 10 x 18 pixels, OOB zero fill = padding
      const b2seg_tensor4& x = d->a_src[i];
      const uint64_t dims[4] = {(uint64_t)x.c, (uint64_t)x.w, (uint64_t)x.h, (uint64_t)x.n};
      const uint64_t strides[3] = {(uint64_t)x.sw * 2, (uint64_t)x.sh * 2, (uint64_t)x.sn * 2};
      const uint32_t box[4] = {64, 10, 18, 1};
      rc = b2_encode_tmap_bf16(&p.tmA[i], x.ptr, 4, dims, strides, box, 1);
    } else {
      rc = b2_view_tmap(&p.tmA[i], &d->a_src[i], g.wt, g.ht, g.nt);
    }
    if (rc != B2SEG_OK) { delete op; return rc; }
  }
  for (int i = d->na_src; i < 4; ++i) p.tmA[i] = p.tmA[0];
  {
    int rc = b2_view_tmap(&p.tmB, &d->b_src, g.wt, g.ht, g.nt);
    if (rc != B2SEG_OK) { delete op; return rc; }
  }
  for (int t = 0; t < d->ntaps; ++t) {
    if (d->taps[t].src < 0 || d->taps[t].src >= d->na_src) {
      b2seg_set_error("wgrad_create: tap %d source out of range", t);
      delete op;
      return B2SEG_EINVAL;
    }
    p.taps[t].src = d->taps[t].src;
    p.taps[t].dh = d->taps[t].dh;
    p.taps[t].dw = d->taps[t].dw;
    p.taps[t].out_base = d->taps[t].out_base;
  }
  p.ntaps = d->ntaps;
  p.cchunks = g.cchunks; p.nchunks = g.nchunks; p.npairs = g.npairs;
  p.a_channels = a_ch; p.b_channels = b_ch;
  p.wt = g.wt; p.ht = g.ht; p.nt = g.nt;
  p.tilesW = g.tilesW; p.tilesH = g.tilesH; p.ktiles = g.ktiles;
  p.nsplit = g.nsplit; p.ncols_pad = g.ncols_pad;
  p.ws = (float*)d->workspace;
  p.grad = d->grad;
  p.row_stride = d->row_stride; p.col_stride = d->col_stride;
  p.accumulate = d->accumulate;
  p.halo = halo ? 1 : 0;
  memcpy(p.slot, slot, sizeof(p.slot));
  op->block_n = g.block_n;
  op->grid = halo ? dim3(g.cchunks, g.ntiles, g.nsplit) : dim3(g.npairs, g.ntiles, g.nsplit);
  op->reduce_grid = dim3(g.cchunks, (b_ch + kRedJ - 1) / kRedJ, 64 / kRedI);
  int rc;
  if (halo) {
    static bool done = false;
    rc = B2SEG_OK;
    if (!done) {
      if (cudaFuncSetAttribute(wgrad_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kHaloSmem) !=
          cudaSuccess) {
        b2seg_set_error("wgrad_create: cannot reserve %d bytes of shared memory", kHaloSmem);
        rc = B2SEG_ECUDA;
      }
      done = rc == B2SEG_OK;
    }
  } else {
    rc = g.block_n == 128 ? set_wg_attr<128, 3>() : set_wg_attr<64, 4>();
  }
  if (rc != B2SEG_OK) { delete op; return rc; }
  *out = op;
  return B2SEG_OK;
}
