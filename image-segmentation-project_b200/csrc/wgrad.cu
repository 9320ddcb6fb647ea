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

// Sum split-K partials in fixed order and scatter into the gradient layout.
// grid (cchunks, ceil(b_channels/16), 64/kRedI); a block owns kRedI channels of the tap side x 16
// columns x all taps.  The splits are dealt round-robin to kRedG thread groups (partial sums in
// split order, combined as ((g0+g1)+(g2+g3)): fixed order, bit-reproducible).  Partials are read
// along the column index (coalesced), transposed through shared memory and written along
// (channel, tap) -- contiguous in OIHW / IOHW parameter layouts.
constexpr int kRedI = 4, kRedJ = 16, kRedG = 4;
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const __grid_constant__ WgradParams p) {
  pdl_prologue();
  __shared__ float sm[kRedG][kRedJ][kRedI * B2SEG_MAX_TAPS + 1];
  const int cchunk = blockIdx.x, j0 = blockIdx.y * kRedJ, i0 = blockIdx.z * kRedI;
  const size_t split_stride = (size_t)p.npairs * 128 * p.ncols_pad;
  const int n1 = p.ntaps * kRedI * kRedJ;
  for (int e = threadIdx.x; e < n1 * kRedG; e += blockDim.x) {
    const int g = e / n1, o = e % n1;
    const int jj = o % kRedJ, i = (o / kRedJ) % kRedI, tap = o / (kRedJ * kRedI);
    const int chunk = tap * p.cchunks + cchunk;
    const int pair = chunk >> 1, row = (chunk & 1) * 64 + i0 + i;
    const int j = j0 + jj;
    float acc = 0.0f;
    if (j < p.b_channels) {
      // eight splits' loads in flight per round trip; the additions keep the split order
      const float* src = p.ws + ((size_t)pair * 128 + row) * p.ncols_pad + j;
      for (int s = g; s < p.nsplit; s += 8 * kRedG) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int ss = s + u * kRedG;
          v[u] = ss < p.nsplit ? __ldg(src + (size_t)ss * split_stride) : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u)
          if (s + u * kRedG < p.nsplit) acc += v[u];
      }
    }
    sm[g][jj][i * p.ntaps + tap] = acc;
  }
  __syncthreads();
  const int run = kRedI * p.ntaps;
  for (int e = threadIdx.x; e < kRedJ * run; e += blockDim.x) {
    const int jj = e / run, r = e % run;
    const int i = r / p.ntaps, tap = r % p.ntaps;
    const int ch = cchunk * 64 + i0 + i, j = j0 + jj;
    if (ch >= p.a_channels || j >= p.b_channels) continue;
    float* dst = p.grad + p.taps[tap].out_base + (long long)ch * p.row_stride +
                 (long long)j * p.col_stride;
    const float v = (sm[0][jj][r] + sm[1][jj][r]) + (sm[2][jj][r] + sm[3][jj][r]);
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
    if (block_n == 128)
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
  return wgrad_geom(a_channels, b_channels, ntaps, pn, ph, pw).ws_bytes;
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
  const WgradGeom g = wgrad_geom(a_ch, b_ch, d->ntaps, d->pn, d->ph, d->pw);
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
    int rc = b2_view_tmap(&p.tmA[i], &d->a_src[i], g.wt, g.ht, g.nt);
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
  op->block_n = g.block_n;
  op->grid = dim3(g.npairs, g.ntiles, g.nsplit);
  op->reduce_grid = dim3(g.cchunks, (b_ch + kRedJ - 1) / kRedJ, 64 / kRedI);
  int rc = g.block_n == 128 ? set_wg_attr<128, 3>() : set_wg_attr<64, 4>();
  if (rc != B2SEG_OK) { delete op; return rc; }
  *out = op;
  return B2SEG_OK;
}
