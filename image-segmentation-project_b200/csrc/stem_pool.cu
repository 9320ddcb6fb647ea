// Encoder stem: 7x7 stride-2 convolution of the 1-channel image (advanced_models.py:76,268),
// its weight gradient, and the 3x3 stride-2 max-pool (:83,272) forward/backward.
//
// K = 49 is too thin for a tensor-core tile and the layer is bandwidth/latency bound, so it is a
// direct convolution on the CUDA cores: a block computes an 8x16 brick of output pixels from a
// 21x37 input patch staged in shared memory; each thread owns one pixel x 32 channels.
#include <stdlib.h>

#include "ew.cuh"

namespace {

constexpr int ST_TH = 8, ST_TW = 16;            // output-pixel brick
constexpr int ST_PH = 2 * ST_TH + 5;            // 21 input rows
constexpr int ST_PW = 2 * ST_TW + 5;            // 37 input cols
constexpr int ST_PWP = ST_PW + 1;               // padded row pitch

struct StemGeom {
  int OH, OW, tilesH, tilesW;
  long long tiles;
};
static inline StemGeom stem_geom(int n, int h, int w) {
  StemGeom g;
  g.OH = h / 2; g.OW = w / 2;
  g.tilesH = (g.OH + ST_TH - 1) / ST_TH;
  g.tilesW = (g.OW + ST_TW - 1) / ST_TW;
  g.tiles = (long long)n * g.tilesH * g.tilesW;
  return g;
}
static inline int stem_blocks(int n, int h, int w) {
  const long long t = stem_geom(n, h, w).tiles;
  return (int)(t < B2SEG_EW_MAX_PARTIAL_ROWS ? t : B2SEG_EW_MAX_PARTIAL_ROWS);
}

__device__ __forceinline__ void stem_load_patch(float (*patch)[ST_PWP], const float* x, int n, int H,
                                                int W, int oh0, int ow0) {
  const int ih0 = 2 * oh0 - 3, iw0 = 2 * ow0 - 3;
  for (int e = threadIdx.x; e < ST_PH * ST_PW; e += blockDim.x) {
    const int r = e / ST_PW, c = e % ST_PW;
    const int ih = ih0 + r, iw = iw0 + c;
    float v = 0.0f;
    if (ih >= 0 && ih < H && iw >= 0 && iw < W) v = __ldg(x + ((size_t)n * H + ih) * W + iw);
    patch[r][c] = v;
  }
}

__global__ void __launch_bounds__(256) stem_fwd_kernel(const EwArgs a) {
  pdl_prologue();
  const float* x = (const float*)a.p[0];
  const float* w = (const float*)a.p[1];
  bf16* y = (bf16*)a.p[2];
  float* partial = (float*)a.p[3];
  const int N = (int)a.i[0], H = (int)a.i[1], W = (int)a.i[2];
  const int OH = H / 2, OW = W / 2;
  const int tilesH = (OH + ST_TH - 1) / ST_TH, tilesW = (OW + ST_TW - 1) / ST_TW;
  const long long tiles = (long long)N * tilesH * tilesW;
  __shared__ __align__(16) float wsm[49][64];
  __shared__ float patch[ST_PH][ST_PWP];
  for (int e = threadIdx.x; e < 64 * 49; e += blockDim.x) {
    const int c = e / 49, t = e % 49;
    wsm[t][c] = w[e];
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int pix = threadIdx.x & 127, half = threadIdx.x >> 7;
  const int ty = pix / ST_TW, tx = pix % ST_TW;
  float ssum = 0.0f, ssq = 0.0f;  // this lane's channel (half*32 + lane), over this warp's pixels
  for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int tw = (int)(tile % tilesW), th = (int)((tile / tilesW) % tilesH);
    const int n = (int)(tile / ((long long)tilesW * tilesH));
    const int oh0 = th * ST_TH, ow0 = tw * ST_TW;
    __syncthreads();
    stem_load_patch(patch, x, n, H, W, oh0, ow0);
    __syncthreads();
    float acc[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) acc[j] = 0.0f;
#pragma unroll 1
    for (int kh = 0; kh < 7; ++kh) {
#pragma unroll
      for (int kw = 0; kw < 7; ++kw) {
        const float xv = patch[2 * ty + kh][2 * tx + kw];
        const float4* wr = reinterpret_cast<const float4*>(&wsm[kh * 7 + kw][half * 32]);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 wv = wr[q];
          acc[4 * q + 0] = fmaf(xv, wv.x, acc[4 * q + 0]);
          acc[4 * q + 1] = fmaf(xv, wv.y, acc[4 * q + 1]);
          acc[4 * q + 2] = fmaf(xv, wv.z, acc[4 * q + 2]);
          acc[4 * q + 3] = fmaf(xv, wv.w, acc[4 * q + 3]);
        }
      }
    }
    const int oh = oh0 + ty, ow = ow0 + tx;
    const bool ok = oh < OH && ow < OW;
    if (ok) {
      bf16* dst = y + (((size_t)n * OH + oh) * OW + ow) * 64 + half * 32;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        F8 o;
#pragma unroll
        for (int j = 0; j < 8; ++j) o.v[j] = acc[q * 8 + j];
        st_bf8(dst + q * 8, o);
      }
    }
    if (partial) {
      float s2[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        acc[j] = ok ? acc[j] : 0.0f;
        s2[j] = acc[j] * acc[j];
      }
      ssum += warp_transpose_sum32(acc, lane);
      ssq += warp_transpose_sum32(s2, lane);
    }
  }
  if (partial) {  // row = block*4 + (warp & 3): [2][64]
    float* row = partial + ((size_t)blockIdx.x * 4 + (warp & 3)) * 128;
    row[half * 32 + lane] = ssum;
    row[64 + half * 32 + lane] = ssq;
  }
}

// dW[c][tap] = sum_pix dy[pix, c] * x[2*oh-3+kh, 2*ow-3+kw]; thread = (16-channel group, tap).
__global__ void __launch_bounds__(256) stem_wgrad_kernel(const EwArgs a) {
  pdl_prologue();
  const float* x = (const float*)a.p[0];
  const bf16* dy = (const bf16*)a.p[1];
  float* ws = (float*)a.p[2];
  const int N = (int)a.i[0], H = (int)a.i[1], W = (int)a.i[2];
  const int OH = H / 2, OW = W / 2;
  const int tilesH = (OH + ST_TH - 1) / ST_TH, tilesW = (OW + ST_TW - 1) / ST_TW;
  const long long tiles = (long long)N * tilesH * tilesW;
  __shared__ __align__(16) float dysm[128][64];
  __shared__ float patch[ST_PH][ST_PWP];
  const int cg = threadIdx.x / 49, tap = threadIdx.x % 49;
  const int kh = tap / 7, kw = tap % 7;
  const bool active = threadIdx.x < 196;
  float acc[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) acc[j] = 0.0f;
  for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int tw = (int)(tile % tilesW), th = (int)((tile / tilesW) % tilesH);
    const int n = (int)(tile / ((long long)tilesW * tilesH));
    const int oh0 = th * ST_TH, ow0 = tw * ST_TW;
    __syncthreads();
    stem_load_patch(patch, x, n, H, W, oh0, ow0);
    for (int e = threadIdx.x; e < 128 * 8; e += blockDim.x) {
      const int px = e >> 3, c8 = (e & 7) * 8;
      const int oh = oh0 + px / ST_TW, ow = ow0 + px % ST_TW;
      F8 v;
      if (oh < OH && ow < OW) {
        v = ld_bf8(dy + (((size_t)n * OH + oh) * OW + ow) * 64 + c8);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v.v[j] = 0.0f;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) dysm[px][c8 + j] = v.v[j];
    }
    __syncthreads();
    if (active) {
#pragma unroll 4
      for (int px = 0; px < 128; ++px) {
        const float xv = patch[2 * (px / ST_TW) + kh][2 * (px % ST_TW) + kw];
        const float4* dr = reinterpret_cast<const float4*>(&dysm[px][cg * 16]);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 dv = dr[q];
          acc[4 * q + 0] = fmaf(xv, dv.x, acc[4 * q + 0]);
          acc[4 * q + 1] = fmaf(xv, dv.y, acc[4 * q + 1]);
          acc[4 * q + 2] = fmaf(xv, dv.z, acc[4 * q + 2]);
          acc[4 * q + 3] = fmaf(xv, dv.w, acc[4 * q + 3]);
        }
      }
    }
  }
  if (active) {
    float* row = ws + (size_t)blockIdx.x * 3136;
#pragma unroll
    for (int j = 0; j < 16; ++j) row[(cg * 16 + j) * 49 + tap] = acc[j];
  }
}

// ------------------------------------------------------------------ max-pool ----
// torch max_pool2d semantics: padding never wins, first maximum in (kh, kw) scan order wins.
__global__ void __launch_bounds__(kEwThreads) maxpool_fwd_kernel(const EwArgs a) {
  pdl_prologue();
  const bf16* x = (const bf16*)a.p[0];
  bf16* out = (bf16*)a.p[1];
  uint8_t* idx = (uint8_t*)a.p[2];
  const int N = (int)a.i[0], H = (int)a.i[1], W = (int)a.i[2], C = (int)a.i[3];
  const long long ldx = a.i[4], ldo = a.i[5];
  const int OH = H / 2, OW = W / 2, cv = C >> 3;
  const long long total = (long long)N * OH * OW * cv;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    long long op, rest;
    int c8, ow, oh;
    ew_divmod(e, cv, op, c8);
    ew_divmod(op, OW, rest, ow);
    ew_divmod(rest, OH, rest, oh);
    const int c0 = c8 * 8, n = (int)rest;
    float best[8];
    uint32_t bi[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; bi[j] = 0; }
    // the nine window loads are issued together (predicated at the borders), then scanned in order
    uint4 raw[9];
    bool ok[9];
#pragma unroll
    for (int q = 0; q < 9; ++q) {
      const int ih = 2 * oh - 1 + q / 3, iw = 2 * ow - 1 + q % 3;
      ok[q] = ih >= 0 && ih < H && iw >= 0 && iw < W;
      raw[q] = ok[q] ? *reinterpret_cast<const uint4*>(x + (((long long)n * H + ih) * W + iw) * ldx + c0)
                     : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int q = 0; q < 9; ++q) {
      if (!ok[q]) continue;
      const F8 v = unpack_bf8(raw[q]);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (v.v[j] > best[j] || v.v[j] != v.v[j]) { best[j] = v.v[j]; bi[j] = q; }
    }
    F8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.v[j] = best[j];
    st_bf8(out + op * ldo + c0, o);
    uint2 pk;
    pk.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
    pk.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
    *reinterpret_cast<uint2*>(idx + op * C + c0) = pk;
  }
}

__global__ void __launch_bounds__(kEwThreads) maxpool_bwd_kernel(const EwArgs a) {
  pdl_prologue();
  const bf16* dout = (const bf16*)a.p[0];
  const uint8_t* idx = (const uint8_t*)a.p[1];
  bf16* dx = (bf16*)a.p[2];
  const int N = (int)a.i[0], H = (int)a.i[1], W = (int)a.i[2], C = (int)a.i[3];
  const long long lddo = a.i[4], lddx = a.i[5];
  const bool accumulate = a.i[6] != 0;
  const int OH = H / 2, OW = W / 2, cv = C >> 3;
  const long long total = (long long)N * H * W * cv;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    long long ip, rest;
    int c8, iw, ih;
    ew_divmod(e, cv, ip, c8);
    ew_divmod(ip, W, rest, iw);
    ew_divmod(rest, H, rest, ih);
    const int c0 = c8 * 8, n = (int)rest;
    // windows oh with 2*oh-1 <= ih <= 2*oh+1: row ih is tap 1 of window ih/2 when even, tap 2 of
    // window (ih-1)/2 and tap 0 of window (ih+1)/2 when odd (same along w): 1, 2 or 4 windows.
    // All their loads are issued together (predicated), then compared: one L2 round trip, not four.
    const int oh0 = ih >> 1, ow0 = iw >> 1;
    const bool vh = (ih & 1) && oh0 + 1 < OH, vw = (iw & 1) && ow0 + 1 < OW;
    const uint32_t kh0 = (ih & 1) ? 2u : 1u, kw0 = (iw & 1) ? 2u : 1u;  // second window: tap 0
    uint2 pk[4];
    uint4 dr[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const bool ok = ((q & 2) == 0 || vh) && ((q & 1) == 0 || vw);
      const long long op = ((long long)n * OH + oh0 + (q >> 1)) * OW + ow0 + (q & 1);
      pk[q] = ok ? *reinterpret_cast<const uint2*>(idx + op * C + c0) : make_uint2(~0u, ~0u);
      dr[q] = ok ? *reinterpret_cast<const uint4*>(dout + op * lddo + c0) : make_uint4(0u, 0u, 0u, 0u);
    }
    F8 g;
#pragma unroll
    for (int j = 0; j < 8; ++j) g.v[j] = 0.0f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t k = ((q & 2) ? 0u : kh0) * 3u + ((q & 1) ? 0u : kw0);
      const uint32_t kk = k * 0x01010101u;
      const uint32_t m0 = __vcmpeq4(pk[q].x, kk), m1 = __vcmpeq4(pk[q].y, kk);
      const F8 d = unpack_bf8(dr[q]);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t m = j < 4 ? m0 : m1;
        if (m & (1u << ((j & 3) * 8))) g.v[j] += d.v[j];
      }
    }
    bf16* dst = dx + ip * lddx + c0;
    if (accumulate) {
      const F8 old = ld_bf8(dst);
#pragma unroll
      for (int j = 0; j < 8; ++j) g.v[j] += old.v[j];
    }
    st_bf8(dst, g);
  }
}

#define EW_LAUNCHER(name, kernel)                                                         \
  int name(const EwArgs& a, dim3 grid, dim3 block, size_t smem, cudaStream_t st) {        \
    B2_CHECK_CUDA(b2_launch(kernel, grid, block, smem, st, a));                                                 \
    B2_CHECK_CUDA(cudaGetLastError());                                                    \
    return B2SEG_OK;                                                                      \
  }
EW_LAUNCHER(launch_stem_fwd, stem_fwd_kernel)
EW_LAUNCHER(launch_stem_wgrad, stem_wgrad_kernel)
EW_LAUNCHER(launch_maxpool_fwd, maxpool_fwd_kernel)
EW_LAUNCHER(launch_maxpool_bwd, maxpool_bwd_kernel)

}  // namespace

extern "C" int b2seg_stem_stat_rows(int n, int h, int w) {
  if (n <= 0 || h <= 1 || w <= 1) return 0;
  return stem_blocks(n, h, w) * 4;
}
extern "C" int b2seg_stem_wgrad_rows(int n, int h, int w) {
  if (n <= 0 || h <= 1 || w <= 1) return 0;
  return stem_blocks(n, h, w);
}

// stem_tc.cu: the same convolution on tcgen05 (B2SEG_STEM_TC=0 keeps the CUDA-core kernel)
int launch_stem_tc_fwd(const EwArgs& a, dim3 grid, dim3 block, size_t smem, cudaStream_t st);
int launch_stem_tc_wgrad(const EwArgs& a, dim3 grid, dim3 block, size_t smem, cudaStream_t st);
static bool stem_tc_enabled() {
  const char* e = getenv("B2SEG_STEM_TC");
  return !(e && e[0] == '0');
}

int ew_plan_stem(const b2seg_ew_desc* d, EwPlan* pl) {
  const int64_t* i = d->i;
  void* const* p = d->p;
  pl->smem = 0;
  pl->launches = 1;
  switch (d->kind) {
    case B2SEG_EW_STEM_FWD:
    case B2SEG_EW_STEM_WGRAD: {
      B2_REQUIRE(p[0] && p[1] && p[2], "stem: null argument");
      B2_REQUIRE(i[0] >= 1 && i[1] >= 2 && i[2] >= 2 && i[1] % 2 == 0 && i[2] % 2 == 0,
                 "stem: image %lldx%lld must have even height and width", i[1], i[2]);
      pl->fn = d->kind == B2SEG_EW_STEM_FWD ? (stem_tc_enabled() ? launch_stem_tc_fwd : launch_stem_fwd)
                                            : (stem_tc_enabled() ? launch_stem_tc_wgrad : launch_stem_wgrad);
      pl->grid = dim3(stem_blocks((int)i[0], (int)i[1], (int)i[2]));
      pl->block = dim3(256);
      return B2SEG_OK;
    }
    case B2SEG_EW_MAXPOOL_FWD:
    case B2SEG_EW_MAXPOOL_BWD: {
      B2_REQUIRE(p[0] && p[1] && p[2], "maxpool: null argument");
      B2_REQUIRE(i[0] >= 1 && i[1] >= 2 && i[2] >= 2 && i[1] % 2 == 0 && i[2] % 2 == 0 &&
                     i[3] >= 8 && i[3] % 8 == 0 && i[4] % 8 == 0 && i[5] % 8 == 0,
                 "maxpool: bad shape");
      const bool fwd = d->kind == B2SEG_EW_MAXPOOL_FWD;
      pl->fn = fwd ? launch_maxpool_fwd : launch_maxpool_bwd;
      const long long px = fwd ? i[0] * (i[1] / 2) * (i[2] / 2) : i[0] * i[1] * i[2];
      pl->grid = dim3(fwd ? ew_resident_blocks(maxpool_fwd_kernel, px * (i[3] / 8))
                          : ew_resident_blocks(maxpool_bwd_kernel, px * (i[3] / 8)));
      pl->block = dim3(kEwThreads);
      return B2SEG_OK;
    }
  }
  b2seg_set_error("ew_plan_stem: kind %d not handled here", d->kind);
  return B2SEG_EINVAL;
}
