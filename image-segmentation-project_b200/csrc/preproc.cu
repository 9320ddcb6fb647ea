// GPU input pipeline of CellSegmentationDataset.__getitem__ (reference dataset.py:30-66; SURVEY 8f
// rank 1): cv2.resize (INTER_AREA image / INTER_NEAREST mask), np.percentile(2, 98) clip, CLAHE
// (clip limit 2.0, 8 x 8 tiles), min-max to [0, 1] -- batched over the images of a fold, uint8
// planes in, fp32 [B,1,H,W] tensors out, no host round trip and no .tif files in between.
//
// Every kernel restates OpenCV's / numpy's arithmetic exactly (oracle/preproc_cv.py pins the same
// restatement against cv2 and numpy themselves): fixed-point bilinear (11-bit coefficients), the
// float tables of ResizeArea, integer histogram clipping / redistribution, float LUT blending with
// no fused multiply-adds, round-half-even saturating casts.  Coefficient tables are built on the
// host once per (source size, target size) (b2seg/preproc.py) and live in device memory.
#include "common.cuh"

namespace {

__device__ __forceinline__ int cv_round_sat_u8(float v) { return max(0, min(255, __float2int_rn(v))); }
__device__ __forceinline__ int reflect101(int p, int n) {
  if (n == 1) return 0;
  while ((unsigned)p >= (unsigned)n) p = p < 0 ? -p : 2 * (n - 1) - p;
  return p;
}

// ---- resize -------------------------------------------------------------------------
__global__ void __launch_bounds__(256) resize_nearest_u8_kernel(const uint8_t* __restrict__ src, int sh, int sw,
                                                                uint8_t* __restrict__ dst, float* __restrict__ dstf,
                                                                int dh, int dw, const int* __restrict__ xofs,
                                                                const int* __restrict__ yofs) {
  const uint8_t* s = src + (size_t)blockIdx.y * sh * sw;
  const long long n = (long long)dh * dw, stride = (long long)gridDim.x * 256;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    const int y = (int)(i / dw), x = (int)(i - (long long)y * dw);
    const uint8_t v = s[(size_t)yofs[y] * sw + xofs[x]];
    if (dst) dst[(size_t)blockIdx.y * n + i] = v;
    if (dstf) dstf[(size_t)blockIdx.y * n + i] = v > 0 ? 1.0f : 0.0f;  // mask: (mask > 0).astype(float32)
  }
}

// INTER_LINEAR (and INTER_AREA when enlarging): HResizeLinear (x 2048) + VResizeLinear fixed point.
__global__ void __launch_bounds__(256) resize_linear_u8_kernel(const uint8_t* __restrict__ src, int sh, int sw,
                                                               uint8_t* __restrict__ dst, int dh, int dw,
                                                               const int* __restrict__ xofs, const short* __restrict__ xa,
                                                               const int* __restrict__ yofs, const short* __restrict__ ya) {
  const uint8_t* s = src + (size_t)blockIdx.y * sh * sw;
  const long long n = (long long)dh * dw, stride = (long long)gridDim.x * 256;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    const int y = (int)(i / dw), x = (int)(i - (long long)y * dw);
    const int x0 = xofs[x], x1 = min(x0 + 1, sw - 1), y0 = yofs[y], y1 = min(y0 + 1, sh - 1);
    const int a0 = xa[2 * x], a1 = xa[2 * x + 1], b0 = ya[2 * y], b1 = ya[2 * y + 1];
    const int r0 = s[(size_t)y0 * sw + x0] * a0 + s[(size_t)y0 * sw + x1] * a1;
    const int r1 = s[(size_t)y1 * sw + x0] * a0 + s[(size_t)y1 * sw + x1] * a1;
    const int v = (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2;
    dst[(size_t)blockIdx.y * n + i] = (uint8_t)max(0, min(255, v));
  }
}

// INTER_AREA with integer ratios (ResizeAreaFast): block sum, times float(1/area), cvRound;
// the 2 x 2 case is (a + b + c + d + 2) >> 2.
__global__ void __launch_bounds__(256) resize_areafast_u8_kernel(const uint8_t* __restrict__ src, int sh, int sw,
                                                                 uint8_t* __restrict__ dst, int dh, int dw, int ix, int iy) {
  const uint8_t* s = src + (size_t)blockIdx.y * sh * sw;
  const long long n = (long long)dh * dw, stride = (long long)gridDim.x * 256;
  const float scale = 1.0f / (float)(ix * iy);
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    const int y = (int)(i / dw), x = (int)(i - (long long)y * dw);
    int sum = 0;
    for (int r = 0; r < iy; ++r) {
      const uint8_t* row = s + (size_t)(y * iy + r) * sw + (size_t)x * ix;
      for (int c = 0; c < ix; ++c) sum += row[c];
    }
    const int v = (ix == 2 && iy == 2) ? (sum + 2) >> 2 : cv_round_sat_u8(__fmul_rn((float)sum, scale));
    dst[(size_t)blockIdx.y * n + i] = (uint8_t)v;
  }
}

// INTER_AREA, fractional ratios (ResizeArea): per destination pixel, for every source row of its
// y-table entries: buf = sum_k S*alpha_k (float, table order, from 0); sum = beta*buf for the first
// row, sum += beta*buf after -- the accumulation order of ResizeArea_Invoker.
__global__ void __launch_bounds__(256) resize_area_u8_kernel(const uint8_t* __restrict__ src, int sh, int sw,
                                                             uint8_t* __restrict__ dst, int dh, int dw,
                                                             const int* __restrict__ xstart, const int* __restrict__ xsi,
                                                             const float* __restrict__ xalpha,
                                                             const int* __restrict__ ystart, const int* __restrict__ ysi,
                                                             const float* __restrict__ ybeta) {
  const uint8_t* s = src + (size_t)blockIdx.y * sh * sw;
  const long long n = (long long)dh * dw, stride = (long long)gridDim.x * 256;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    const int y = (int)(i / dw), x = (int)(i - (long long)y * dw);
    const int k0 = xstart[x], k1 = xstart[x + 1];
    float sum = 0.0f;
    for (int j = ystart[y]; j < ystart[y + 1]; ++j) {
      const uint8_t* row = s + (size_t)ysi[j] * sw;
      float buf = 0.0f;
      for (int k = k0; k < k1; ++k) buf = __fadd_rn(buf, __fmul_rn((float)row[xsi[k]], xalpha[k]));
      const float t = __fmul_rn(ybeta[j], buf);
      sum = j == ystart[y] ? t : __fadd_rn(sum, t);
    }
    dst[(size_t)blockIdx.y * n + i] = (uint8_t)cv_round_sat_u8(sum);
  }
}

// ---- np.percentile(image, [lo, hi]) -> clip LUT ----------------------------------------
// One block per image.  Order statistic k of a uint8 image = smallest v with cdf(v) > k.
// p = a + (b - a)*t (t < 0.5) or b - (b - a)*(1 - t), as numpy's _lerp; lut[v] = uint8(clip(v, p_lo, p_hi)).
__global__ void __launch_bounds__(256) percentile_lut_kernel(const unsigned* __restrict__ hist, long long npix,
                                                             double qlo, double qhi, uint8_t* __restrict__ lut,
                                                             double* __restrict__ pout) {
  __shared__ unsigned long long cdf[256];
  __shared__ double pv[2];
  const unsigned* h = hist + (size_t)blockIdx.x * 256;
  cdf[threadIdx.x] = h[threadIdx.x];
  __syncthreads();
  for (int o = 1; o < 256; o <<= 1) {
    const unsigned long long add = threadIdx.x >= o ? cdf[threadIdx.x - o] : 0ull;
    __syncthreads();
    cdf[threadIdx.x] += add;
    __syncthreads();
  }
  if (threadIdx.x < 2) {
    const double q = threadIdx.x == 0 ? qlo : qhi;
    const double vi = q / 100.0 * (double)(npix - 1);
    const long long k = (long long)floor(vi);
    const double t = vi - (double)k;
    const long long k1 = min(k + 1, npix - 1);
    int a = 0, b = 0;
    for (int v = 255; v >= 0; --v) {
      if (cdf[v] > (unsigned long long)k) a = v;
      if (cdf[v] > (unsigned long long)k1) b = v;
    }
    const double d = (double)(b - a);
    double p = (double)a + d * t;
    if (t >= 0.5) p = (double)b - d * (1.0 - t);
    pv[threadIdx.x] = p;
    if (pout) pout[(size_t)blockIdx.x * 2 + threadIdx.x] = p;
  }
  __syncthreads();
  double v = (double)threadIdx.x;
  v = v < pv[0] ? pv[0] : (v > pv[1] ? pv[1] : v);
  lut[(size_t)blockIdx.x * 256 + threadIdx.x] = (uint8_t)v;
}

// ---- CLAHE ----------------------------------------------------------------------------
// dst = lut_img[src]; histogram of every CLAHE tile of the (reflect-101 extended) clipped image.
// grid (tiles_x * tiles_y, images), one block per tile.
__global__ void __launch_bounds__(256) clip_tilehist_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                            int h, int w, const uint8_t* __restrict__ lut,
                                                            int tx, int th, int tw, unsigned* __restrict__ tilehist) {
  __shared__ unsigned hist[256];
  __shared__ uint8_t l[256];
  const int img = blockIdx.y, tile = blockIdx.x, tj = tile / tx, ti = tile - tj * tx;
  hist[threadIdx.x] = 0;
  l[threadIdx.x] = lut[(size_t)img * 256 + threadIdx.x];
  __syncthreads();
  const uint8_t* s = src + (size_t)img * h * w;
  uint8_t* d = dst + (size_t)img * h * w;
  for (int e = threadIdx.x; e < th * tw; e += 256) {
    const int ey = tj * th + e / tw, ex = ti * tw + e % tw;
    const int y = reflect101(ey, h), x = reflect101(ex, w);
    const uint8_t v = l[s[(size_t)y * w + x]];
    atomicAdd(&hist[v], 1u);
    if (ey < h && ex < w) d[(size_t)ey * w + ex] = v;
  }
  __syncthreads();
  tilehist[((size_t)img * gridDim.x + tile) * 256 + threadIdx.x] = hist[threadIdx.x];
}

// clip / redistribute / cumulative sum / scale of one tile histogram (CLAHE_CalcLut_Body).
__global__ void __launch_bounds__(256) clahe_lut_kernel(const unsigned* __restrict__ tilehist, int area, int clip,
                                                        uint8_t* __restrict__ luts) {
  __shared__ int hs[256];
  __shared__ int red[256];
  const unsigned* hsrc = tilehist + (size_t)blockIdx.x * 256;
  int v = (int)hsrc[threadIdx.x];
  if (clip > 0) {
    const int over = v > clip ? v - clip : 0;
    v = min(v, clip);
    red[threadIdx.x] = over;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
      __syncthreads();
    }
    const int clipped = red[0];
    const int batch = clipped / 256;
    int residual = clipped - batch * 256;
    v += batch;
    if (residual != 0) {
      const int step = max(256 / residual, 1);
      // bins 0, step, 2*step, ... get +1 while the residual lasts
      if (threadIdx.x % step == 0 && threadIdx.x / step < residual) v += 1;
    }
  }
  hs[threadIdx.x] = v;
  __syncthreads();
  for (int o = 1; o < 256; o <<= 1) {
    const int add = threadIdx.x >= o ? hs[threadIdx.x - o] : 0;
    __syncthreads();
    hs[threadIdx.x] += add;
    __syncthreads();
  }
  const float lut_scale = __fdiv_rn(255.0f, (float)area);
  luts[(size_t)blockIdx.x * 256 + threadIdx.x] = (uint8_t)cv_round_sat_u8(__fmul_rn((float)hs[threadIdx.x], lut_scale));
}

// bilinear blend of the four neighbouring tile LUTs (CLAHE_Interpolation_Body); per-image min / max.
__global__ void __launch_bounds__(256) clahe_apply_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                          int h, int w, const uint8_t* __restrict__ luts, int tx, int ty,
                                                          int th, int tw, unsigned* __restrict__ vmin,
                                                          unsigned* __restrict__ vmax) {
  const int img = blockIdx.y;
  const uint8_t* s = src + (size_t)img * h * w;
  uint8_t* d = dst + (size_t)img * h * w;
  const uint8_t* L = luts + (size_t)img * tx * ty * 256;
  const float inv_tw = __fdiv_rn(1.0f, (float)tw), inv_th = __fdiv_rn(1.0f, (float)th);
  const long long n = (long long)h * w, stride = (long long)gridDim.x * 256;
  unsigned lo = 255, hi = 0;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    const int y = (int)(i / w), x = (int)(i - (long long)y * w);
    const float txf = __fsub_rn(__fmul_rn((float)x, inv_tw), 0.5f), tyf = __fsub_rn(__fmul_rn((float)y, inv_th), 0.5f);
    int tx1 = (int)floorf(txf), ty1 = (int)floorf(tyf);
    const float xa = __fsub_rn(txf, (float)tx1), ya = __fsub_rn(tyf, (float)ty1);
    const float xa1 = __fsub_rn(1.0f, xa), ya1 = __fsub_rn(1.0f, ya);
    const int tx2 = min(tx1 + 1, tx - 1), ty2 = min(ty1 + 1, ty - 1);
    tx1 = max(tx1, 0);
    ty1 = max(ty1, 0);
    const int v = s[i];
    const float l11 = (float)L[((size_t)ty1 * tx + tx1) * 256 + v], l12 = (float)L[((size_t)ty1 * tx + tx2) * 256 + v];
    const float l21 = (float)L[((size_t)ty2 * tx + tx1) * 256 + v], l22 = (float)L[((size_t)ty2 * tx + tx2) * 256 + v];
    const float top = __fadd_rn(__fmul_rn(l11, xa1), __fmul_rn(l12, xa));
    const float bot = __fadd_rn(__fmul_rn(l21, xa1), __fmul_rn(l22, xa));
    const unsigned r = (unsigned)cv_round_sat_u8(__fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya)));
    d[i] = (uint8_t)r;
    lo = min(lo, r);
    hi = max(hi, r);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(vmin + img, lo);
    atomicMax(vmax + img, hi);
  }
}

// out = (v - min) / (uint8(max - min) + 1e-8) in float64, stored as float32 (dataset.py:41-42,62);
// without min/max tables: v / 255 in float32 (normalize=False, dataset.py:55).
__global__ void __launch_bounds__(256) normalize_f32_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst,
                                                            long long n, const unsigned* __restrict__ vmin,
                                                            const unsigned* __restrict__ vmax) {
  const int img = blockIdx.y;
  const uint8_t* s = src + (size_t)img * n;
  float* d = dst + (size_t)img * n;
  __shared__ float lut[256];
  if (vmin) {
    const int mn = (int)vmin[img], mx = (int)vmax[img];
    const double den = (double)((mx - mn) & 255) + 1e-8;
    lut[threadIdx.x] = (float)((double)((threadIdx.x - mn) & 255) / den);
  } else {
    lut[threadIdx.x] = __fdiv_rn((float)threadIdx.x, 255.0f);
  }
  __syncthreads();
  const long long stride = (long long)gridDim.x * 256;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) d[i] = lut[s[i]];
}

int pblocks(long long items, int images) {
  long long b = (items + 1023) / 1024;
  long long cap = ((long long)b2_num_sms() * 16 + images - 1) / images;
  if (cap < 1) cap = 1;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace

#define PRE_SHAPE(planes, sh, sw, dh, dw)                                                        \
  B2_REQUIRE((planes) > 0 && (planes) <= 65535 && (sh) > 0 && (sw) > 0 && (dh) > 0 && (dw) > 0,    \
             "preproc: bad shape %d x (%d x %d) -> (%d x %d)", planes, sh, sw, dh, dw)

extern "C" int b2seg_resize_nearest_u8(const uint8_t* src, int planes, int sh, int sw, uint8_t* dst,
                                       float* dst_binary_f32, int dh, int dw, const int32_t* xofs,
                                       const int32_t* yofs, void* stream) {
  B2_REQUIRE(src && (dst || dst_binary_f32) && xofs && yofs, "resize_nearest: null pointer");
  PRE_SHAPE(planes, sh, sw, dh, dw);
  resize_nearest_u8_kernel<<<dim3(pblocks((long long)dh * dw, planes), planes), 256, 0, (cudaStream_t)stream>>>(
      src, sh, sw, dst, dst_binary_f32, dh, dw, xofs, yofs);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_resize_linear_u8(const uint8_t* src, int planes, int sh, int sw, uint8_t* dst, int dh,
                                      int dw, const int32_t* xofs, const int16_t* xcoef, const int32_t* yofs,
                                      const int16_t* ycoef, void* stream) {
  B2_REQUIRE(src && dst && xofs && xcoef && yofs && ycoef, "resize_linear: null pointer");
  PRE_SHAPE(planes, sh, sw, dh, dw);
  resize_linear_u8_kernel<<<dim3(pblocks((long long)dh * dw, planes), planes), 256, 0, (cudaStream_t)stream>>>(
      src, sh, sw, dst, dh, dw, xofs, xcoef, yofs, ycoef);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_resize_areafast_u8(const uint8_t* src, int planes, int sh, int sw, uint8_t* dst, int dh,
                                        int dw, void* stream) {
  B2_REQUIRE(src && dst, "resize_areafast: null pointer");
  PRE_SHAPE(planes, sh, sw, dh, dw);
  B2_REQUIRE(sw % dw == 0 && sh % dh == 0, "resize_areafast: %d x %d is not an integer multiple of %d x %d", sh, sw,
             dh, dw);
  resize_areafast_u8_kernel<<<dim3(pblocks((long long)dh * dw, planes), planes), 256, 0, (cudaStream_t)stream>>>(
      src, sh, sw, dst, dh, dw, sw / dw, sh / dh);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_resize_area_u8(const uint8_t* src, int planes, int sh, int sw, uint8_t* dst, int dh, int dw,
                                    const int32_t* xstart, const int32_t* xsi, const float* xalpha,
                                    const int32_t* ystart, const int32_t* ysi, const float* ybeta, void* stream) {
  B2_REQUIRE(src && dst && xstart && xsi && xalpha && ystart && ysi && ybeta, "resize_area: null pointer");
  PRE_SHAPE(planes, sh, sw, dh, dw);
  resize_area_u8_kernel<<<dim3(pblocks((long long)dh * dw, planes), planes), 256, 0, (cudaStream_t)stream>>>(
      src, sh, sw, dst, dh, dw, xstart, xsi, xalpha, ystart, ysi, ybeta);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_pre_percentile_lut(const uint32_t* hist, int n_images, int64_t pixels_per_image, double q_lo,
                                        double q_hi, uint8_t* lut, double* percentiles, void* stream) {
  B2_REQUIRE(hist && lut && n_images > 0 && pixels_per_image > 0, "percentile_lut: bad arguments");
  B2_REQUIRE(q_lo >= 0.0 && q_lo <= q_hi && q_hi <= 100.0, "percentile_lut: bad percentiles");
  percentile_lut_kernel<<<n_images, 256, 0, (cudaStream_t)stream>>>(hist, pixels_per_image, q_lo, q_hi, lut,
                                                                     percentiles);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

/* scratch layout: tilehist u32 [n][tiles][256] | luts u8 [n][tiles][256] | vmin u32 [n] | vmax u32 [n] */
extern "C" size_t b2seg_pre_clahe_workspace_bytes(int n_images, int tiles_x, int tiles_y) {
  if (n_images <= 0 || tiles_x <= 0 || tiles_y <= 0) return 0;
  const size_t t = (size_t)n_images * tiles_x * tiles_y * 256;
  return t * 4 + t + (size_t)n_images * 8 + 256;
}

/* clipped = lut[src] (per image); eq = CLAHE(clipped); out = min-max normalised fp32.
 * `clipped` and `eq` are uint8 [n][h][w] buffers of the caller (eq may alias src). */
extern "C" int b2seg_pre_clahe_normalize(const uint8_t* src, int n_images, int h, int w, const uint8_t* clip_lut,
                                         double clip_limit, int tiles_x, int tiles_y, uint8_t* clipped,
                                         uint8_t* eq, float* out, void* workspace, size_t workspace_bytes,
                                         void* stream) {
  B2_REQUIRE(src && clip_lut && clipped && eq && out && workspace, "clahe_normalize: null pointer");
  B2_REQUIRE(n_images > 0 && n_images <= 65535 && h > 0 && w > 0 && tiles_x > 0 && tiles_y > 0 &&
                 tiles_x * tiles_y <= 65535, "clahe_normalize: bad shape");
  B2_REQUIRE(workspace_bytes >= b2seg_pre_clahe_workspace_bytes(n_images, tiles_x, tiles_y),
             "clahe_normalize: workspace too small");
  B2_REQUIRE(clipped != src && clipped != eq, "clahe_normalize: `clipped` must be its own buffer");
  cudaStream_t st = (cudaStream_t)stream;
  const int tiles = tiles_x * tiles_y;
  const int eh = h % tiles_y == 0 ? h : h + (tiles_y - h % tiles_y);
  const int ew = w % tiles_x == 0 ? w : w + (tiles_x - w % tiles_x);
  const int th = eh / tiles_y, tw = ew / tiles_x, area = th * tw;
  int clip = 0;
  if (clip_limit > 0.0) {
    clip = (int)(clip_limit * area / 256);
    if (clip < 1) clip = 1;
  }
  unsigned* tilehist = (unsigned*)workspace;
  uint8_t* luts = (uint8_t*)(tilehist + (size_t)n_images * tiles * 256);
  unsigned* vmin = (unsigned*)(luts + (((size_t)n_images * tiles * 256 + 15) & ~(size_t)15));
  unsigned* vmax = vmin + n_images;
  clip_tilehist_kernel<<<dim3(tiles, n_images), 256, 0, st>>>(src, clipped, h, w, clip_lut, tiles_x, th, tw, tilehist);
  B2_CHECK_CUDA(cudaGetLastError());
  clahe_lut_kernel<<<n_images * tiles, 256, 0, st>>>(tilehist, area, clip, luts);
  B2_CHECK_CUDA(cudaGetLastError());
  B2_CHECK_CUDA(cudaMemsetAsync(vmin, 0xFF, (size_t)n_images * 4, st));
  B2_CHECK_CUDA(cudaMemsetAsync(vmax, 0, (size_t)n_images * 4, st));
  clahe_apply_kernel<<<dim3(pblocks((long long)h * w, n_images), n_images), 256, 0, st>>>(
      clipped, eq, h, w, luts, tiles_x, tiles_y, th, tw, vmin, vmax);
  B2_CHECK_CUDA(cudaGetLastError());
  normalize_f32_kernel<<<dim3(pblocks((long long)h * w, n_images), n_images), 256, 0, st>>>(
      eq, out, (long long)h * w, vmin, vmax);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_pre_scale255_f32(const uint8_t* src, int n_images, int64_t pixels_per_image, float* out,
                                      void* stream) {
  B2_REQUIRE(src && out && n_images > 0 && n_images <= 65535 && pixels_per_image > 0, "scale255: bad arguments");
  normalize_f32_kernel<<<dim3(pblocks(pixels_per_image, n_images), n_images), 256, 0, (cudaStream_t)stream>>>(
      src, out, pixels_per_image, nullptr, nullptr);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}
