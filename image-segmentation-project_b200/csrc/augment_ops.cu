// Non-geometric augmentation kernels on uint8 planes and the helpers of the
// device-side map builders.  They stand behind the albumentations names the reference's
// augmentation search instantiates (forward_selection_integration.py:107-178): InvertImg, Solarize,
// RandomGamma, ColorJitter (look-up tables), GaussNoise (counter-based Philox + Box-Muller in the
// kernel), CoarseDropout / GridDropout (rectangle fills), AdvancedBlur (cv2.filter2D arithmetic),
// CropAndPad (cv2.copyMakeBorder), and ElasticTransform's displacement field (separable Gaussian of
// uniform noise).  All are HBM-bound streaming kernels: 16-byte accesses where the layout allows.
#include "common.cuh"

namespace {

// ------------------------------------------------------------------- LUT ----
__device__ __forceinline__ uint32_t lut4(const uint8_t* lut, uint32_t v) {
  return (uint32_t)lut[v & 255] | ((uint32_t)lut[(v >> 8) & 255] << 8) |
         ((uint32_t)lut[(v >> 16) & 255] << 16) | ((uint32_t)lut[v >> 24] << 24);
}
__global__ void __launch_bounds__(256) lut_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                     long long n, const uint8_t* __restrict__ lut_g, int vec) {
  __shared__ uint8_t lut[256];
  lut[threadIdx.x] = lut_g[threadIdx.x];
  __syncthreads();
  const long long stride = (long long)gridDim.x * 256, t = (long long)blockIdx.x * 256 + threadIdx.x;
  long long done = 0;
  if (vec) {
    const long long n16 = n >> 4;
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    for (long long i = t; i < n16; i += stride) {
      uint4 v = ldg_stream_u4(s4 + i);
      v.x = lut4(lut, v.x); v.y = lut4(lut, v.y); v.z = lut4(lut, v.z); v.w = lut4(lut, v.w);
      stg_stream_u4(d4 + i, v);
    }
    done = n16 << 4;
  }
  for (long long i = done + t; i < n; i += stride) dst[i] = lut[src[i]];
}

// 256-bin histograms of `nimg` images of `n` bytes each: per-block shared-memory privatisation,
// one global atomic per (block, bin).  out[nimg][256] must be zeroed by the caller.
__global__ void __launch_bounds__(256) hist_u8_kernel(const uint8_t* __restrict__ src, long long n,
                                                      unsigned* __restrict__ out, int vec) {
  __shared__ unsigned h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const uint8_t* s = src + (size_t)blockIdx.y * n;
  const long long stride = (long long)gridDim.x * 256, t = (long long)blockIdx.x * 256 + threadIdx.x;
  long long done = 0;
  if (vec) {
    const long long n16 = n >> 4;
    const uint4* s4 = reinterpret_cast<const uint4*>(s);
    for (long long i = t; i < n16; i += stride) {
      const uint4 v = ldg_stream_u4(s4 + i);
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        atomicAdd(&h[w[k] & 255], 1u); atomicAdd(&h[(w[k] >> 8) & 255], 1u);
        atomicAdd(&h[(w[k] >> 16) & 255], 1u); atomicAdd(&h[w[k] >> 24], 1u);
      }
    }
    done = n16 << 4;
  }
  for (long long i = done + t; i < n; i += stride) atomicAdd(&h[s[i]], 1u);
  __syncthreads();
  if (h[threadIdx.x]) atomicAdd(out + (size_t)blockIdx.y * 256 + threadIdx.x, h[threadIdx.x]);
}

// lut[v] = uint8(clip(v*factor + mean*mean_coef, 0, 255)) with mean from the histogram (exact
// integer sums) -- albumentations' uint8 brightness (mean_coef = 0) and contrast
// (mean_coef = 1 - factor) adjustments -- applied in the same launch.
__global__ void __launch_bounds__(256) affine_lut_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                            long long n, const unsigned* __restrict__ hist,
                                                            double factor, double mean_coef) {
  __shared__ uint8_t lut[256];
  __shared__ double red[256];
  __shared__ double cnt[256];
  red[threadIdx.x] = hist ? (double)hist[threadIdx.x] * (double)threadIdx.x : 0.0;
  cnt[threadIdx.x] = hist ? (double)hist[threadIdx.x] : 0.0;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { red[threadIdx.x] += red[threadIdx.x + o]; cnt[threadIdx.x] += cnt[threadIdx.x + o]; }
    __syncthreads();
  }
  const double mean = cnt[0] > 0.0 ? red[0] / cnt[0] : 0.0;
  double v = (double)threadIdx.x * factor;
  v = v + mean * mean_coef;
  v = v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v);
  lut[threadIdx.x] = (uint8_t)v;  // truncation, as ndarray.astype(uint8)
  __syncthreads();
  const long long stride = (long long)gridDim.x * 256;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) dst[i] = lut[src[i]];
}

// ------------------------------------------------------------- rectangles ----
__global__ void __launch_bounds__(256) fill_rects_u8_kernel(uint8_t* __restrict__ plane, int h, int w,
                                                            const int* __restrict__ rects, int value) {
  const int y1 = max(0, rects[4 * blockIdx.x]), x1 = max(0, rects[4 * blockIdx.x + 1]);
  const int y2 = min(h, rects[4 * blockIdx.x + 2]), x2 = min(w, rects[4 * blockIdx.x + 3]);
  const int rw = x2 - x1, rh = y2 - y1;
  if (rw <= 0 || rh <= 0) return;
  for (int e = threadIdx.x; e < rw * rh; e += 256) plane[(size_t)(y1 + e / rw) * w + x1 + e % rw] = (uint8_t)value;
}

// ------------------------------------------------------------------ noise ----
// Philox-4x32-10 (Salmon et al.), counter = element block index, key = seed.
__device__ __forceinline__ uint4 philox4x32(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ float u01(uint32_t x) { return ((float)(x >> 8) + 0.5f) * (1.0f / 16777216.0f); }
__device__ __forceinline__ void normal4(long long block, unsigned long long seed, float (&z)[4]) {
  const uint4 r = philox4x32(make_uint4((uint32_t)block, (uint32_t)(block >> 32), 0x5eed0001u, 0u),
                             make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
  const float r0 = sqrtf(-2.0f * logf(u01(r.x))), r1 = sqrtf(-2.0f * logf(u01(r.z)));
  float s0, c0, s1, c1;
  sincospif(2.0f * u01(r.y), &s0, &c0);
  sincospif(2.0f * u01(r.w), &s1, &c1);
  z[0] = r0 * c0; z[1] = r0 * s0; z[2] = r1 * c1; z[3] = r1 * s1;
}
// dst = uint8(clip(float(src) + mean + sigma*N(0,1), 0, 255)) -- A.GaussNoise on a uint8 plane
// (float32 sum, clip, truncating cast).
__global__ void __launch_bounds__(256) gauss_noise_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                             long long n, float mean, float sigma,
                                                             unsigned long long seed) {
  const long long nb = (n + 3) >> 2, stride = (long long)gridDim.x * 256;
  for (long long b = (long long)blockIdx.x * 256 + threadIdx.x; b < nb; b += stride) {
    float z[4];
    normal4(b, seed, z);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const long long i = 4 * b + e;
      if (i < n) {
        float v = (float)src[i] + (mean + sigma * z[e]);
        v = fminf(fmaxf(v, 0.0f), 255.0f);
        dst[i] = (uint8_t)v;
      }
    }
  }
}
// uniform(-1, 1) float field (ElasticTransform's raw displacement noise)
__global__ void __launch_bounds__(256) uniform_f32_kernel(float* __restrict__ dst, long long n,
                                                          unsigned long long seed) {
  const long long nb = (n + 3) >> 2, stride = (long long)gridDim.x * 256;
  for (long long b = (long long)blockIdx.x * 256 + threadIdx.x; b < nb; b += stride) {
    const uint4 r = philox4x32(make_uint4((uint32_t)b, (uint32_t)(b >> 32), 0x5eed0002u, 0u),
                               make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (4 * b + e < n) dst[4 * b + e] = 2.0f * u01(w[e]) - 1.0f;
  }
}

// ----------------------------------------------------------------- border ----
__device__ __forceinline__ int reflect_idx(int p, int len, int delta /*0 REFLECT, 1 REFLECT_101*/) {
  if (len == 1) return 0;
  while ((unsigned)p >= (unsigned)len) {
    if (p < 0) p = -p - 1 + delta;
    else p = len - 1 - (p - len) - delta;
  }
  return p;
}

// ---------------------------------------------------------------- filter2D ----
// cv2.filter2D(src, -1, kernel) for a k x k float32 kernel (k odd, <= 7), anchor at the centre,
// BORDER_REFLECT_101: correlation accumulated in float in kernel raster order (fma chain, as
// OpenCV's vectorised FilterVec_8u does), rounded to nearest-even and saturated.
constexpr int FT = 32, FK = 7;
__global__ void __launch_bounds__(256) filter2d_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                          int h, int w, const float* __restrict__ kern, int k) {
  __shared__ uint8_t tile[FT + FK - 1][FT + FK - 1 + 2];
  __shared__ float kf[FK * FK];
  const int rad = k >> 1, x0 = blockIdx.x * FT, y0 = blockIdx.y * FT;
  const uint8_t* s = src + (size_t)blockIdx.z * h * w;
  uint8_t* d = dst + (size_t)blockIdx.z * h * w;
  if (threadIdx.x < k * k) kf[threadIdx.x] = kern[threadIdx.x];
  const int span = FT + 2 * rad;
  for (int e = threadIdx.x; e < span * span; e += 256) {
    const int ty = e / span, tx = e - ty * span;
    tile[ty][tx] = s[(size_t)reflect_idx(y0 + ty - rad, h, 1) * w + reflect_idx(x0 + tx - rad, w, 1)];
  }
  __syncthreads();
  for (int e = threadIdx.x; e < FT * FT; e += 256) {
    const int ly = e / FT, lx = e - ly * FT;
    if (y0 + ly >= h || x0 + lx >= w) continue;
    float acc = 0.0f;
    for (int i = 0; i < k; ++i)
      for (int j = 0; j < k; ++j) {
        const float c = kf[i * k + j];
        if (c != 0.0f) acc = __fmaf_rn((float)tile[ly + i][lx + j], c, acc);
      }
    const int r = __float2int_rn(acc);
    d[(size_t)(y0 + ly) * w + x0 + lx] = (uint8_t)max(0, min(255, r));
  }
}

// ----------------------------------------------------------- copyMakeBorder ----
__global__ void __launch_bounds__(256) pad_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                     int h, int w, int top, int left, int oh, int ow,
                                                     int border /*B2SEG_BORDER_*/, int cval) {
  const uint8_t* s = src + (size_t)blockIdx.y * h * w;
  uint8_t* d = dst + (size_t)blockIdx.y * oh * ow;
  const long long n = (long long)oh * ow, stride = (long long)gridDim.x * 256;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    const int y = (int)(i / ow) - top, x = (int)(i % ow) - left;
    uint8_t v;
    if ((unsigned)y < (unsigned)h && (unsigned)x < (unsigned)w) v = s[(size_t)y * w + x];
    else if (border == B2SEG_BORDER_CONSTANT) v = (uint8_t)cval;
    else v = s[(size_t)reflect_idx(y, h, border == B2SEG_BORDER_REFLECT101) * w +
               reflect_idx(x, w, border == B2SEG_BORDER_REFLECT101)];
    d[i] = v;
  }
}

// ------------------------------------------------- separable Gaussian, float ----
// scipy.ndimage.gaussian_filter(mode='reflect', truncate=4): 1-D correlation with
// w[i] = exp(-0.5 i^2 / sigma^2) / sum, radius = int(4 sigma + 0.5), along x then along y.
// out = base(x or y coordinate) * use_coord + gain * blur(in)  (the elastic map in one pass).
__global__ void __launch_bounds__(256) blur1d_f32_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                         int h, int w, const float* __restrict__ wts, int rad,
                                                         int along_x, float gain, int add_coord) {
  const long long n = (long long)h * w, stride = (long long)gridDim.x * 256;
  const float* s = in + (size_t)blockIdx.y * n;
  float* d = out + (size_t)blockIdx.y * n;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride) {
    const int y = (int)(i / w), x = (int)(i % w);
    float acc = 0.0f;
    if (along_x) {
      for (int t = -rad; t <= rad; ++t) acc = fmaf(s[(size_t)y * w + reflect_idx(x + t, w, 0)], wts[t + rad], acc);
    } else {
      for (int t = -rad; t <= rad; ++t) acc = fmaf(s[(size_t)reflect_idx(y + t, h, 0) * w + x], wts[t + rad], acc);
    }
    float v = gain * acc;
    if (add_coord == 1) v += (float)x;
    else if (add_coord == 2) v += (float)y;
    d[i] = v;
  }
}

int blocks_for(long long items, int per_thread) {
  long long b = (items + 256LL * per_thread - 1) / (256LL * per_thread);
  const long long cap = (long long)b2_num_sms() * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}
bool al16(const void* a, const void* b) {
  return ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0;
}

}  // namespace

extern "C" int b2seg_aug_lut_u8(const uint8_t* src, uint8_t* dst, int64_t n, const uint8_t* lut, void* stream) {
  B2_REQUIRE(src && dst && lut && n > 0, "aug_lut: bad arguments");
  lut_u8_kernel<<<blocks_for(n, 64), 256, 0, (cudaStream_t)stream>>>(src, dst, n, lut, al16(src, dst) ? 1 : 0);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_aug_hist_u8(const uint8_t* src, int64_t n_per_image, int n_images, uint32_t* hist,
                                 void* stream) {
  B2_REQUIRE(src && hist && n_per_image > 0 && n_images > 0 && n_images <= 65535, "aug_hist: bad arguments");
  B2_CHECK_CUDA(cudaMemsetAsync(hist, 0, (size_t)n_images * 256 * sizeof(uint32_t), (cudaStream_t)stream));
  int bx = blocks_for(n_per_image, 256);
  const int cap = (b2_num_sms() * 16 + n_images - 1) / n_images;
  if (bx > cap) bx = cap < 1 ? 1 : cap;
  const int vec = (n_per_image % 16 == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
  hist_u8_kernel<<<dim3(bx, n_images), 256, 0, (cudaStream_t)stream>>>(src, n_per_image, hist, vec);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_aug_affine_lut_u8(const uint8_t* src, uint8_t* dst, int64_t n, const uint32_t* hist,
                                       double factor, double mean_coef, void* stream) {
  B2_REQUIRE(src && dst && n > 0, "aug_affine_lut: bad arguments");
  B2_REQUIRE(hist || mean_coef == 0.0, "aug_affine_lut: a mean term needs the histogram");
  affine_lut_u8_kernel<<<blocks_for(n, 16), 256, 0, (cudaStream_t)stream>>>(src, dst, n, hist, factor, mean_coef);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_aug_fill_rects_u8(uint8_t* plane, int h, int w, const int32_t* rects, int n_rects,
                                       int value, void* stream) {
  B2_REQUIRE(plane && h > 0 && w > 0 && (rects || n_rects == 0) && n_rects >= 0, "aug_fill_rects: bad arguments");
  if (n_rects == 0) return B2SEG_OK;
  fill_rects_u8_kernel<<<n_rects, 256, 0, (cudaStream_t)stream>>>(plane, h, w, rects, value & 255);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_aug_gauss_noise_u8(const uint8_t* src, uint8_t* dst, int64_t n, float mean, float sigma,
                                        uint64_t seed, void* stream) {
  B2_REQUIRE(src && dst && n > 0 && sigma >= 0.0f, "aug_gauss_noise: bad arguments");
  gauss_noise_u8_kernel<<<blocks_for(n, 16), 256, 0, (cudaStream_t)stream>>>(src, dst, n, mean, sigma, seed);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_aug_filter2d_u8(const uint8_t* src, uint8_t* dst, int planes, int h, int w,
                                     const float* kernel, int k, void* stream) {
  B2_REQUIRE(src && dst && kernel && src != dst, "aug_filter2d: bad pointers");
  B2_REQUIRE(planes > 0 && planes <= 65535 && h > 0 && w > 0, "aug_filter2d: bad shape");
  B2_REQUIRE(k >= 1 && k <= FK && (k & 1), "aug_filter2d: kernel size %d must be odd and <= %d", k, FK);
  dim3 grid((w + FT - 1) / FT, (h + FT - 1) / FT, planes);
  filter2d_u8_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst, h, w, kernel, k);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_aug_pad_u8(const uint8_t* src, uint8_t* dst, int planes, int h, int w, int top, int bottom,
                                int left, int right, int border, int border_value, void* stream) {
  B2_REQUIRE(src && dst && planes > 0 && planes <= 65535 && h > 0 && w > 0, "aug_pad: bad arguments");
  B2_REQUIRE(top >= 0 && bottom >= 0 && left >= 0 && right >= 0, "aug_pad: negative padding");
  B2_REQUIRE(border >= 0 && border <= 2, "aug_pad: border %d unsupported", border);
  const int oh = h + top + bottom, ow = w + left + right;
  pad_u8_kernel<<<dim3(blocks_for((long long)oh * ow, 8), planes), 256, 0, (cudaStream_t)stream>>>(
      src, dst, h, w, top, left, oh, ow, border, border_value & 255);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_aug_uniform_f32(float* dst, int64_t n, uint64_t seed, void* stream) {
  B2_REQUIRE(dst && n > 0, "aug_uniform: bad arguments");
  uniform_f32_kernel<<<blocks_for(n, 16), 256, 0, (cudaStream_t)stream>>>(dst, n, seed);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_aug_blur1d_f32(const float* src, float* dst, int planes, int h, int w, const float* weights,
                                    int radius, int along_x, float gain, int add_coord, void* stream) {
  B2_REQUIRE(src && dst && weights && src != dst, "aug_blur1d: bad pointers");
  B2_REQUIRE(planes > 0 && planes <= 65535 && h > 0 && w > 0 && radius >= 0, "aug_blur1d: bad shape");
  blur1d_f32_kernel<<<dim3(blocks_for((long long)h * w, 4), planes), 256, 0, (cudaStream_t)stream>>>(
      src, dst, h, w, weights, radius, along_x, gain, add_coord);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}
