// Target-only weight maps of the boundary-style losses, on the device.
//
//   BoundaryLoss (losses.py:283-301):  1 + 10 * gauss_theta( dilate_theta0(t) - erode_theta0(t) )
//     -- one kernel: the tile and its halo are staged in shared memory once, the morphological
//        gradient is formed there and the theta x theta Gaussian (zero padding, conv2d semantics)
//        is applied from shared memory.  8 B/px algorithmic (read t, write w).
//   DistanceMapPenalizedCrossEntropy (losses.py:216-232) and HausdorffDistanceLoss._compute_dtm
//     (losses.py:262-273): scipy.ndimage.distance_transform_edt, the exact Euclidean distance of
//     every non-zero pixel to the nearest zero pixel.  Two separable passes on integers (column
//     scan, then a pruned exact lower envelope per row), so the squared distances are exact and the
//     result is bit-identical to scipy's float64 computation rounded to float32.  scipy's
//     behaviour for a plane WITHOUT any zero pixel (distances to an imaginary pixel at row -1,
//     column 0) is reproduced, because Hausdorff feeds it sigmoid outputs, which have no zeros.
#include "common.cuh"

namespace {

constexpr int kTile = 32;
constexpr int kMaxTheta = 15;
constexpr int kInf = 1 << 14;  // "no zero in this column": (1<<14)^2 + w^2 still fits int32

struct GaussK {
  float w[kMaxTheta * kMaxTheta];
};

__global__ void __launch_bounds__(256)
    boundary_weight_kernel(const float* __restrict__ t, float* __restrict__ out, int h, int w, int r0,
                           int r1, GaussK gk, float gain) {
  extern __shared__ float sm[];
  const int R = r0 + r1, tw = kTile + 2 * R, ew = kTile + 2 * r1, th = 2 * r1 + 1;
  float* T = sm;             // [tw][tw] target with halo
  float* E = sm + tw * tw;   // [ew][ew] dilate - erode, 0 outside the image
  const long long plane = blockIdx.z;
  const float* tp = t + plane * h * w;
  float* op = out + plane * h * w;
  const int x0 = blockIdx.x * kTile, y0 = blockIdx.y * kTile;
  for (int i = threadIdx.x; i < tw * tw; i += blockDim.x) {
    const int yy = y0 - R + i / tw, xx = x0 - R + i % tw;
    T[i] = (yy >= 0 && yy < h && xx >= 0 && xx < w) ? tp[(long long)yy * w + xx] : 0.0f;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < ew * ew; i += blockDim.x) {
    const int ey = i / ew, ex = i % ew;
    const int gy = y0 - r1 + ey, gx = x0 - r1 + ex;
    float v = 0.0f;
    if (gy >= 0 && gy < h && gx >= 0 && gx < w) {
      // max_pool2d pads with -inf: the window is clipped to the image
      const int ya = max(gy - r0, 0), yb = min(gy + r0, h - 1);
      const int xa = max(gx - r0, 0), xb = min(gx + r0, w - 1);
      float mx = -INFINITY, mn = INFINITY;
      for (int yy = ya; yy <= yb; ++yy)
        for (int xx = xa; xx <= xb; ++xx) {
          const float q = T[(yy - (y0 - R)) * tw + (xx - (x0 - R))];
          mx = fmaxf(mx, q);
          mn = fminf(mn, q);
        }
      v = mx - mn;
    }
    E[i] = v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kTile * kTile; i += blockDim.x) {
    const int oy = i / kTile, ox = i % kTile;
    const int gy = y0 + oy, gx = x0 + ox;
    if (gy >= h || gx >= w) continue;
    float s = 0.0f;
    for (int ky = 0; ky < th; ++ky)
      for (int kx = 0; kx < th; ++kx) s = fmaf(gk.w[ky * th + kx], E[(oy + ky) * ew + ox + kx], s);
    op[(long long)gy * w + gx] = __fadd_rn(1.0f, __fmul_rn(gain, s));
  }
}

// ---- exact Euclidean distance transform ---------------------------------------
struct EdtPlane {
  int max_d2;
  int has_zero;
};

// Pass 1: one thread per column; g(y, x) = distance to the nearest zero pixel of the same column.
__global__ void __launch_bounds__(128)
    edt_columns_kernel(const float* __restrict__ src, int* __restrict__ g, EdtPlane* __restrict__ pl,
                       int planes, int h, int w) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, p = blockIdx.y;
  if (x >= w) return;
  const float* s = src + (long long)p * h * w + x;
  int* gp = g + (long long)p * h * w + x;
  int d = kInf;
  bool any = false;
  for (int y = 0; y < h; ++y) {
    const bool zero = s[(long long)y * w] == 0.0f;
    any |= zero;
    d = zero ? 0 : min(d + 1, kInf);
    gp[(long long)y * w] = d;
  }
  d = kInf;
  for (int y = h - 1; y >= 0; --y) {
    const int cur = gp[(long long)y * w];
    d = cur == 0 ? 0 : min(d + 1, kInf);
    if (d < cur) gp[(long long)y * w] = d;
  }
  if (any) pl[p].has_zero = 1;
}

// Pass 2: one block per row; d2(x) = min over x' of (x - x')^2 + g(x')^2, searched outwards from x
// and stopped as soon as the horizontal offset alone exceeds the best squared distance.
__global__ void __launch_bounds__(256)
    edt_rows_kernel(int* __restrict__ g, EdtPlane* __restrict__ pl, int h, int w) {
  extern __shared__ int row[];
  const int y = blockIdx.x, p = blockIdx.y;
  int* gp = g + ((long long)p * h + y) * w;
  for (int x = threadIdx.x; x < w; x += blockDim.x) row[x] = gp[x];
  __syncthreads();
  const bool has_zero = pl[p].has_zero != 0;
  int best_max = 0;
  for (int x = threadIdx.x; x < w; x += blockDim.x) {
    int best;
    if (!has_zero) {
      best = (y + 1) * (y + 1) + x * x;  // scipy: an imaginary background pixel at (-1, 0)
    } else {
      const int g0 = row[x];
      best = g0 * g0;
      for (int d = 1; d * d < best; ++d) {
        if (x - d >= 0) {
          const int q = row[x - d];
          best = min(best, d * d + q * q);
        }
        if (x + d < w) {
          const int q = row[x + d];
          best = min(best, d * d + q * q);
        }
        if (x - d < 0 && x + d >= w) break;
      }
    }
    gp[x] = best;
    best_max = max(best_max, best);
  }
  for (int o = 16; o > 0; o >>= 1) best_max = max(best_max, __shfl_xor_sync(0xffffffffu, best_max, o));
  if ((threadIdx.x & 31) == 0 && best_max > 0) atomicMax(&pl[p].max_d2, best_max);
}

// Pass 3: mode 0 -> float32(sqrt(d2)); mode 1 -> 1 + alpha * float32(sqrt(d2) / (sqrt(max_d2) + 1e-7))
// with the division in float64 exactly as numpy does it before the cast (losses.py:226-231).
__global__ void __launch_bounds__(256)
    edt_finish_kernel(const int* __restrict__ d2, const EdtPlane* __restrict__ pl, float* __restrict__ out,
                      long long plane_px, long long total, int mode, float alpha) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const double d = sqrt((double)d2[i]);
    if (mode == 0) {
      out[i] = (float)d;
    } else {
      const double mx = sqrt((double)pl[i / plane_px].max_d2);
      out[i] = __fadd_rn(1.0f, __fmul_rn(alpha, (float)(d / (mx + 1e-7))));
    }
  }
}

}  // namespace

extern "C" int b2seg_loss_boundary_weight_map(const float* target, float* weight, int planes, int h,
                                              int w, int theta0, int theta, const float* gauss,
                                              float gain, void* stream) {
  B2_REQUIRE(target && weight && gauss, "boundary_weight_map: null argument");
  B2_REQUIRE(planes > 0 && h > 0 && w > 0, "boundary_weight_map: bad shape %d x %d x %d", planes, h, w);
  B2_REQUIRE(theta0 >= 1 && theta >= 1 && (theta0 & 1) && (theta & 1) && theta0 <= kMaxTheta && theta <= kMaxTheta,
             "boundary_weight_map: theta0 / theta must be odd and <= %d (got %d, %d); even sizes change "
             "the map's shape in the reference as well", kMaxTheta, theta0, theta);
  B2_REQUIRE(planes <= 65535, "boundary_weight_map: at most 65535 planes per call");
  GaussK gk;
  for (int i = 0; i < theta * theta; ++i) gk.w[i] = gauss[i];
  const int r0 = theta0 / 2, r1 = theta / 2;
  const int tw = kTile + 2 * (r0 + r1), ew = kTile + 2 * r1;
  const size_t smem = sizeof(float) * (size_t)(tw * tw + ew * ew);
  dim3 grid((w + kTile - 1) / kTile, (h + kTile - 1) / kTile, planes);
  boundary_weight_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(target, weight, h, w, r0, r1, gk, gain);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" size_t b2seg_loss_edt_workspace_bytes(int planes, int h, int w) {
  if (planes <= 0 || h <= 0 || w <= 0) return 0;
  return sizeof(int) * (size_t)planes * h * w + sizeof(EdtPlane) * (size_t)planes;
}

extern "C" int b2seg_loss_edt_f32(const float* src, float* out, int planes, int h, int w, int mode,
                                  float alpha, void* workspace, void* stream) {
  B2_REQUIRE(src && out && workspace, "edt: null argument");
  B2_REQUIRE(planes > 0 && planes <= 65535 && h > 0 && w > 0 && h < kInf && w < kInf,
             "edt: bad shape %d x %d x %d (sides < %d, planes <= 65535)", planes, h, w, kInf);
  B2_REQUIRE(mode == 0 || mode == 1, "edt: mode must be 0 (distance) or 1 (1 + alpha * d / max), got %d", mode);
  cudaStream_t st = (cudaStream_t)stream;
  int* g = (int*)workspace;
  EdtPlane* pl = (EdtPlane*)(g + (size_t)planes * h * w);
  B2_CHECK_CUDA(cudaMemsetAsync(pl, 0, sizeof(EdtPlane) * (size_t)planes, st));
  edt_columns_kernel<<<dim3((w + 127) / 128, planes), 128, 0, st>>>(src, g, pl, planes, h, w);
  edt_rows_kernel<<<dim3(h, planes), 256, sizeof(int) * (size_t)w, st>>>(g, pl, h, w);
  const long long total = (long long)planes * h * w;
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)b2_num_sms() * 16) blocks = (long long)b2_num_sms() * 16;
  edt_finish_kernel<<<(int)blocks, 256, 0, st>>>(g, pl, out, (long long)h * w, total, mode, alpha);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}
