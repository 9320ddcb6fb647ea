// Geometric augmentation as shared-memory-tiled gathers on uint8 planes.
//
// The reference applies albumentations transforms (dataset.py:148-154,184;
// forward_selection_integration.py:73-105) that bottom out in cv2.flip /
// np.rot90 / transpose / cv2.warpAffine / cv2.remap.  These kernels restate the
// cv2 arithmetic exactly: index permutations are bit-exact; warps use cv2's
// fixed-point scheme (coordinates in 1/32 px, 15-bit bilinear weights, constant
// / reflect / reflect101 borders).
//
// One CTA produces a 32x32 destination tile.  It first computes the source
// footprint of the tile, stages that footprint (with the border rule already
// applied) in shared memory with coalesced loads, and then gathers from shared
// memory.  Footprints that do not fit fall back to direct global gathers.
#include "common.cuh"

namespace {

constexpr int TILE = 32;
constexpr int FOOT = 72;  // staged footprint is at most FOOT x FOOT source pixels

// ---- exact permutations -----------------------------------------------------------
// dst(i,j) = src(si,sj); for transposing modes the tile is staged through smem so
// that both the global read and the global write are row-contiguous.
__global__ void __launch_bounds__(256)
    permute_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int h, int w,
                      int mode /*0 h,1 v,2 rot180,3 rot90(k=1),4 rot270(k=3),5 transpose*/) {
  __shared__ uint8_t tile[TILE][TILE + 4];
  const int plane = blockIdx.z;
  const bool swap = mode >= 3;
  const int dh = swap ? w : h, dw = swap ? h : w;
  const uint8_t* s = src + (size_t)plane * h * w;
  uint8_t* d = dst + (size_t)plane * h * w;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const int dj0 = blockIdx.x * TILE, di0 = blockIdx.y * TILE;

  if (!swap) {
#pragma unroll
    for (int r = ty; r < TILE; r += 8) {
      const int di = di0 + r, dj = dj0 + tx;
      if (di < dh && dj < dw) {
        const int si = (mode == 0) ? di : h - 1 - di;
        const int sj = (mode == 1) ? dj : w - 1 - dj;
        d[(size_t)di * dw + dj] = s[(size_t)si * w + sj];
      }
    }
    return;
  }
  // Transposing family.  dst(i,j) = src(si,sj) with
  //   rot90 k=1 : si = j,       sj = w-1-i
  //   rot90 k=3 : si = h-1-j,   sj = i
  //   transpose : si = j,       sj = i
  // Source tile rows run along dst columns: read src rows contiguously.
  // For the tile, dst i in [di0,di0+32), dst j in [dj0,dj0+32).
#pragma unroll
  for (int r = ty; r < TILE; r += 8) {
    // r indexes dst j (-> src row), tx indexes dst i (-> src column)
    const int dj = dj0 + r, di = di0 + tx;
    int si, sj;
    if (mode == 3) { si = dj; sj = w - 1 - di; }
    else if (mode == 4) { si = h - 1 - dj; sj = di; }
    else { si = dj; sj = di; }
    uint8_t v = 0;
    if (di < dh && dj < dw) v = s[(size_t)si * w + sj];
    tile[r][tx] = v;  // tile[dst j][dst i]
  }
  __syncthreads();
#pragma unroll
  for (int r = ty; r < TILE; r += 8) {
    const int di = di0 + r, dj = dj0 + tx;
    if (di < dh && dj < dw) d[(size_t)di * dw + dj] = tile[tx][r];
  }
}

// ---- cv2 border rule (borderInterpolate) ------------------------------------------
__device__ __forceinline__ int border_index(int p, int len, int border) {
  if ((unsigned)p < (unsigned)len) return p;
  if (border == B2SEG_BORDER_CONSTANT) return -1;
  if (len == 1) return 0;
  const int delta = border == B2SEG_BORDER_REFLECT101 ? 1 : 0;
  do {
    if (p < 0) p = -p - 1 + delta;
    else p = len - 1 - (p - len) - delta;
  } while ((unsigned)p >= (unsigned)len);
  return p;
}

struct WarpArgs {
  double m[6];          // affine: dst->src matrix (already inverted, as cv2 does)
  const float* map_x;   // remap: absolute float maps (or null for affine)
  const float* map_y;
  int h, w, interp, border, cval;
};

// Fixed-point source coordinate of destination pixel (x,y): integer part and, for
// bilinear, the 5-bit fractions.  Mirrors cv2's WarpAffineInvoker / remap().
__device__ __forceinline__ void src_coord(const WarpArgs& a, int x, int y, int& sx, int& sy,
                                          int& fx, int& fy) {
  if (a.map_x == nullptr) {
    const int round_delta = a.interp == B2SEG_INTERP_NEAREST ? 512 : 16;
    const int ad = __double2int_rn(__dmul_rn(__dmul_rn(a.m[0], (double)x), 1024.0));
    const int bd = __double2int_rn(__dmul_rn(__dmul_rn(a.m[3], (double)x), 1024.0));
    const int X0 =
        __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(a.m[1], (double)y), a.m[2]), 1024.0)) +
        round_delta;
    const int Y0 =
        __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(a.m[4], (double)y), a.m[5]), 1024.0)) +
        round_delta;
    if (a.interp == B2SEG_INTERP_NEAREST) {
      sx = (X0 + ad) >> 10;
      sy = (Y0 + bd) >> 10;
      fx = fy = 0;
    } else {
      const int X = (X0 + ad) >> 5, Y = (Y0 + bd) >> 5;
      sx = X >> 5;
      sy = Y >> 5;
      fx = X & 31;
      fy = Y & 31;
    }
  } else {
    const size_t o = (size_t)y * a.w + x;
    const float mx = a.map_x[o], my = a.map_y[o];
    if (a.interp == B2SEG_INTERP_NEAREST) {
      sx = __float2int_rn(mx);
      sy = __float2int_rn(my);
      fx = fy = 0;
    } else {
      const int X = __float2int_rn(mx * 32.0f), Y = __float2int_rn(my * 32.0f);
      sx = X >> 5;
      sy = Y >> 5;
      fx = X & 31;
      fy = Y & 31;
    }
  }
  // saturate_cast<short>
  sx = max(-32768, min(32767, sx));
  sy = max(-32768, min(32767, sy));
}

__device__ __forceinline__ int fetch_global(const uint8_t* s, int h, int w, int x, int y,
                                            int border, int cval) {
  const int xx = border_index(x, w, border), yy = border_index(y, h, border);
  if (xx < 0 || yy < 0) return cval;
  return s[(size_t)yy * w + xx];
}

__global__ void __launch_bounds__(256)
    warp_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, WarpArgs a) {
  __shared__ uint8_t foot[FOOT][FOOT + 4];
  __shared__ int s_box[4];
  const int h = a.h, w = a.w;
  const uint8_t* s = src + (size_t)blockIdx.z * h * w;
  uint8_t* d = dst + (size_t)blockIdx.z * h * w;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int x = blockIdx.x * TILE + tx;
  const bool lin = a.interp == B2SEG_INTERP_LINEAR;

  int sx[4], sy[4], fx[4], fy[4];
  int mnx = INT_MAX, mny = INT_MAX, mxx = INT_MIN, mxy = INT_MIN;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int y = blockIdx.y * TILE + ty + 8 * k;
    if (x < w && y < h) {
      src_coord(a, x, y, sx[k], sy[k], fx[k], fy[k]);
      if (a.border == B2SEG_BORDER_CONSTANT) {
        // anything at or beyond one pixel outside reads only the constant
        sx[k] = max(-2, min(w, sx[k]));
        sy[k] = max(-2, min(h, sy[k]));
      }
      mnx = min(mnx, sx[k]); mxx = max(mxx, sx[k]);
      mny = min(mny, sy[k]); mxy = max(mxy, sy[k]);
    }
  }
  // block-wide footprint
  if (threadIdx.x == 0) { s_box[0] = INT_MAX; s_box[1] = INT_MAX; s_box[2] = INT_MIN; s_box[3] = INT_MIN; }
  __syncthreads();
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mnx = min(mnx, __shfl_xor_sync(0xffffffffu, mnx, o));
    mny = min(mny, __shfl_xor_sync(0xffffffffu, mny, o));
    mxx = max(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
    mxy = max(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
  }
  if (tx == 0) {
    atomicMin(&s_box[0], mnx); atomicMin(&s_box[1], mny);
    atomicMax(&s_box[2], mxx); atomicMax(&s_box[3], mxy);
  }
  __syncthreads();
  const int bx0 = s_box[0], by0 = s_box[1];
  const int bw = s_box[2] - bx0 + (lin ? 2 : 1), bh = s_box[3] - by0 + (lin ? 2 : 1);
  const bool staged = s_box[0] != INT_MAX && bw <= FOOT && bh <= FOOT && bw > 0 && bh > 0;

  if (staged) {
    // coalesced footprint load with the border rule applied once per source pixel
    for (int r = ty; r < bh; r += 8) {
      const int yy = border_index(by0 + r, h, a.border);
      for (int c = tx; c < bw; c += 32) {
        const int xx = border_index(bx0 + c, w, a.border);
        foot[r][c] = (xx < 0 || yy < 0) ? (uint8_t)a.cval : s[(size_t)yy * w + xx];
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int y = blockIdx.y * TILE + ty + 8 * k;
    if (x >= w || y >= h) continue;
    int v00, v01 = 0, v10 = 0, v11 = 0;
    if (staged) {
      const int lx = sx[k] - bx0, ly = sy[k] - by0;
      v00 = foot[ly][lx];
      if (lin) { v01 = foot[ly][lx + 1]; v10 = foot[ly + 1][lx]; v11 = foot[ly + 1][lx + 1]; }
    } else {
      v00 = fetch_global(s, h, w, sx[k], sy[k], a.border, a.cval);
      if (lin) {
        v01 = fetch_global(s, h, w, sx[k] + 1, sy[k], a.border, a.cval);
        v10 = fetch_global(s, h, w, sx[k], sy[k] + 1, a.border, a.cval);
        v11 = fetch_global(s, h, w, sx[k] + 1, sy[k] + 1, a.border, a.cval);
      }
    }
    int out = v00;
    if (lin) {
      // cv2 BilinearTab_i: (32-fy)(32-fx)*32 etc., sum 32768; FixedPtCast<15>
      const int w00 = (32 - fy[k]) * (32 - fx[k]) * 32, w01 = (32 - fy[k]) * fx[k] * 32;
      const int w10 = fy[k] * (32 - fx[k]) * 32, w11 = fy[k] * fx[k] * 32;
      out = (v00 * w00 + v01 * w01 + v10 * w10 + v11 * w11 + (1 << 14)) >> 15;
      out = max(0, min(255, out));
    }
    d[(size_t)y * w + x] = (uint8_t)out;
  }
}

int launch_warp(const uint8_t* src, uint8_t* dst, int planes, const WarpArgs& a, void* stream) {
  dim3 grid((a.w + TILE - 1) / TILE, (a.h + TILE - 1) / TILE, planes);
  warp_u8_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst, a);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

}  // namespace

extern "C" int b2seg_aug_permute_u8(const uint8_t* src, uint8_t* dst, int planes, int h, int w,
                                    int mode, int k, void* stream) {
  B2_REQUIRE(src && dst, "aug_permute: null pointer");
  B2_REQUIRE(planes > 0 && h > 0 && w > 0, "aug_permute: bad shape %d x %d x %d", planes, h, w);
  B2_REQUIRE(planes <= 65535, "aug_permute: too many planes");
  int km;
  switch (mode) {
    case B2SEG_AUG_HFLIP: km = 0; break;
    case B2SEG_AUG_VFLIP: km = 1; break;
    case B2SEG_AUG_TRANSPOSE: km = 5; break;
    case B2SEG_AUG_ROT90: {
      const int kk = ((k % 4) + 4) % 4;
      if (kk == 0) {
        B2_CHECK_CUDA(cudaMemcpyAsync(dst, src, (size_t)planes * h * w, cudaMemcpyDeviceToDevice,
                                      (cudaStream_t)stream));
        return B2SEG_OK;
      }
      km = kk == 1 ? 3 : (kk == 2 ? 2 : 4);
      break;
    }
    default:
      B2_REQUIRE(false, "aug_permute: unknown mode %d", mode);
  }
  const bool swap = km >= 3;
  const int dh = swap ? w : h, dw = swap ? h : w;
  dim3 grid((dw + TILE - 1) / TILE, (dh + TILE - 1) / TILE, planes);
  permute_u8_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst, h, w, km);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_aug_warp_affine_u8(const uint8_t* src, uint8_t* dst, int planes, int h, int w,
                                        const double* minv, int interp, int border,
                                        int border_value, void* stream) {
  B2_REQUIRE(src && dst && minv, "aug_warp_affine: null pointer");
  B2_REQUIRE(planes > 0 && planes <= 65535 && h > 0 && w > 0, "aug_warp_affine: bad shape");
  B2_REQUIRE(interp == B2SEG_INTERP_NEAREST || interp == B2SEG_INTERP_LINEAR,
             "aug_warp_affine: interp %d unsupported", interp);
  B2_REQUIRE(border >= 0 && border <= 2, "aug_warp_affine: border %d unsupported", border);
  B2_REQUIRE(src != dst, "aug_warp_affine: in-place not supported");
  WarpArgs a;
  for (int i = 0; i < 6; ++i) a.m[i] = minv[i];
  a.map_x = a.map_y = nullptr;
  a.h = h; a.w = w; a.interp = interp; a.border = border; a.cval = border_value & 255;
  return launch_warp(src, dst, planes, a, stream);
}

extern "C" int b2seg_aug_remap_u8(const uint8_t* src, uint8_t* dst, int planes, int h, int w,
                                  const float* map_x, const float* map_y, int interp, int border,
                                  int border_value, void* stream) {
  B2_REQUIRE(src && dst && map_x && map_y, "aug_remap: null pointer");
  B2_REQUIRE(planes > 0 && planes <= 65535 && h > 0 && w > 0, "aug_remap: bad shape");
  B2_REQUIRE(interp == B2SEG_INTERP_NEAREST || interp == B2SEG_INTERP_LINEAR,
             "aug_remap: interp %d unsupported", interp);
  B2_REQUIRE(border >= 0 && border <= 2, "aug_remap: border %d unsupported", border);
  B2_REQUIRE(src != dst, "aug_remap: in-place not supported");
  WarpArgs a;
  for (int i = 0; i < 6; ++i) a.m[i] = 0.0;
  a.map_x = map_x; a.map_y = map_y;
  a.h = h; a.w = w; a.interp = interp; a.border = border; a.cval = border_value & 255;
  return launch_warp(src, dst, planes, a, stream);
}
