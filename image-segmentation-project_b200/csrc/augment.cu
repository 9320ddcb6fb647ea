// Geometric augmentation as shared-memory-tiled gathers on uint8 planes.
//
// The reference applies albumentations transforms (dataset.py:148-154,184;
// forward_selection_integration.py:73-105) that bottom out in cv2.flip /
// np.rot90 / transpose / cv2.warpAffine / cv2.remap.  These kernels restate the
// cv2 arithmetic exactly: index permutations are bit-exact; warps use cv2's
// fixed-point scheme (coordinates in 1/32 px, 15-bit bilinear weights, constant
// / reflect / reflect101 borders).
//
// Warps: one CTA produces a 32x32 destination tile of several planes.  It computes the
// source footprint of the tile, stages it in shared memory (border rule already applied, next
// planes prefetched in registers, bank-conflict-aware pitch) and gathers from shared memory with
// one dp4a per pixel row; each thread emits 4 adjacent pixels as one 32-bit store.  Footprints
// that do not fit fall back to direct global gathers.
// Permutations: 16-byte chunk moves (flips) and register/shared-memory 64x64 byte-tile
// transposes (rot90 / transpose); odd sizes take the scalar tile kernel.
#include "common.cuh"
#include <type_traits>

namespace {

constexpr int TILE = 32;

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

// ---- vectorised permutations (the common case: 16-byte aligned planes, sides % 64 == 0) -----
__device__ __forceinline__ uint32_t brev4(uint32_t v) { return __byte_perm(v, 0, 0x0123); }
__device__ __forceinline__ uint4 rev16(uint4 v) {
  return make_uint4(brev4(v.w), brev4(v.z), brev4(v.y), brev4(v.x));
}

// Flips move whole 16-byte chunks: vflip copies chunk (i, q) from row h-1-i, hflip / rot180 read
// the mirrored chunk and reverse its bytes in registers.  Pure streaming: 16 B in, 16 B out.
template <typename IDX>
__global__ void __launch_bounds__(256)
    flip16_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, int h, int wq, long long total_,
                  int mode /*0 h, 1 v, 2 rot180*/) {
  const IDX total = (IDX)total_, stride = (IDX)gridDim.x * 256;
  const IDX uwq = (IDX)wq, uh = (IDX)h;
  for (IDX idx = (IDX)blockIdx.x * 256 + threadIdx.x; idx < total; idx += stride) {
    const IDX row = idx / uwq, q = idx - row * uwq;
    const IDX plane = row / uh, i = row - plane * uh;
    const IDX si = mode == 0 ? i : uh - 1 - i;
    const IDX sq = mode == 1 ? q : uwq - 1 - q;
    uint4 v = __ldcs(src + ((size_t)plane * uh + si) * uwq + sq);
    if (mode != 1) v = rev16(v);
    __stcs(dst + idx, v);
  }
}

// Transposing family on 64x64 byte tiles.  Each thread loads a 4x4 byte block (four coalesced
// 32-bit loads), transposes it in registers (8 PRMT) and parks it in shared memory as words of
// the transposed tile; the tile is then written as 16-byte chunks of destination rows.
//   transpose : dst(i,j) = src(j,i)
//   rot90 k=1 : dst(i,j) = src(j, w-1-i)   -> rows of the transposed tile land in reverse order
//   rot90 k=3 : dst(i,j) = src(h-1-j, i)   -> each row of the transposed tile is byte-reversed
__global__ void __launch_bounds__(256)
    transpose64_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int h, int w,
                       int mode /*3 rot90 k=1, 4 rot90 k=3, 5 transpose*/) {
  __shared__ uint32_t T[64][17];
  const int t = threadIdx.x;
  const size_t plane_off = (size_t)blockIdx.z * h * w;
  const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 64;  // source tile origin
  {
    const int l = t & 15, g = t >> 4;
    const uint8_t* s = src + plane_off + (size_t)(r0 + 4 * g) * w + c0 + 4 * l;
    const uint32_t a0 = __ldcs(reinterpret_cast<const uint32_t*>(s));
    const uint32_t a1 = __ldcs(reinterpret_cast<const uint32_t*>(s + w));
    const uint32_t a2 = __ldcs(reinterpret_cast<const uint32_t*>(s + 2 * (size_t)w));
    const uint32_t a3 = __ldcs(reinterpret_cast<const uint32_t*>(s + 3 * (size_t)w));
    const uint32_t t0 = __byte_perm(a0, a1, 0x5140), t1 = __byte_perm(a2, a3, 0x5140);
    const uint32_t t2 = __byte_perm(a0, a1, 0x7362), t3 = __byte_perm(a2, a3, 0x7362);
    T[4 * l + 0][g] = __byte_perm(t0, t1, 0x5410);  // source column 4l+0, rows 4g..4g+3
    T[4 * l + 1][g] = __byte_perm(t0, t1, 0x7632);
    T[4 * l + 2][g] = __byte_perm(t2, t3, 0x5410);
    T[4 * l + 3][g] = __byte_perm(t2, t3, 0x7632);
  }
  __syncthreads();
  const int c = t >> 2, q = t & 3;  // transposed-tile row (= source column), 16-byte quarter
  uint4 v = make_uint4(T[c][4 * q], T[c][4 * q + 1], T[c][4 * q + 2], T[c][4 * q + 3]);
  size_t drow, dcol;  // destination is [w][h]
  if (mode == 5) { drow = c0 + c; dcol = r0 + 16 * q; }
  else if (mode == 3) { drow = w - 1 - (c0 + c); dcol = r0 + 16 * q; }
  else { drow = c0 + c; dcol = h - r0 - 16 * q - 16; v = rev16(v); }
  __stcs(reinterpret_cast<uint4*>(dst + plane_off + drow * h + dcol), v);
}

// Same family on 64 (rows) x 128 (columns) byte tiles: a warp reads 128 contiguous bytes of a source row
// (full lines instead of 64-byte halves) and every thread has eight 32-bit loads in flight instead of four.
__global__ void __launch_bounds__(256)
    transpose64x128_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int h, int w, int mode) {
  __shared__ uint32_t T[128][17];
  const int t = threadIdx.x;
  const size_t plane_off = (size_t)blockIdx.z * h * w;
  const int r0 = blockIdx.y * 64, c0 = blockIdx.x * 128;  // source tile origin
  {
    const int l = t & 31, g = t >> 5;
    uint32_t a[2][4];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const uint8_t* s = src + plane_off + (size_t)(r0 + 4 * (g + 8 * u)) * w + c0 + 4 * l;
#pragma unroll
      for (int k = 0; k < 4; ++k) a[u][k] = __ldcs(reinterpret_cast<const uint32_t*>(s + k * (size_t)w));
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const uint32_t t0 = __byte_perm(a[u][0], a[u][1], 0x5140), t1 = __byte_perm(a[u][2], a[u][3], 0x5140);
      const uint32_t t2 = __byte_perm(a[u][0], a[u][1], 0x7362), t3 = __byte_perm(a[u][2], a[u][3], 0x7362);
      T[4 * l + 0][g + 8 * u] = __byte_perm(t0, t1, 0x5410);  // source column 4l+0, rows 4g'..4g'+3
      T[4 * l + 1][g + 8 * u] = __byte_perm(t0, t1, 0x7632);
      T[4 * l + 2][g + 8 * u] = __byte_perm(t2, t3, 0x5410);
      T[4 * l + 3][g + 8 * u] = __byte_perm(t2, t3, 0x7632);
    }
  }
  __syncthreads();
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int e = t + 256 * u;
    const int c = e >> 2, q = e & 3;  // transposed-tile row (= source column), 16-byte quarter
    uint4 v = make_uint4(T[c][4 * q], T[c][4 * q + 1], T[c][4 * q + 2], T[c][4 * q + 3]);
    size_t drow, dcol;  // destination is [w][h]
    if (mode == 5) { drow = c0 + c; dcol = r0 + 16 * q; }
    else if (mode == 3) { drow = w - 1 - (c0 + c); dcol = r0 + 16 * q; }
    else { drow = c0 + c; dcol = h - r0 - 16 * q - 16; v = rev16(v); }
    __stcs(reinterpret_cast<uint4*>(dst + plane_off + drow * h + dcol), v);
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

// Affine coordinates of pixel (x,y) from the per-tile integer tables (exactly the values
// src_coord computes: ad/bd depend only on x, X0/Y0 only on y).
__device__ __forceinline__ void affine_from_tables(int interp, int adx, int bdx, int X0y, int Y0y,
                                                   int& sx, int& sy, int& fx, int& fy) {
  if (interp == B2SEG_INTERP_NEAREST) {
    sx = (X0y + adx) >> 10;
    sy = (Y0y + bdx) >> 10;
    fx = fy = 0;
  } else {
    const int X = (X0y + adx) >> 5, Y = (Y0y + bdx) >> 5;
    sx = X >> 5;
    sy = Y >> 5;
    fx = X & 31;
    fy = Y & 31;
  }
  sx = max(-32768, min(32767, sx));
  sy = max(-32768, min(32767, sy));
}

// ---- warp kernel ----------------------------------------------------------------------------------
// One CTA = one TW x TH = 32 x 32 (or 64 x 16) destination tile of `pz` consecutive planes (all planes of a call
// share the transform, so source coordinates, shared-memory addresses and blend weights are computed once and kept
// in registers).  Thread = 4 horizontally adjacent pixels -> one 32-bit store.  Affine coordinates come from four
// integer tables (the double-precision products are evaluated once per tile row / column, not per pixel); the
// source footprint of an affine tile is the bounding box of its four corners.
//
// Table mode (`items` != NULL): blockIdx.z = item * ppi + plane; every item brings its own matrix (or its own pair
// of maps, `map_stride` floats apart) and every plane of an item its own interpolation and constant -- image
// (bilinear) and mask (nearest) of ALL augmented copies of a batch are one launch.
//
// Built around the instruction count and the shared-memory wavefronts (the round-1 kernel: byte gathers, 59
// instructions per pixel, issue bound at 9-11 % of the HBM roofline; profiles/r02_warp_kernel_notes.txt):
//  * every thread owns up to three fixed 32-bit words of the footprint (word index -> row by ONE division per
//    CTA, offsets kept in registers) and loads them unpredicated into registers FOUR planes ahead of the plane
//    that is being gathered (cp.async instead of registers was measured and lost: a 4-byte LDGSTS costs ~4.5
//    shared-memory wavefronts per warp, an STS one);
//  * two staging buffers: plane k + 1 is stored while plane k is gathered, one barrier per plane;
//  * the row pitch of the staged footprint is chosen per CTA: every warp counts, for two of the pitches that
//    fit, the bank conflicts of its own gather addresses (match.any on bank and on address) and the CTA keeps the
//    best.  With a fixed pitch a rotated tile spreads a warp's 32 words pseudo-randomly over the banks (3.5
//    wavefronts per LDS measured, shared memory 75 % busy; 1.8 with the chosen pitch);
//  * bilinear planes keep a second copy of the footprint shifted by two bytes (built with one shuffle per word).
//    Whatever the parity of the source column, one aligned 32-bit shared-memory word then holds both horizontal
//    neighbours: copy (lx >> 1) & 1, word lx >> 2, bytes (lx & 1) and (lx & 1) + 1.  The horizontal blend
//    v0 * (32 - fx) + v1 * fx is ONE dp4a against a per-pixel weight word that is zero in the other two byte
//    lanes, so a pixel costs 2 LDS.32 + 2 IDP4A + 2 IMAD + shift + pack instead of 4 LDS.U8 + 9 integer ops;
//  * the border rule is folded into the description of a word once per CTA (see `esel` below).
// Footprints of more than 744 words (strong zoom-out) are gathered from global memory.
constexpr int NSLOT = 3;      // footprint words per thread
constexpr int SEGW = 31;      // new words per warp and slot: lane 31 re-loads its neighbour's word for the shifted copy
constexpr int FWORDS = 1280;  // words of one staged footprint (rows x chosen pitch); word FWORDS is a dump for idle lanes
constexpr int CWORDS = 1312;  // distance of the shifted copy in words (a multiple of 32: same banks)
constexpr int PDEPTH = 4;     // planes whose footprint words are in flight in registers (6 measured: no gain)
constexpr int FBUF = 2 * CWORDS;  // words of one staging buffer (footprint + shifted copy); there are two

// Footprint word with the border rule applied per byte (tiles that touch the image edge, or unaligned planes).
__device__ __noinline__ uint32_t foot_word_border(const uint8_t* __restrict__ s, int h, int w, int gy, int gx,
                                                  int border, int cval) {
  const int yy = border_index(gy, h, border);
  uint32_t word = 0;
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    const int xx = border_index(gx + b, w, border);
    const uint32_t v = (xx < 0 || yy < 0) ? (uint32_t)cval : (uint32_t)s[(size_t)yy * w + xx];
    word |= v << (8 * b);
  }
  return word;
}

template <int TW, int TH>
__global__ void __launch_bounds__(256, 4)
    warp_u8_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, WarpArgs a, int planes, int pz,
                   int vec_ok, const b2seg_warp_item* __restrict__ items, int ppi, long long map_stride) {
  static_assert(TW * TH == 1024 && TW % 4 == 0 && TW + TH <= 256, "256 threads x 4 pixels");
  constexpr int TPR = TW / 4;  // threads per destination row
  __shared__ __align__(16) uint32_t foot[2 * FBUF];  // two buffers of (footprint | footprint shifted by two bytes)
  __shared__ int s_pitch;
  if (items != nullptr) {
    const int item = blockIdx.z / ppi, pl = blockIdx.z - item * ppi;
    const b2seg_warp_item& it = items[item];
#pragma unroll
    for (int i = 0; i < 6; ++i) a.m[i] = it.minv[i];
    a.interp = it.interp[pl];
    a.cval = it.cval[pl] & 255;
    a.border = it.border;
    if (a.map_x != nullptr) {
      a.map_x += (size_t)item * map_stride;
      a.map_y += (size_t)item * map_stride;
    }
  }
  __shared__ int s_ad[TW], s_bd[TW], s_X0[TH], s_Y0[TH];
  __shared__ int s_box[4];
  const int h = a.h, w = a.w;
  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
  const int r = tid / TPR, xg = (tid % TPR) * 4;
  const int y = y0 + r, x = x0 + xg;
  const bool lin = a.interp == B2SEG_INTERP_LINEAR;
  const bool affine = a.map_x == nullptr;

  int sx[4], sy[4], fx[4], fy[4];
  int bx0, by0, bx1, by1;  // inclusive box of the integer source coordinates
  if (affine) {
    if (tid < TW) {
      const double xd = (double)(x0 + tid);
      s_ad[tid] = __double2int_rn(__dmul_rn(__dmul_rn(a.m[0], xd), 1024.0));
      s_bd[tid] = __double2int_rn(__dmul_rn(__dmul_rn(a.m[3], xd), 1024.0));
    } else if (tid < TW + TH) {
      const double yd = (double)(y0 + tid - TW);
      const int round_delta = a.interp == B2SEG_INTERP_NEAREST ? 512 : 16;
      s_X0[tid - TW] =
          __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(a.m[1], yd), a.m[2]), 1024.0)) + round_delta;
      s_Y0[tid - TW] =
          __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(a.m[4], yd), a.m[5]), 1024.0)) + round_delta;
    }
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 4; ++e)
      affine_from_tables(a.interp, s_ad[xg + e], s_bd[xg + e], s_X0[r], s_Y0[r], sx[e], sy[e], fx[e], fy[e]);
    // corners of the valid part of the tile bound every pixel's coordinate (monotone in x and y)
    const int cx1 = min(TW - 1, w - 1 - x0), cy1 = min(TH - 1, h - 1 - y0);
    bx0 = by0 = INT_MAX; bx1 = by1 = INT_MIN;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int cx = (c & 1) ? cx1 : 0, cy = (c & 2) ? cy1 : 0;
      int qx, qy, q0, q1;
      affine_from_tables(a.interp, s_ad[cx], s_bd[cx], s_X0[cy], s_Y0[cy], qx, qy, q0, q1);
      bx0 = min(bx0, qx); bx1 = max(bx1, qx);
      by0 = min(by0, qy); by1 = max(by1, qy);
    }
  } else {
    int mnx = INT_MAX, mny = INT_MAX, mxx = INT_MIN, mxy = INT_MIN;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (x + e < w && y < h) {
        src_coord(a, x + e, y, sx[e], sy[e], fx[e], fy[e]);
        mnx = min(mnx, sx[e]); mxx = max(mxx, sx[e]);
        mny = min(mny, sy[e]); mxy = max(mxy, sy[e]);
      } else {
        sx[e] = sy[e] = fx[e] = fy[e] = 0;
      }
    }
    if (tid == 0) { s_box[0] = INT_MAX; s_box[1] = INT_MAX; s_box[2] = INT_MIN; s_box[3] = INT_MIN; }
    __syncthreads();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mnx = min(mnx, __shfl_xor_sync(0xffffffffu, mnx, o));
      mny = min(mny, __shfl_xor_sync(0xffffffffu, mny, o));
      mxx = max(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
      mxy = max(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
    }
    if ((tid & 31) == 0) {
      atomicMin(&s_box[0], mnx); atomicMin(&s_box[1], mny);
      atomicMax(&s_box[2], mxx); atomicMax(&s_box[3], mxy);
    }
    __syncthreads();
    bx0 = s_box[0]; by0 = s_box[1]; bx1 = s_box[2]; by1 = s_box[3];
  }
  if (a.border == B2SEG_BORDER_CONSTANT) {
    // anything at or beyond one pixel outside the image reads only the constant
    bx0 = max(-2, min(w, bx0)); bx1 = max(-2, min(w, bx1));
    by0 = max(-2, min(h, by0)); by1 = max(-2, min(h, by1));
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      sx[e] = max(-2, min(w, sx[e]));
      sy[e] = max(-2, min(h, sy[e]));
    }
  }
  const int bxa = bx0 & ~3;  // footprint columns start on a 4-byte boundary of the source row
  const int bw = bx1 + (lin ? 1 : 0) - bxa + 1, bh = by1 + (lin ? 1 : 0) - by0 + 1;
  const int bwords = (bw + 3) >> 2;
  const int nwords = bh * bwords;
  const bool staged = bx0 != INT_MAX && bw > 0 && bh > 0 && nwords <= 8 * SEGW * NSLOT;
  const bool valid_row = y < h;

  const int p_begin = items ? (int)blockIdx.z : blockIdx.z * pz;
  const int p_end = items ? p_begin + 1 : min(planes, p_begin + pz);
  if (!staged) {  // footprint larger than the staging buffer (strong zoom-out): gather from global memory
    if (!valid_row) return;
    for (int plane = p_begin; plane < p_end; ++plane) {
      const uint8_t* s = src + (size_t)plane * h * w;
      uint8_t* d = dst + (size_t)plane * h * w;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        if (x + e >= w) continue;
        int out = fetch_global(s, h, w, sx[e], sy[e], a.border, a.cval);
        if (lin) {
          const int v01 = fetch_global(s, h, w, sx[e] + 1, sy[e], a.border, a.cval);
          const int v10 = fetch_global(s, h, w, sx[e], sy[e] + 1, a.border, a.cval);
          const int v11 = fetch_global(s, h, w, sx[e] + 1, sy[e] + 1, a.border, a.cval);
          const int top = out * (32 - fx[e]) + v01 * fx[e], bot = v10 * (32 - fx[e]) + v11 * fx[e];
          out = ((32 - fy[e]) * top + fy[e] * bot + 512) >> 10;
        }
        d[(size_t)y * w + x + e] = (uint8_t)out;
      }
    }
    return;
  }

  // Row pitch of the staged footprint (words): every warp tries two of the (up to 16) pitches that fit and counts,
  // over its own gather addresses of pixels 0 and 2, the lanes that share a bank without sharing the word;
  // the CTA keeps the pitch with the fewest (ties: the smallest).
  const int wi = tid >> 5, lane = tid & 31;
  if (tid == 0) s_pitch = INT_MAX;
  __syncthreads();
  {
    const int pmax = min(bwords + 15, FWORDS / bh);
    for (int P = bwords + wi; P <= pmax; P += 8) {
      int c = 0;
#pragma unroll
      for (int e = 0; e < 4; e += 2) {
        const bool v = valid_row && x + e < w;
        const unsigned vm = __ballot_sync(0xffffffffu, v);
        const int lx = sx[e] - bxa, ly = sy[e] - by0;
        const int aw = (lin ? ((lx >> 1) & 1) * CWORDS : 0) + ly * P + (lx >> 2);
        const unsigned same_bank = __match_any_sync(0xffffffffu, aw & 31);
        const unsigned same_word = __match_any_sync(0xffffffffu, aw);
        c += v ? __popc(same_bank & vm) - __popc(same_word & vm) : 0;
      }
      c = __reduce_add_sync(0xffffffffu, c);
      if (lane == 0) atomicMin(&s_pitch, (c << 12) | P);
    }
  }
  __syncthreads();
  const int P = s_pitch & 4095;

  // Per-pixel shared-memory byte offset and (bilinear) dp4a weight word; fixed for every plane.
  uint32_t goff[4], wsel[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    int lx = sx[e] - bxa, ly = sy[e] - by0;
    if (!valid_row || x + e >= w) lx = ly = 0;  // never stored; keep the address inside the array
    if (lin) {
      goff[e] = (uint32_t)((((lx >> 1) & 1) * CWORDS + ly * P + (lx >> 2)) * 4);
      wsel[e] = (uint32_t)((32 - fx[e]) | (fx[e] << 8)) << ((lx & 1) * 8);
    } else {
      goff[e] = (uint32_t)(ly * P * 4 + lx);
      wsel[e] = 0;
    }
  }
  const uint8_t* const f0 = reinterpret_cast<const uint8_t*>(foot);
  const int pitch_b = P * 4;
  auto gather_store = [&](const uint8_t* __restrict__ fb, uint8_t* __restrict__ d, auto lin_c) {  // d -> pixel (y, x)
    constexpr bool LIN = decltype(lin_c)::value;
    if (!valid_row) return;
    uint32_t packed = 0;
    if (LIN) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const uint32_t wt = *reinterpret_cast<const uint32_t*>(fb + goff[e]);
        const uint32_t wb = *reinterpret_cast<const uint32_t*>(fb + goff[e] + pitch_b);
        // cv2 BilinearTab_i weights are (32-fy)(32-fx)*32 etc. (sum 32768) with FixedPtCast<15>:
        //   (sum v*w + 2^14) >> 15  ==  ((32-fy)*top + fy*bot + 2^9) >> 10   exactly (the sum is 32 * (...)),
        // a convex combination of bytes, so no clamp is needed.
        const uint32_t top = __dp4a(wt, wsel[e], 0u), bot = __dp4a(wb, wsel[e], 0u);
        const uint32_t out = ((32u - (uint32_t)fy[e]) * top + (uint32_t)fy[e] * bot + 512u) >> 10;
        packed |= out << (8 * e);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) packed |= (uint32_t)fb[goff[e]] << (8 * e);
    }
    if (vec_ok && x + 3 < w) {
      __stcs(reinterpret_cast<uint32_t*>(d), packed);
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (x + e < w) d[e] = (uint8_t)(packed >> (8 * e));
    }
  };

  // Footprint words of this thread: in slot j, warp wi owns words (8 j + wi) * 31 .. + 30; lane 31 loads the
  // first word of the next run only to hand it to lane 30 for the shifted copy.  Idle lanes re-load the plane's
  // first word and store to the dump word, so the per-plane loop of an interior tile carries no predicates.
  uint32_t gsrc[NSLOT], sdst[NSLOT];
  uint32_t oob = 0;
#pragma unroll
  for (int j = 0; j < NSLOT; ++j) {
    const int i = (8 * j + wi) * SEGW + lane;
    const int fr = i / bwords, fc = i - fr * bwords;
    const int gy = by0 + fr, gx = bxa + 4 * fc;
    gsrc[j] = 0;
    if (i < nwords) {
      if (vec_ok && (unsigned)gy < (unsigned)h && gx >= 0 && gx + 3 < w) gsrc[j] = (uint32_t)(gy * w + gx);
      else oob |= 1u << j;
    }
    sdst[j] = (i < nwords && lane < SEGW) ? (uint32_t)(fr * P + fc) : (uint32_t)FWORDS;
  }
  const bool interior = __syncthreads_and(oob == 0) != 0;
  const int nslots = (nwords + 8 * SEGW - 1) / (8 * SEGW);  // 1..NSLOT, CTA-uniform

  // Tiles on the image edge.  The border rule is the same for every plane, so it is folded into the word's
  // description once: a row outside the image is another row (reflect) or the constant, and the four bytes of a
  // word that crosses the left / right edge come from at most two aligned words of that row -- two loads and
  // two byte permutes per plane, prefetched like any other word (calling the per-byte rule inside the plane
  // loop exposed a full memory latency per word and plane: edge CTAs set the kernel's duration).
  //   gsrc bit 0: the second aligned word is the next one (else the same);  esel: low half = PRMT selector over
  //   the two words, high half = PRMT selector that replaces bytes by the constant.
  // Words that do not fit this form (planes without 4-byte alignment, images narrower than the overhang) keep
  // their oob flag and take foot_word_border.
  uint32_t esel[NSLOT];
  if (!interior) {
#pragma unroll
    for (int j = 0; j < NSLOT; ++j) {
      esel[j] = 0x32103210u;
      if (!((oob >> j) & 1) || !vec_ok) continue;
      const int i = (8 * j + wi) * SEGW + lane;
      const int fr = i / bwords, fc = i - fr * bwords;
      const int gy = by0 + fr, gx = bxa + 4 * fc;
      const int yy = border_index(gy, h, a.border);
      int xx[4], lo = INT_MAX, hi = INT_MIN;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        xx[q] = yy < 0 ? -1 : border_index(gx + q, w, a.border);
        if (xx[q] >= 0) { lo = min(lo, xx[q]); hi = max(hi, xx[q]); }
      }
      if (hi < 0) {  // nothing but the constant
        esel[j] = 0x76543210u;
        oob &= ~(1u << j);
        continue;
      }
      const int base = lo & ~3;
      if (hi > base + 7) continue;  // stays on the per-byte path
      uint32_t s1 = 0, s2 = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        s1 |= (uint32_t)(xx[q] >= 0 ? xx[q] - base : 0) << (4 * q);
        s2 |= (uint32_t)(xx[q] >= 0 ? q : 4 + q) << (4 * q);
      }
      esel[j] = s1 | (s2 << 16);
      gsrc[j] = (uint32_t)(yy * w + base) | (hi >= base + 4 ? 1u : 0u);
      oob &= ~(1u << j);
    }
  }
  const uint32_t cval4 = (uint32_t)a.cval * 0x01010101u;

  // One barrier per plane: plane k + 1 is stored into the other buffer while plane k is gathered.
  auto run = [&](auto lin_c, auto ns_c, auto edge_c) {
    constexpr bool LIN = decltype(lin_c)::value;
    constexpr int NS = decltype(ns_c)::value;
    constexpr bool EDGE = decltype(edge_c)::value;
    const int np = p_end - p_begin;
    const size_t hw = (size_t)h * w;
    const uint8_t* sp = src + (size_t)p_begin * hw;                // plane whose words are loaded next
    uint8_t* dp = dst + (size_t)p_begin * hw + (size_t)y * w + x;  // this thread's four pixels
    constexpr int PD = PDEPTH;  // planes in flight; even (buffer parity = plane parity)
    uint32_t pre[PD][NS];
    auto load = [&](uint32_t (&r)[NS]) {
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        if (EDGE && ((oob >> j) & 1)) {
          const int i = (8 * j + wi) * SEGW + lane;
          const int fr = i / bwords, fc = i - fr * bwords;
          r[j] = foot_word_border(sp, h, w, by0 + fr, bxa + 4 * fc, a.border, a.cval);
        } else if (!EDGE) {
          r[j] = __ldg(reinterpret_cast<const uint32_t*>(sp + gsrc[j]));
        } else if (vec_ok) {
          const uint32_t* g = reinterpret_cast<const uint32_t*>(sp + (gsrc[j] & ~3u));
          const uint32_t two = __byte_perm(__ldg(g), __ldg(g + (gsrc[j] & 1u)), esel[j] & 0xffffu);
          r[j] = __byte_perm(two, cval4, esel[j] >> 16);
        } else {
          r[j] = 0;  // idle lane of an unaligned plane (every real word took the border path)
        }
      }
      sp += hw;
    };
    auto store = [&](const uint32_t (&r)[NS], int buf) {
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        foot[buf * FBUF + sdst[j]] = r[j];
        if (LIN) foot[buf * FBUF + sdst[j] + CWORDS] = __funnelshift_r(r[j], __shfl_down_sync(0xffffffffu, r[j], 1), 16);
      }
    };
#pragma unroll
    for (int u = 0; u < PD; ++u)
      if (u < np) load(pre[u]);
    store(pre[0], 0);
    if (PD < np) load(pre[0]);
    __syncthreads();
    for (int k0 = 0; k0 < np; k0 += PD) {
#pragma unroll
      for (int u = 0; u < PD; ++u) {
        const int k = k0 + u;
        if (k < np) {
          if (k + 1 < np) {
            store(pre[(u + 1) % PD], (u + 1) & 1);
            if (k + 1 + PD < np) load(pre[(u + 1) % PD]);
          }
          gather_store(f0 + (u & 1) * (FBUF * 4), dp, lin_c);
          dp += hw;
          __syncthreads();
        }
      }
    }
  };
  auto pick_ns = [&](auto lin_c, auto edge_c) {
    if (nslots == 1) run(lin_c, std::integral_constant<int, 1>{}, edge_c);
    else if (nslots == 2) run(lin_c, std::integral_constant<int, 2>{}, edge_c);
    else run(lin_c, std::integral_constant<int, 3>{}, edge_c);
  };
  if (interior) {
    if (lin) pick_ns(std::true_type{}, std::false_type{}); else pick_ns(std::false_type{}, std::false_type{});
  } else {
    if (lin) pick_ns(std::true_type{}, std::true_type{}); else pick_ns(std::false_type{}, std::true_type{});
  }
}

// B2SEG_WARP_TILE = 64x16 selects the wide tile (same speed at small angles, 194 vs 191 us at 128 x 1024^2; its
// footprint outgrows the staging buffer beyond ~30 degrees); default 32 x 32.
bool warp_wide_tile() {
  static const bool wide = [] {
    const char* e = getenv("B2SEG_WARP_TILE");
    return e && e[0] == '6';
  }();
  return wide;
}

int launch_warp_any(const uint8_t* src, uint8_t* dst, const WarpArgs& a, int planes, int pz, int gz, int vec_ok,
                    const b2seg_warp_item* items, int ppi, long long map_stride, void* stream) {
  const bool wide = warp_wide_tile();
  const int tw = wide ? 64 : 32, th = wide ? 16 : 32;
  const int tx = (a.w + tw - 1) / tw, ty = (a.h + th - 1) / th;
  B2_REQUIRE(ty <= 65535 && gz <= 65535, "aug_warp: image too tall for the launch grid (%d rows)", a.h);
  dim3 grid(tx, ty, gz);
  cudaStream_t st = (cudaStream_t)stream;
  if (wide) warp_u8_kernel<64, 16><<<grid, 256, 0, st>>>(src, dst, a, planes, pz, vec_ok, items, ppi, map_stride);
  else warp_u8_kernel<32, 32><<<grid, 256, 0, st>>>(src, dst, a, planes, pz, vec_ok, items, ppi, map_stride);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

int launch_warp(const uint8_t* src, uint8_t* dst, int planes, const WarpArgs& a, void* stream) {
  // ~16 CTAs per SM over the call (4 are resident: four waves) as long as a CTA still walks >= 8 planes with one
  // set of source coordinates and one pitch search (~590 instructions per warp, as much as 9 planes of the loop;
  // 8 / 32 CTAs per SM measured: 192 / 196 us against 191)
  const int tiles = ((a.w + 31) / 32) * ((a.h + 31) / 32);
  int gz = (148 * 16 + tiles - 1) / tiles;
  gz = gz < 1 ? 1 : (gz > planes ? planes : gz);
  int pz = (planes + gz - 1) / gz;
  if (pz < 8) {
    for (int cand = 8; cand > pz; cand >>= 1)
      if ((long long)tiles * ((planes + cand - 1) / cand) >= 148 * 8) { pz = cand; break; }
  }
  gz = (planes + pz - 1) / pz;
  const int vec_ok = (a.w % 4 == 0) && (reinterpret_cast<uintptr_t>(src) % 4 == 0) &&
                     (reinterpret_cast<uintptr_t>(dst) % 4 == 0);
  return launch_warp_any(src, dst, a, planes, pz, gz, vec_ok, nullptr, 1, 0, stream);
}

int launch_warp_batch(const uint8_t* src, uint8_t* dst, int n_items, int ppi, WarpArgs a,
                      const b2seg_warp_item* items, long long map_stride, void* stream) {
  const size_t plane_bytes = (size_t)a.h * a.w;
  const int vec_ok = (a.w % 4 == 0) && (plane_bytes % 4 == 0) && (reinterpret_cast<uintptr_t>(src) % 4 == 0) &&
                     (reinterpret_cast<uintptr_t>(dst) % 4 == 0);
  return launch_warp_any(src, dst, a, n_items * ppi, 1, n_items * ppi, vec_ok, items, ppi, map_stride, stream);
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
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
  if (aligned && !swap && w % 16 == 0) {
    const long long total = (long long)planes * h * (w / 16);
    const long long want = (total + 255) / 256;
    const int blocks = (int)(want < 148 * 16 ? want : 148 * 16);
    if (total < (1ll << 31))
      flip16_kernel<unsigned><<<blocks, 256, 0, (cudaStream_t)stream>>>(
          reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), h, w / 16, total, km);
    else
      flip16_kernel<unsigned long long><<<blocks, 256, 0, (cudaStream_t)stream>>>(
          reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), h, w / 16, total, km);
    B2_CHECK_CUDA(cudaGetLastError());
    return B2SEG_OK;
  }
  if (aligned && swap && w % 128 == 0 && h % 64 == 0) {
    dim3 grid128(w / 128, h / 64, planes);
    transpose64x128_kernel<<<grid128, 256, 0, (cudaStream_t)stream>>>(src, dst, h, w, km);
    B2_CHECK_CUDA(cudaGetLastError());
    return B2SEG_OK;
  }
  if (aligned && swap && w % 64 == 0 && h % 64 == 0) {
    dim3 grid64(w / 64, h / 64, planes);
    transpose64_kernel<<<grid64, 256, 0, (cudaStream_t)stream>>>(src, dst, h, w, km);
    B2_CHECK_CUDA(cudaGetLastError());
    return B2SEG_OK;
  }
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

// Batched warps: `items` is a DEVICE table of n_items entries; planes are [n_items][ppi][h][w].
extern "C" int b2seg_aug_warp_affine_batch_u8(const uint8_t* src, uint8_t* dst, int n_items, int ppi, int h,
                                              int w, const b2seg_warp_item* items, void* stream) {
  B2_REQUIRE(src && dst && items && src != dst, "aug_warp_affine_batch: bad pointers");
  B2_REQUIRE(n_items > 0 && ppi >= 1 && ppi <= 4 && (long long)n_items * ppi <= 65535 && h > 0 && w > 0,
             "aug_warp_affine_batch: bad shape (%d items x %d planes)", n_items, ppi);
  WarpArgs a;
  for (int i = 0; i < 6; ++i) a.m[i] = 0.0;
  a.map_x = a.map_y = nullptr;
  a.h = h; a.w = w; a.interp = B2SEG_INTERP_LINEAR; a.border = B2SEG_BORDER_CONSTANT; a.cval = 0;
  return launch_warp_batch(src, dst, n_items, ppi, a, items, 0, stream);
}

extern "C" int b2seg_aug_remap_batch_u8(const uint8_t* src, uint8_t* dst, int n_items, int ppi, int h, int w,
                                        const float* map_x, const float* map_y, int64_t map_stride,
                                        const b2seg_warp_item* items, void* stream) {
  B2_REQUIRE(src && dst && items && map_x && map_y && src != dst, "aug_remap_batch: bad pointers");
  B2_REQUIRE(n_items > 0 && ppi >= 1 && ppi <= 4 && (long long)n_items * ppi <= 65535 && h > 0 && w > 0,
             "aug_remap_batch: bad shape (%d items x %d planes)", n_items, ppi);
  B2_REQUIRE(map_stride == 0 || map_stride >= (int64_t)h * w, "aug_remap_batch: map stride %lld", (long long)map_stride);
  WarpArgs a;
  for (int i = 0; i < 6; ++i) a.m[i] = 0.0;
  a.map_x = map_x; a.map_y = map_y;
  a.h = h; a.w = w; a.interp = B2SEG_INTERP_LINEAR; a.border = B2SEG_BORDER_REFLECT; a.cval = 0;
  return launch_warp_batch(src, dst, n_items, ppi, a, items, map_stride, stream);
}
