// Encoder stem on the tensor cores: 7x7 stride-2 convolution of the 1-channel image
// (advanced_models.py:76,268) as an implicit GEMM  [128 pixels] x [K = 49 -> 64] x [64 channels].
//
// There is no channel dimension to feed TMA with, so the A operand is built by the CTA itself:
// a 21x37 fp32 input patch is staged in shared memory, every thread gathers the 49 taps of ITS
// output pixel, converts them to bf16 and writes the 128-byte K-major row at the position the
// 128-byte swizzle expects (16-byte chunk c of row r lives at chunk c ^ (r & 7)); the weights are
// converted once per CTA into the same layout.  After a generic->async proxy fence one elected
// thread issues four tcgen05.mma (128x64x16) into a 64-column TMEM accumulator; the epilogue reads
// it back with tcgen05.ld, stores bf16 NHWC rows (128 B per pixel) and folds the BatchNorm
// (sum, sum of squares) partials exactly like the FMA kernel it replaces (stem_pool.cu).
// 4 CTAs of 128 threads per SM hide each other's patch-load and epilogue latency.
#include "ew.cuh"

namespace {

constexpr int TC_TH = 8, TC_TW = 16;              // output-pixel brick = 128 GEMM rows
constexpr int TC_PH = 2 * TC_TH + 5;              // 21 input rows
constexpr int TC_PW = 2 * TC_TW + 5;              // 37 input cols
constexpr int TC_PWP = TC_PW + 1;
constexpr int TC_PATCH = TC_PH * TC_PW;           // 777 input values per tile
constexpr int TC_PER_THREAD = (TC_PATCH + 127) / 128;  // 7
constexpr int TC_A_BYTES = 128 * 128, TC_B_BYTES = 64 * 128;
constexpr int TC_SMEM = TC_A_BYTES + TC_B_BYTES + TC_PH * TC_PWP * 4 + 64 + 1024;

struct StemTile {
  int n, oh0, ow0;
};
__device__ __forceinline__ StemTile stem_tile(long long tile, int tilesW, int tilesH) {
  StemTile s;
  const int tw = (int)(tile % tilesW), th = (int)((tile / tilesW) % tilesH);
  s.n = (int)(tile / ((long long)tilesW * tilesH));
  s.oh0 = th * TC_TH;
  s.ow0 = tw * TC_TW;
  return s;
}
// this thread's share of a tile's input patch, zero outside the image (= the conv padding)
__device__ __forceinline__ void stem_fetch(float (&v)[TC_PER_THREAD], const float* __restrict__ x,
                                           const StemTile& s, int H, int W) {
  const int ih0 = 2 * s.oh0 - 3, iw0 = 2 * s.ow0 - 3;
#pragma unroll
  for (int q = 0; q < TC_PER_THREAD; ++q) {
    const int e = threadIdx.x + q * 128;
    const int r = e / TC_PW, c = e - r * TC_PW;
    const int ih = ih0 + r, iw = iw0 + c;
    v[q] = (e < TC_PATCH && ih >= 0 && ih < H && iw >= 0 && iw < W)
               ? __ldg(x + ((size_t)s.n * H + ih) * W + iw) : 0.0f;
  }
}
__device__ __forceinline__ void stem_stash(const float (&v)[TC_PER_THREAD], float* patch) {
#pragma unroll
  for (int q = 0; q < TC_PER_THREAD; ++q) {
    const int e = threadIdx.x + q * 128;
    if (e < TC_PATCH) {
      const int r = e / TC_PW, c = e - r * TC_PW;
      patch[r * TC_PWP + c] = v[q];
    }
  }
}
// 128-byte K-major row `r` of a swizzled tile: words[32] = 64 bf16 along K
__device__ __forceinline__ void store_swizzled_row(uint8_t* tile, int r, const uint32_t (&w)[32]) {
  uint8_t* row = tile + r * 128;
#pragma unroll
  for (int c = 0; c < 8; ++c)
    *reinterpret_cast<uint4*>(row + ((c ^ (r & 7)) << 4)) =
        make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
}

__global__ void __launch_bounds__(128, 4) stem_tc_fwd_kernel(const EwArgs a) {
  pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint8_t* As = smem;
  uint8_t* Bs = smem + TC_A_BYTES;
  float* patch = reinterpret_cast<float*>(Bs + TC_B_BYTES);
  uint64_t* mma_bar = reinterpret_cast<uint64_t*>(patch + TC_PH * TC_PWP);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mma_bar + 1);
  const float* __restrict__ x = (const float*)a.p[0];
  const float* __restrict__ w = (const float*)a.p[1];
  bf16* __restrict__ y = (bf16*)a.p[2];
  float* __restrict__ partial = (float*)a.p[3];
  const int N = (int)a.i[0], H = (int)a.i[1], W = (int)a.i[2];
  const int OH = H / 2, OW = W / 2;
  const int tilesH = (OH + TC_TH - 1) / TC_TH, tilesW = (OW + TC_TW - 1) / TC_TW;
  const long long tiles = (long long)N * tilesH * tilesW;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;

  if (t == 0) {
    mbar_init(mma_bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<64>(tmem_slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_wait();  // the weights were written by the previous step's optimizer kernel

  {  // B operand: row = output channel, K = tap (49 -> 64, zero padded); two threads per row
    const int n = t >> 1, h = t & 1;
    uint8_t* row = Bs + n * 128;
#pragma unroll
    for (int c = 4 * h; c < 4 * h + 4; ++c) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = 8 * c + j;
        f[j] = k < 49 ? w[n * 49 + k] : 0.0f;
      }
      *reinterpret_cast<uint4*>(row + ((c ^ (n & 7)) << 4)) =
          make_uint4(f2_to_bf2(f[0], f[1]), f2_to_bf2(f[2], f[3]), f2_to_bf2(f[4], f[5]), f2_to_bf2(f[6], f[7]));
    }
  }
  constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);
  const uint64_t adesc = umma_smem_desc(smem_u32(As), 16, 1024);
  const uint64_t bdesc = umma_smem_desc(smem_u32(Bs), 16, 1024);
  const int ty = t / TC_TW, tx = t % TC_TW;
  float ssum[2] = {0.0f, 0.0f}, ssq[2] = {0.0f, 0.0f};  // channels lane, lane + 32 over this warp's pixels
  uint32_t phase = 0;
  float pre[TC_PER_THREAD];
  long long tile = blockIdx.x;
  if (tile < tiles) stem_fetch(pre, x, stem_tile(tile, tilesW, tilesH), H, W);
  for (; tile < tiles; tile += gridDim.x) {
    const StemTile s = stem_tile(tile, tilesW, tilesH);
    stem_stash(pre, patch);
    __syncthreads();
    if (tile + gridDim.x < tiles)  // next tile's patch travels while this one is built and multiplied
      stem_fetch(pre, x, stem_tile(tile + gridDim.x, tilesW, tilesH), H, W);
    {
      uint32_t words[32];
      const float* p0 = patch + (2 * ty) * TC_PWP + 2 * tx;
#pragma unroll
      for (int i = 0; i < 24; ++i) {
        const int k0 = 2 * i, k1 = 2 * i + 1;
        words[i] = f2_to_bf2(p0[(k0 / 7) * TC_PWP + (k0 % 7)], p0[(k1 / 7) * TC_PWP + (k1 % 7)]);
      }
      words[24] = f2_to_bf2(p0[6 * TC_PWP + 6], 0.0f);
#pragma unroll
      for (int i = 25; i < 32; ++i) words[i] = 0u;
      store_swizzled_row(As, t, words);
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (warp == 0) {
      tcgen05_fence_after();
      if (elect_one_sync()) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16_ss(tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, k != 0 ? 1u : 0u);
        tcgen05_commit(mma_bar);
      }
      __syncwarp();
    }
    mbar_wait(mma_bar, phase);
    phase ^= 1u;
    tcgen05_fence_after();
    const int oh = s.oh0 + ty, ow = s.ow0 + tx;
    const bool ok = oh < OH && ow < OW;
    bf16* dst = y + (((size_t)s.n * OH + oh) * OW + ow) * 64;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      float v[32];
      tmem_ld_32x32b_x32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(c * 32), v);
      if (ok) {
#pragma unroll
        for (int g = 0; g < 4; ++g)
          *reinterpret_cast<uint4*>(dst + c * 32 + g * 8) =
              make_uint4(f2_to_bf2(v[8 * g], v[8 * g + 1]), f2_to_bf2(v[8 * g + 2], v[8 * g + 3]),
                         f2_to_bf2(v[8 * g + 4], v[8 * g + 5]), f2_to_bf2(v[8 * g + 6], v[8 * g + 7]));
      }
      if (partial) {
        float s2[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          v[j] = ok ? v[j] : 0.0f;
          s2[j] = v[j] * v[j];
        }
        ssum[c] += warp_transpose_sum32(v, lane);
        ssq[c] += warp_transpose_sum32(s2, lane);
      }
    }
    tcgen05_fence_before();  // the next tile's MMA overwrites the accumulator after the barrier above
  }
  if (partial) {  // row = block*4 + warp: [2][64], same layout as the FMA kernel's rows
    float* row = partial + ((size_t)blockIdx.x * 4 + warp) * 128;
    row[lane] = ssum[0];
    row[32 + lane] = ssum[1];
    row[64 + lane] = ssq[0];
    row[96 + lane] = ssq[1];
  }
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<64>(tmem);
  }
}

// ---- weight gradient ------------------------------------------------------------------------
// dW[c][tap] = sum_pix dy[pix][c] * x[2*oh-3+kh, 2*ow-3+kw]:  D[tap][c] += A^T B with the pixels as
// K.  Both operands are MN-major tiles of 128 pixel rows x 128 B: A = the same im2col rows as the
// forward kernel (taps along M), B = the dy rows of the brick (channels along N), copied from
// global memory into the swizzled layout by the thread that owns the pixel.  The MMA is 128 rows
// tall; the second 64-row block of A is pointed at the B tile (LBO = 16 KB) and its products are
// never read.  The accumulator stays in TMEM across all bricks of the CTA; one fp32 partial row
// [64][49] per CTA goes to the workspace COLSUM_FINALIZE reduces (same contract as stem_pool.cu).
__global__ void __launch_bounds__(128, 4) stem_tc_wgrad_kernel(const EwArgs a) {
  pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint8_t* As = smem;
  uint8_t* Bs = smem + TC_A_BYTES;  // 128 pixel rows x 64 channels
  float* patch = reinterpret_cast<float*>(Bs + TC_A_BYTES);
  uint64_t* mma_bar = reinterpret_cast<uint64_t*>(patch + TC_PH * TC_PWP);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mma_bar + 1);
  const float* __restrict__ x = (const float*)a.p[0];
  const bf16* __restrict__ dy = (const bf16*)a.p[1];
  float* __restrict__ ws = (float*)a.p[2];
  const int N = (int)a.i[0], H = (int)a.i[1], W = (int)a.i[2];
  const int OH = H / 2, OW = W / 2;
  const int tilesH = (OH + TC_TH - 1) / TC_TH, tilesW = (OW + TC_TW - 1) / TC_TW;
  const long long tiles = (long long)N * tilesH * tilesW;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;

  if (t == 0) {
    mbar_init(mma_bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<64>(tmem_slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_wait();

  constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 1, 1);
  const uint64_t adesc = umma_smem_desc(smem_u32(As), TC_A_BYTES, 1024);
  const uint64_t bdesc = umma_smem_desc(smem_u32(Bs), TC_A_BYTES, 1024);
  const int ty = t / TC_TW, tx = t % TC_TW;
  uint32_t phase = 0;
  bool first = true;
  float pre[TC_PER_THREAD];
  long long tile = blockIdx.x;
  if (tile < tiles) stem_fetch(pre, x, stem_tile(tile, tilesW, tilesH), H, W);
  for (; tile < tiles; tile += gridDim.x) {
    const StemTile s = stem_tile(tile, tilesW, tilesH);
    stem_stash(pre, patch);
    __syncthreads();
    if (tile + gridDim.x < tiles)
      stem_fetch(pre, x, stem_tile(tile + gridDim.x, tilesW, tilesH), H, W);
    const int oh = s.oh0 + ty, ow = s.ow0 + tx;
    const bool ok = oh < OH && ow < OW;
    uint4 dyr[8];
    {
      const uint4* src = reinterpret_cast<const uint4*>(dy + (((size_t)s.n * OH + oh) * OW + ow) * 64);
#pragma unroll
      for (int c = 0; c < 8; ++c) dyr[c] = ok ? __ldg(src + c) : make_uint4(0u, 0u, 0u, 0u);
    }
    {
      uint32_t words[32];
      const float* p0 = patch + (2 * ty) * TC_PWP + 2 * tx;
#pragma unroll
      for (int i = 0; i < 24; ++i) {
        const int k0 = 2 * i, k1 = 2 * i + 1;
        words[i] = f2_to_bf2(p0[(k0 / 7) * TC_PWP + (k0 % 7)], p0[(k1 / 7) * TC_PWP + (k1 % 7)]);
      }
      words[24] = f2_to_bf2(p0[6 * TC_PWP + 6], 0.0f);
#pragma unroll
      for (int i = 25; i < 32; ++i) words[i] = 0u;
      store_swizzled_row(As, t, words);
    }
    {
      uint8_t* row = Bs + t * 128;
#pragma unroll
      for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(row + ((c ^ (t & 7)) << 4)) = dyr[c];
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (warp == 0) {
      tcgen05_fence_after();
      if (elect_one_sync()) {
#pragma unroll
        for (int k = 0; k < 8; ++k)  // 16 pixel rows (2 KB) per MMA
          umma_bf16_ss(tmem, adesc + (uint64_t)(k * 128), bdesc + (uint64_t)(k * 128), idesc,
                       (first && k == 0) ? 0u : 1u);
        tcgen05_commit(mma_bar);
      }
      __syncwarp();
    }
    first = false;
    mbar_wait(mma_bar, phase);  // operands may be overwritten once the MMAs have read them
    phase ^= 1u;
  }
  tcgen05_fence_after();
  float* row = ws + (size_t)blockIdx.x * 3136;
  const int tap = warp * 32 + lane;  // TMEM lane = accumulator row = tap
  if (warp < 2) {
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      float v[32];
      if (!first) {
        tmem_ld_32x32b_x32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(c * 32), v);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.0f;
      }
      if (tap < 49) {
#pragma unroll
        for (int j = 0; j < 32; ++j) row[(c * 32 + j) * 49 + tap] = v[j];
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<64>(tmem);
  }
}

}  // namespace

int launch_stem_tc_fwd(const EwArgs& a, dim3 grid, dim3 block, size_t smem, cudaStream_t st) {
  (void)block;
  (void)smem;
  B2_CHECK_CUDA(b2_launch(stem_tc_fwd_kernel, grid, dim3(128), (size_t)TC_SMEM, st, a));
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

int launch_stem_tc_wgrad(const EwArgs& a, dim3 grid, dim3 block, size_t smem, cudaStream_t st) {
  (void)block;
  (void)smem;
  B2_CHECK_CUDA(b2_launch(stem_tc_wgrad_kernel, grid, dim3(128),
                          (size_t)(2 * TC_A_BYTES + TC_PH * TC_PWP * 4 + 64 + 1024), st, a));
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}
