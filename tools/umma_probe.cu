// Hardware probe (not product code): does tcgen05.mma read a 128B-swizzled shared-memory operand
// correctly when the descriptor's start address is NOT 1024-byte aligned and the 8-row group
// stride (SBO) is not a multiple of 1024 B?  That is what a 3x3 convolution needs to run all nine
// taps out of ONE halo brick in shared memory instead of nine separate TMA loads.
//
//   K-major  A (fprop/dgrad): row m of the MMA = smem row  off + (m/8)*sbo_rows + m%8
//   MN-major A (wgrad)      : K index k (pixel)  = smem row  off + (k/8)*sbo_rows + k%8
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 tools/umma_probe.cu \
//        -o tools/umma_probe -L image-segmentation-project_b200/b2seg -lb2seg -Xlinker -rpath=...
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#include "../image-segmentation-project_b200/csrc/common.cuh"

constexpr int ROWS = 512;  // rows of the A source matrix landed in smem (two 256-row TMA boxes)
constexpr int A_TILE = ROWS * 128;

struct ProbeParams {
  CUtensorMap tmA, tmA2, tmB;
  int off_rows, sbo_rows, bo_mode, mn_major;
  float* out;  // [128][64]
};

__device__ __forceinline__ uint64_t desc_with_bo(uint32_t saddr, uint32_t lbo, uint32_t sbo, int bo_mode) {
  uint64_t d = umma_smem_desc(saddr, lbo, sbo);
  if (bo_mode == 1) d |= (uint64_t)((saddr >> 7) & 7) << 49;
  return d;
}

__global__ void __launch_bounds__(128) probe_kernel(const __grid_constant__ ProbeParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                 // [512 rows][128 B]
  uint8_t* sA2 = smem + A_TILE;       // second 64-channel chunk (MN-major test)
  uint8_t* sB = smem + 2 * A_TILE;    // K-major: [64 n][128 B]; MN-major: [128 k rows][128 B]
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 2 * A_TILE + 16384);
  uint64_t* done = bar + 1;
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<64>(slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar, 2 * A_TILE + (p.mn_major ? 16384 : 8192));
    tma_load_2d(sA, &p.tmA, bar, 0, 0);
    tma_load_2d(sA + A_TILE / 2, &p.tmA, bar, 0, 256);
    tma_load_2d(sA2, &p.tmA2, bar, 0, 0);
    tma_load_2d(sA2 + A_TILE / 2, &p.tmA2, bar, 0, 256);
    tma_load_2d(sB, &p.tmB, bar, 0, 0);
    mbar_wait(bar, 0);
    tcgen05_fence_after();
    const uint32_t a0 = smem_u32(sA) + p.off_rows * 128;
    const uint32_t sbo = p.sbo_rows * 128;
    if (!p.mn_major) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 0, 0);
      const uint64_t ad = desc_with_bo(a0, 16, sbo, p.bo_mode);
      const uint64_t bd = umma_smem_desc(smem_u32(sB), 16, 1024);
      for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem, ad + 2 * k, bd + 2 * k, idesc, k != 0);
    } else {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 64, 1, 1);
      for (int k = 0; k < 8; ++k) {  // 16 pixel rows per MMA = two 8-row groups
        const uint64_t ad = desc_with_bo(a0 + k * 2 * sbo, A_TILE, sbo, p.bo_mode);
        const uint64_t bd = umma_smem_desc(smem_u32(sB) + k * 2048, 16384, 1024);
        umma_bf16_ss(tmem, ad, bd, idesc, k != 0);
      }
    }
    tcgen05_commit(done);
  }
  mbar_wait(done, 0);
  tcgen05_fence_after();
  const int r = warp * 32 + lane;
  for (int c = 0; c < 2; ++c) {
    float v[32];
    tmem_ld_32x32b_x32(tmem + ((uint32_t)(warp * 32) << 16) + c * 32, v);
    for (int j = 0; j < 32; ++j) p.out[r * 64 + c * 32 + j] = v[j];
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<64>(tmem);
  }
}

static float bf(float x) { return __bfloat162float(__float2bfloat16(x)); }

int main() {
  std::vector<float> A(ROWS * 64), A2(ROWS * 64), B(128 * 64);
  srand(7);
  for (auto* v : {&A, &A2, &B})
    for (auto& x : *v) x = bf((rand() % 2001 - 1000) / 1000.0f);
  std::vector<__nv_bfloat16> hA(A.size()), hA2(A2.size()), hB(B.size());
  for (size_t i = 0; i < A.size(); ++i) hA[i] = __float2bfloat16(A[i]), hA2[i] = __float2bfloat16(A2[i]);
  for (size_t i = 0; i < B.size(); ++i) hB[i] = __float2bfloat16(B[i]);
  __nv_bfloat16 *dA, *dA2, *dB;
  float* dOut;
  cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dA2, hA2.size() * 2); cudaMalloc(&dB, hB.size() * 2);
  cudaMalloc(&dOut, 128 * 64 * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dA2, hA2.data(), hA2.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  const int SMEM = 2 * A_TILE + 16384 + 64 + 1024;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
  for (int mn = 0; mn < 2; ++mn) {
    ProbeParams p;
    const uint64_t dimsA[2] = {64, ROWS}, strA[1] = {128};
    const uint32_t boxA[2] = {64, 256};
    if (b2_encode_tmap_bf16(&p.tmA, dA, 2, dimsA, strA, boxA, 1) ||
        b2_encode_tmap_bf16(&p.tmA2, dA2, 2, dimsA, strA, boxA, 1)) { printf("tmap A failed\n"); return 1; }
    const uint64_t dimsB[2] = {64, (uint64_t)(mn ? 128 : 64)};
    const uint32_t boxB[2] = {64, (uint32_t)(mn ? 128 : 64)};
    if (b2_encode_tmap_bf16(&p.tmB, dB, 2, dimsB, strA, boxB, 1)) { printf("tmap B failed\n"); return 1; }
    p.out = dOut;
    p.mn_major = mn;
    const int offs[] = {0, 8, 1, 3, 11, 19}, sbos[] = {8, 16, 10, 18};
    for (int sbo : sbos)
      for (int off : offs)
        for (int bo = 0; bo < 2; ++bo) {
          p.off_rows = off; p.sbo_rows = sbo; p.bo_mode = bo;
          cudaMemset(dOut, 0, 128 * 64 * 4);
          probe_kernel<<<1, 128, SMEM>>>(p);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); return 2; }
          std::vector<float> out(128 * 64);
          cudaMemcpy(out.data(), dOut, out.size() * 4, cudaMemcpyDeviceToHost);
          double worst = 0;
          for (int m = 0; m < 128; ++m)
            for (int n = 0; n < 64; ++n) {
              double ref = 0;
              if (!mn) {
                const int row = off + (m / 8) * sbo + m % 8;
                for (int k = 0; k < 64; ++k) ref += (double)A[row * 64 + k] * B[n * 64 + k];
              } else {
                const std::vector<float>& S = m < 64 ? A : A2;
                for (int k = 0; k < 128; ++k) {
                  const int row = off + (k / 8) * sbo + k % 8;
                  ref += (double)S[row * 64 + (m & 63)] * B[k * 64 + n];
                }
              }
              const double d = fabs(ref - out[m * 64 + n]);
              if (d > worst) worst = d;
            }
          printf("%s off=%2d sbo_rows=%2d base_offset=%s  max|err|=%.4g  %s\n", mn ? "MN-major" : "K-major ",
                 off, sbo, bo ? "(addr>>7)&7" : "0", worst, worst < 1e-2 ? "OK" : "WRONG");
        }
  }
  return 0;
}
