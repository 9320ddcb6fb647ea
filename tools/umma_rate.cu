// Hardware probe (not product code): cycles per tcgen05.mma (kind::f16, bf16 in, fp32 out, M = 128,
// K = 16) as a function of N and of the operands' major-ness, operands resident in shared memory
// (128B swizzle), one CTA per SM.  Answers what bounds the weight-gradient GEMMs (both operands
// MN-major): the tensor core's transposed operand reads or something else.
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 tools/umma_rate.cu -o tools/umma_rate \
//        -L image-segmentation-project_b200/b2seg -lb2seg -lcuda -Xlinker -rpath=$PWD/image-segmentation-project_b200/b2seg
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#include "../image-segmentation-project_b200/csrc/common.cuh"

struct RateParams {
  int iters;       // MMAs issued per CTA
  long long* out;  // [grid] cycles from first issue to completion
};

// A: 32 KB (two 64-wide M atoms x 128 K rows, or 128 M rows x 64 K), B: 64 KB (up to N = 256)
constexpr int kABytes = 32768, kBBytes = 65536;

template <int N, int A_MN, int B_MN>
__global__ void __launch_bounds__(128) rate_kernel(const RateParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* done = reinterpret_cast<uint64_t*>(smem + kABytes + kBBytes);
  uint32_t* slot = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5;
  // finite operand bits (the values do not matter, NaN payloads might)
  for (int i = threadIdx.x; i < (kABytes + kBBytes) / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + (uint32_t)(i & 0xff);
  if (threadIdx.x == 0) {
    mbar_init(done, 1);
    fence_barrier_init();
  }
  fence_proxy_async_smem();
  if (warp == 0) tmem_alloc<(N < 32 ? 32 : N)>(slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *slot;
  long long t0 = 0, t1 = 0;
  if (warp == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N, A_MN, B_MN);
    // K-major: rows of 128 B (64 K elements), 8-row groups 1024 B apart, K step (16 el) = +32 B
    // MN-major: K rows of 128 B (64 MN elements), 64-wide atoms 16 KB apart, K step (16 rows) = +2048 B
    const uint64_t ad0 = A_MN ? umma_smem_desc(smem_u32(smem), 16384, 1024) : umma_smem_desc(smem_u32(smem), 16, 1024);
    const uint64_t bd0 = B_MN ? umma_smem_desc(smem_u32(smem) + kABytes, 16384, 1024)
                              : umma_smem_desc(smem_u32(smem) + kABytes, 16, 1024);
    const uint64_t astep = A_MN ? 128 : 2, bstep = B_MN ? 128 : 2;
    t0 = clock64();
    if (elect_one_sync()) {
      for (int i = 0; i < p.iters; i += 4) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16_ss(tmem, ad0 + astep * k, bd0 + bstep * k, idesc, (i | k) != 0);
      }
      tcgen05_commit(done);
    }
    __syncwarp();
  }
  mbar_wait(done, 0);
  tcgen05_fence_after();
  t1 = clock64();
  if (threadIdx.x == 0) p.out[blockIdx.x] = t1 - t0;
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<(N < 32 ? 32 : N)>(tmem);
  }
}

template <int N, int A_MN, int B_MN>
static void run(long long* dOut, int grid, int iters) {
  const int SMEM = kABytes + kBBytes + 64 + 1024;
  cudaFuncSetAttribute(rate_kernel<N, A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
  RateParams p{iters, dOut};
  for (int rep = 0; rep < 2; ++rep) rate_kernel<N, A_MN, B_MN><<<grid, 128, SMEM>>>(p);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); exit(2); }
  std::vector<long long> h(grid);
  cudaMemcpy(h.data(), dOut, grid * sizeof(long long), cudaMemcpyDeviceToHost);
  double s = 0;
  long long mx = 0;
  for (long long v : h) { s += (double)v; if (v > mx) mx = v; }
  const double cyc = s / grid / iters;
  const double flop = 2.0 * 128 * N * 16;
  printf("M=128 N=%3d  A %s  B %s   %7.1f cycles/MMA (max CTA %7.1f)   %6.0f FLOP/cycle/SM   operand bytes/MMA %5d\n", N,
         A_MN ? "MN-major" : "K-major ", B_MN ? "MN-major" : "K-major ", cyc, (double)mx / iters, flop / cyc,
         (128 + N) * 16 * 2);
}

int main() {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long* dOut;
  cudaMalloc(&dOut, 1024 * sizeof(long long));
  const int iters = 4096;
  printf("tcgen05.mma kind::f16 rate probe: %d SMs, %d MMAs per CTA, all SMs busy\n", sms, iters);
  run<64, 0, 0>(dOut, sms, iters);
  run<128, 0, 0>(dOut, sms, iters);
  run<256, 0, 0>(dOut, sms, iters);
  run<64, 1, 1>(dOut, sms, iters);
  run<128, 1, 1>(dOut, sms, iters);
  run<256, 1, 1>(dOut, sms, iters);
  run<64, 1, 0>(dOut, sms, iters);
  run<64, 0, 1>(dOut, sms, iters);
  run<128, 1, 0>(dOut, sms, iters);
  run<128, 0, 1>(dOut, sms, iters);
  run<32, 0, 0>(dOut, sms, iters);
  run<32, 1, 1>(dOut, sms, iters);
  printf("one SM alone (no power / clock interaction):\n");
  run<64, 0, 0>(dOut, 1, iters);
  run<64, 1, 1>(dOut, 1, iters);
  run<128, 1, 1>(dOut, 1, iters);
  return 0;
}
