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
  int a_row_off;   // A operand starts this many 128-byte rows into the buffer (halo-tile taps: conv3x3.cu)
  int a_sbo;       // A 8-row-group stride in bytes (1024 = dense tile, 1280 = 10-pixel halo pitch); 0 = default
};

// A: 32 KB (two 64-wide M atoms x 128 K rows, or 128 M rows x 64 K), B: 64 KB (up to N = 256)
constexpr int kABytes = 65536, kBBytes = 65536;

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
    const uint32_t asbo = p.a_sbo ? (uint32_t)p.a_sbo : 1024u;
    const uint32_t abase = smem_u32(smem) + (uint32_t)p.a_row_off * 128u;
    const uint64_t ad0 = A_MN ? umma_smem_desc(abase, 16384, asbo) : umma_smem_desc(abase, 16, asbo);
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

// Burst probe: groups of `burst` MMAs separated by `gap` idle cycles, each group committed and awaited on its own
// (what a per-tile accumulator hand-off does).  Reports cycles from the first issue of a group to the
// arrival of its commit: burst * (steady MMA time) + whatever a cold start of the tensor pipe costs.
template <int N>
__global__ void __launch_bounds__(128) burst_kernel(const RateParams p, int burst, int gap, int groups, int overlap) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* done = reinterpret_cast<uint64_t*>(smem + kABytes + kBBytes);
  uint32_t* slot = reinterpret_cast<uint32_t*>(done + 2);
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (kABytes + kBBytes) / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + (uint32_t)(i & 0xff);
  if (threadIdx.x == 0) {
    mbar_init(&done[0], 1);
    mbar_init(&done[1], 1);
    fence_barrier_init();
  }
  fence_proxy_async_smem();
  if (warp == 0) tmem_alloc<2 * N>(slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *slot;
  long long total = 0, issue = 0;
  if (warp == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N, 0, 0);
    const uint64_t ad0 = umma_smem_desc(smem_u32(smem), 16, 1024);
    const uint64_t bd0 = umma_smem_desc(smem_u32(smem) + kABytes, 16, 1024);
    for (int g = 0; g < groups; ++g) {
      const long long tg = clock64();
      while (clock64() - tg < gap) {}
      const long long t0 = clock64();
      if (elect_one_sync()) {
        for (int i = 0; i < burst; i += 4) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem + (g & 1) * N, ad0 + 2 * k, bd0 + 2 * k, idesc, (i | k) != 0);
        }
        tcgen05_commit(&done[g & 1]);
      }
      __syncwarp();
      const long long t1 = clock64();
      // overlap = 1: wait for the PREVIOUS group's commit only (two accumulators in flight)
      if (!overlap) mbar_wait(&done[g & 1], (g >> 1) & 1);
      else if (g > 0) mbar_wait(&done[(g - 1) & 1], ((g - 1) >> 1) & 1);
      tcgen05_fence_after();
      const long long t2 = clock64();
      issue += t1 - t0;
      total += t2 - t0;
    }
    if (overlap) mbar_wait(&done[(groups - 1) & 1], ((groups - 1) >> 1) & 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) { p.out[blockIdx.x] = total; p.out[blockIdx.x + 512] = issue; }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<2 * N>(tmem);
  }
}

// Poll probe: groups of `burst` MMAs with a wait on an ALREADY COMPLETED mbarrier between groups (what a
// pipeline stage hand-off costs the issuing warp when the data is long there).  mode 0: nothing between
// groups, 1: mbarrier.try_wait, 2: mbarrier.test_wait, 3: volatile ld.shared of the barrier word.
template <int N>
__global__ void __launch_bounds__(128) poll_kernel(const RateParams p, int burst, int groups, int mode) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* done = reinterpret_cast<uint64_t*>(smem + kABytes + kBBytes);
  uint32_t* slot = reinterpret_cast<uint32_t*>(done + 2);
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (kABytes + kBBytes) / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + (uint32_t)(i & 0xff);
  if (threadIdx.x == 0) {
    mbar_init(&done[0], 1);
    mbar_init(&done[1], 1);
    fence_barrier_init();
    mbar_arrive(&done[1]);  // phase 0 of done[1] is complete from now on
  }
  fence_proxy_async_smem();
  if (warp == 0) tmem_alloc<N>(slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *slot;
  long long total = 0;
  if (warp == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N, 0, 0);
    const uint64_t ad0 = umma_smem_desc(smem_u32(smem), 16, 1024);
    const uint64_t bd0 = umma_smem_desc(smem_u32(smem) + kABytes, 16, 1024);
    uint32_t sink = 0;
    const long long t0 = clock64();
    for (int g = 0; g < groups; ++g) {
      if (mode == 1) { while (!mbar_try_wait(&done[1], 0)) {} }
      else if (mode == 2) { while (!mbar_test_wait(&done[1], 0)) {} }
      else if (mode == 3) { sink += *reinterpret_cast<volatile uint32_t*>(&done[1]); }
      if (elect_one_sync()) {
        for (int i = 0; i < burst; i += 4) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem, ad0 + 2 * k, bd0 + 2 * k, idesc, (g | i | k) != 0);
        }
      }
      __syncwarp();
    }
    if (elect_one_sync()) tcgen05_commit(&done[0]);
    __syncwarp();
    mbar_wait(&done[0], 0);
    tcgen05_fence_after();
    total = clock64() - t0 + (sink == 0x12345u);
  }
  __syncthreads();
  if (threadIdx.x == 0) p.out[blockIdx.x] = total;
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<N>(tmem);
  }
}

// Single-thread role probe: ONE elected thread runs  [try_wait on a completed barrier] x nwait,
// `burst` MMAs, tcgen05.commit x ncommit  per group, all inside one elect region (conv3x3.cu's MMA role).
template <int N>
__global__ void __launch_bounds__(128) role_kernel(const RateParams p, int burst, int groups, int nwait, int ncommit) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* done = reinterpret_cast<uint64_t*>(smem + kABytes + kBBytes);  // [0] final, [1] complete, [2..3] sinks
  uint32_t* slot = reinterpret_cast<uint32_t*>(done + 4);
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (kABytes + kBBytes) / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + (uint32_t)(i & 0xff);
  if (threadIdx.x == 0) {
    mbar_init(&done[0], 1);
    mbar_init(&done[1], 1);
    mbar_init(&done[2], 1);
    mbar_init(&done[3], 1);
    fence_barrier_init();
    mbar_arrive(&done[1]);
  }
  fence_proxy_async_smem();
  if (warp == 0) tmem_alloc<2 * N>(slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *slot;
  long long total = 0;
  if (warp == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(128, N, 0, 0);
    const uint64_t ad0 = umma_smem_desc(smem_u32(smem), 16, 1024);
    const uint64_t bd0 = umma_smem_desc(smem_u32(smem) + kABytes, 16, 1024);
    const long long t0 = clock64();
    if (elect_one_sync()) {
      for (int g = 0; g < groups; ++g) {
        for (int w = 0; w < nwait; ++w) { while (!mbar_try_wait(&done[1], 0)) {} }
        tcgen05_fence_after();
        for (int i = 0; i < burst; i += 4) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem + (g & 1) * N, ad0 + 2 * k, bd0 + 2 * k, idesc, (i | k) != 0);
        }
        for (int c = 0; c < ncommit; ++c) tcgen05_commit(&done[2 + c]);
      }
      tcgen05_commit(&done[0]);
    }
    __syncwarp();
    mbar_wait(&done[0], 0);
    tcgen05_fence_after();
    total = clock64() - t0;
  }
  __syncthreads();
  if (threadIdx.x == 0) p.out[blockIdx.x] = total;
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<2 * N>(tmem);
  }
}

template <int N>
static void run_role(long long* dOut, int grid, int burst, int nwait, int ncommit) {
  const int SMEM = kABytes + kBBytes + 128 + 1024, groups = 256;
  cudaFuncSetAttribute(role_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
  RateParams p{0, dOut, 0, 0};
  for (int rep = 0; rep < 2; ++rep) role_kernel<N><<<grid, 128, SMEM>>>(p, burst, groups, nwait, ncommit);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); exit(2); }
  std::vector<long long> h(1024);
  cudaMemcpy(h.data(), dOut, 1024 * sizeof(long long), cudaMemcpyDeviceToHost);
  double s = 0;
  for (int i = 0; i < grid; ++i) s += (double)h[i];
  printf("N=%3d one thread: %d waits + %2d MMAs + %d commits per group: %7.1f cycles per group (%5.1f per MMA)\n", N, nwait, burst,
         ncommit, s / grid / groups, s / grid / groups / burst);
}

template <int N>
static void run_poll(long long* dOut, int grid, int burst, int mode) {
  const int SMEM = kABytes + kBBytes + 64 + 1024, groups = 256;
  cudaFuncSetAttribute(poll_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
  RateParams p{0, dOut, 0, 0};
  for (int rep = 0; rep < 2; ++rep) poll_kernel<N><<<grid, 128, SMEM>>>(p, burst, groups, mode);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); exit(2); }
  std::vector<long long> h(1024);
  cudaMemcpy(h.data(), dOut, 1024 * sizeof(long long), cudaMemcpyDeviceToHost);
  double s = 0;
  for (int i = 0; i < grid; ++i) s += (double)h[i];
  const char* names[4] = {"nothing", "mbarrier.try_wait (complete)", "mbarrier.test_wait (complete)", "ld.shared of the barrier word"};
  printf("N=%3d groups of %2d MMAs, between groups: %-30s %7.1f cycles per group (%5.1f per MMA)\n", N, burst, names[mode],
         s / grid / groups, s / grid / groups / burst);
}

template <int N>
static void run_burst(long long* dOut, int grid, int burst, int gap, int overlap) {
  const int SMEM = kABytes + kBBytes + 64 + 1024, groups = 64;
  cudaFuncSetAttribute(burst_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
  RateParams p{0, dOut, 0, 0};
  for (int rep = 0; rep < 2; ++rep) burst_kernel<N><<<grid, 128, SMEM>>>(p, burst, gap, groups, overlap);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); exit(2); }
  std::vector<long long> h(1024);
  cudaMemcpy(h.data(), dOut, 1024 * sizeof(long long), cudaMemcpyDeviceToHost);
  double s = 0, si = 0;
  for (int i = 0; i < grid; ++i) { s += (double)h[i]; si += (double)h[i + 512]; }
  printf("N=%3d burst %3d MMAs, idle gap %5d cycles, %s: %7.0f cycles per group to completion (%5.1f per MMA), issue loop %7.0f\n",
         N, burst, gap, overlap ? "next group issued before this one completes" : "group awaited before the next", s / grid / groups,
         s / grid / groups / burst, si / grid / groups);
}

template <int N, int A_MN, int B_MN>
static void run(long long* dOut, int grid, int iters, int a_row_off = 0, int a_sbo = 0) {
  const int SMEM = kABytes + kBBytes + 64 + 1024;
  cudaFuncSetAttribute(rate_kernel<N, A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
  RateParams p{iters, dOut, a_row_off, a_sbo};
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
  if (a_row_off || a_sbo) printf("[A start +%d rows, SBO %d] ", a_row_off, a_sbo ? a_sbo : 1024);
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
  printf("A operand read out of a halo tile (unaligned start row and / or 8-row-group stride != 1024 B):\n");
  const int offs[6] = {0, 1, 2, 4, 8, 11};
  const int sbos[3] = {1024, 1280, 2048};
  for (int si = 0; si < 3; ++si)
    for (int oi = 0; oi < 6; ++oi) {
      run<64, 0, 0>(dOut, sms, iters, offs[oi], sbos[si]);
    }
  for (int oi = 0; oi < 6; ++oi) run<32, 0, 0>(dOut, sms, iters, offs[oi], 1280);
  for (int oi = 0; oi < 6; ++oi) run<128, 0, 0>(dOut, sms, iters, offs[oi], 1280);
  for (int oi = 0; oi < 6; ++oi) run<64, 1, 1>(dOut, sms, iters, offs[oi], 1280);
  printf("one elected thread running the whole issue loop:\n");
  for (int b = 4; b <= 36; b *= 3) {
    run_role<64>(dOut, sms, b, 0, 0);
    run_role<64>(dOut, sms, b, 2, 0);
    run_role<64>(dOut, sms, b, 0, 1);
    run_role<64>(dOut, sms, b, 0, 2);
    run_role<64>(dOut, sms, b, 2, 2);
  }
  run_role<128>(dOut, sms, 4, 1, 1);
  run_role<128>(dOut, sms, 36, 2, 2);
  printf("cost of polling a completed mbarrier between MMA groups (issuing warp):\n");
  for (int b = 4; b <= 36; b *= 3)
    for (int m = 0; m < 4; ++m) run_poll<64>(dOut, sms, b, m);
  for (int m = 0; m < 4; ++m) run_poll<128>(dOut, sms, 4, m);
  printf("bursts of MMAs with idle gaps (does an idle tensor pipe cost a restart?):\n");
  {
    const int gaps[6] = {0, 100, 300, 1000, 3000, 10000};
    for (int gi = 0; gi < 6; ++gi) run_burst<64>(dOut, sms, 36, gaps[gi], 0);
    for (int gi = 0; gi < 6; ++gi) run_burst<64>(dOut, sms, 108, gaps[gi], 0);
    for (int gi = 0; gi < 6; ++gi) run_burst<64>(dOut, sms, 36, gaps[gi], 1);
    for (int gi = 0; gi < 3; ++gi) run_burst<64>(dOut, sms, 4, gaps[gi], 0);
    for (int gi = 0; gi < 3; ++gi) run_burst<64>(dOut, sms, 8, gaps[gi], 1);
    for (int gi = 0; gi < 3; ++gi) run_burst<128>(dOut, sms, 36, gaps[gi], 0);
    run_burst<64>(dOut, 1, 36, 0, 0);
    run_burst<64>(dOut, 1, 36, 1000, 0);
  }
  printf("one SM alone (no power / clock interaction):\n");
  run<64, 0, 0>(dOut, 1, iters);
  run<64, 1, 1>(dOut, 1, iters);
  run<128, 1, 1>(dOut, 1, iters);
  return 0;
}
