// ChannelAttention (advanced_models.py:43-61): x * sigmoid(fc(avgpool(x)) + fc(maxpool(x))),
// fc = 1x1 C -> C/r -> C without bias, shared by both branches.  Forward = pool (two-phase
// reduction with first-argmax), a tiny per-sample MLP kernel, and a scale pass; backward mirrors it.
#include "ew.cuh"

namespace {

__device__ __forceinline__ float sigm(float x) { return 1.0f / (1.0f + __expf(-x)); }

// grid (rows, N).  part [N][rows][3][C] = sum, max, argmax (int bits) over this block's pixels.
__global__ void __launch_bounds__(kEwThreads) se_pool_kernel(const EwArgs a) {
  pdl_prologue();
  extern __shared__ float red[];  // [3][pl][C]
  const bf16* x = (const bf16*)a.p[0];
  float* part = (float*)a.p[1];
  const int HW = (int)a.i[1], C = (int)a.i[2];
  const long long ld = a.i[3];
  const int n = blockIdx.y, rows = gridDim.x;
  const int cv = C >> 3, pl = blockDim.x / cv > 0 ? blockDim.x / cv : 1;
  const int lc = threadIdx.x % cv, lp = threadIdx.x / cv;
  float s[8], m[8];
  int am[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s[j] = 0.0f; m[j] = -INFINITY; am[j] = 0; }
  if (lp < pl) {
    for (int p = blockIdx.x * pl + lp; p < HW; p += rows * pl) {
      const F8 v = ld_bf8(x + ((long long)n * HW + p) * ld + lc * 8);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s[j] += v.v[j];
        if (v.v[j] > m[j]) { m[j] = v.v[j]; am[j] = p; }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = lc * 8 + j;
      red[((size_t)0 * pl + lp) * C + c] = s[j];
      red[((size_t)1 * pl + lp) * C + c] = m[j];
      red[((size_t)2 * pl + lp) * C + c] = __int_as_float(am[j]);
    }
  }
  __syncthreads();
  float* row = part + ((size_t)n * rows + blockIdx.x) * 3 * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float ss = 0.0f, mm = -INFINITY;
    int aa = 0x7fffffff;
    for (int l = 0; l < pl; ++l) {
      ss += red[((size_t)0 * pl + l) * C + c];
      const float v = red[((size_t)1 * pl + l) * C + c];
      const int ai = __float_as_int(red[((size_t)2 * pl + l) * C + c]);
      if (v > mm || (v == mm && ai < aa)) { mm = v; aa = ai; }
    }
    row[c] = ss;
    row[C + c] = mm;
    row[2 * C + c] = __int_as_float(aa);
  }
}

// grid N, 256 threads.  saved [N][4][C] = avg, max, gate, argmax(int bits); hidden [N][2][Ch].
__global__ void __launch_bounds__(kEwThreads) se_mlp_kernel(const EwArgs a) {
  pdl_prologue();
  extern __shared__ float sm[];  // avg[C], mx[C], hid[2*Ch]
  const float* part = (const float*)a.p[0];
  const float* w1 = (const float*)a.p[1];
  const float* w2 = (const float*)a.p[2];
  float* saved = (float*)a.p[3];
  float* hidden = (float*)a.p[4];
  const int HW = (int)a.i[1], C = (int)a.i[2], Ch = (int)a.i[3], rows = (int)a.i[4];
  const int n = blockIdx.x;
  float* avg = sm;
  float* mx = sm + C;
  float* hid = sm + 2 * C;
  float* sv = saved + (size_t)n * 4 * C;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float ss = 0.0f, mm = -INFINITY;
    int aa = 0x7fffffff;
    for (int r = 0; r < rows; ++r) {
      const float* row = part + ((size_t)n * rows + r) * 3 * C;
      ss += row[c];
      const float v = row[C + c];
      const int ai = __float_as_int(row[2 * C + c]);
      if (v > mm || (v == mm && ai < aa)) { mm = v; aa = ai; }
    }
    avg[c] = ss / (float)HW;
    mx[c] = mm;
    sv[c] = avg[c];
    sv[C + c] = mm;
    sv[3 * C + c] = __int_as_float(aa);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int o = warp; o < 2 * Ch; o += nw) {  // hidden unit o % Ch of branch o / Ch
    const float* src = o < Ch ? avg : mx;
    const float* wr = w1 + (size_t)(o % Ch) * C;
    float s = 0.0f;
    for (int c = lane; c < C; c += 32) s = fmaf(wr[c], src[c], s);
    s = warp_sum(s);
    if (lane == 0) {
      s = fmaxf(s, 0.0f);
      hid[o] = s;
      hidden[(size_t)n * 2 * Ch + o] = s;
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.0f;
    for (int h = 0; h < Ch; ++h) s = fmaf(w2[(size_t)c * Ch + h], hid[h] + hid[Ch + h], s);
    sv[2 * C + c] = sigm(s);
  }
}

__global__ void __launch_bounds__(kEwThreads) se_scale_kernel(const EwArgs a) {
  pdl_prologue();
  const bf16* x = (const bf16*)a.p[0];
  const float* saved = (const float*)a.p[1];
  bf16* out = (bf16*)a.p[2];
  const int N = (int)a.i[0], HW = (int)a.i[1], C = (int)a.i[2];
  const long long ldx = a.i[3], ldo = a.i[4];
  const int cv = C >> 3;
  const long long total = (long long)N * HW * cv;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    long long p, nn;
    int c8, pl_unused;
    ew_divmod(idx, cv, p, c8);
    ew_divmod(p, HW, nn, pl_unused);
    const int c0 = c8 * 8, n = (int)nn;
    F8 v = ld_bf8(x + p * ldx + c0);
    const F8 g = ld_f8(saved + ((size_t)n * 4 + 2) * C + c0);
#pragma unroll
    for (int j = 0; j < 8; ++j) v.v[j] *= g.v[j];
    st_bf8(out + p * ldo + c0, v);
  }
}

// grid (rows, N): part [N][rows][C] = sum_p dout*x
__global__ void __launch_bounds__(kEwThreads) se_bwd_reduce_kernel(const EwArgs a) {
  pdl_prologue();
  extern __shared__ float red[];
  const bf16* dout = (const bf16*)a.p[0];
  const bf16* x = (const bf16*)a.p[1];
  const int HW = (int)a.i[1], C = (int)a.i[2];
  const long long ldd = a.i[3], ldx = a.i[4];
  const int n = blockIdx.y;
  float* part = (float*)a.p[2] + (size_t)n * gridDim.x * C;
  ew_channel_reduce<1>((long long)n * HW, (long long)(n + 1) * HW, C, part, red,
                       [&](long long p, int c0, float (&acc)[1][8]) {
                         const F8 d = ld_bf8(dout + p * ldd + c0), v = ld_bf8(x + p * ldx + c0);
#pragma unroll
                         for (int j = 0; j < 8; ++j) acc[0][j] = fmaf(d.v[j], v.v[j], acc[0][j]);
                       });
}

// Backward of the shared MLP, two launches.  (1) grid = N samples: gate gradient ds[C], hidden
// gradients dha/dhm[Ch] and the pooled-input gradients dpool; ds / dh are parked behind the
// [N][rows][C] block partials inside `part` (its forward shape was [N][rows][3][C]).
// (2) grid over the weight elements: dW2[c][h] and dW1[h][c] summed over samples in sample
// order (deterministic).
__global__ void __launch_bounds__(kEwThreads) se_mlp_bwd_sample_kernel(const EwArgs a) {
  pdl_prologue();
  extern __shared__ float sm[];  // ds[C], dha[Ch], dhm[Ch]
  float* part = (float*)a.p[0];
  const float* w1 = (const float*)a.p[1];
  const float* w2 = (const float*)a.p[2];
  const float* saved = (const float*)a.p[3];
  const float* hidden = (const float*)a.p[4];
  float* dpool = (float*)a.p[7];
  const int N = (int)a.i[0], C = (int)a.i[2], Ch = (int)a.i[3], rows = (int)a.i[4];
  const int n = blockIdx.x;
  float* ds = sm;
  float* dha = sm + C;
  float* dhm = dha + Ch;
  float* ds_out = part + (size_t)N * rows * C + (size_t)n * C;
  float* dh_out = part + (size_t)2 * N * rows * C + (size_t)n * 2 * Ch;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const float* sv = saved + (size_t)n * 4 * C;
  const float* hid = hidden + (size_t)n * 2 * Ch;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.0f;
    for (int r = 0; r < rows; ++r) s += part[((size_t)n * rows + r) * C + c];
    const float g = sv[2 * C + c];
    const float v = s * g * (1.0f - g);
    ds[c] = v;
    ds_out[c] = v;
  }
  __syncthreads();
  for (int h = warp; h < Ch; h += nw) {
    float s = 0.0f;
    for (int c = lane; c < C; c += 32) s = fmaf(ds[c], w2[(size_t)c * Ch + h], s);
    s = warp_sum(s);
    if (lane == 0) {
      dha[h] = hid[h] > 0.0f ? s : 0.0f;
      dhm[h] = hid[Ch + h] > 0.0f ? s : 0.0f;
      dh_out[h] = dha[h];
      dh_out[Ch + h] = dhm[h];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float da = 0.0f, dm = 0.0f;
    for (int h = 0; h < Ch; ++h) {
      const float w = w1[(size_t)h * C + c];
      da = fmaf(dha[h], w, da);
      dm = fmaf(dhm[h], w, dm);
    }
    dpool[((size_t)n * 2) * C + c] = da;
    dpool[((size_t)n * 2 + 1) * C + c] = dm;
  }
}

__global__ void __launch_bounds__(kEwThreads) se_mlp_bwd_weight_kernel(const EwArgs a) {
  pdl_prologue();
  const float* __restrict__ part = (const float*)a.p[0];
  const float* __restrict__ saved = (const float*)a.p[3];
  const float* __restrict__ hidden = (const float*)a.p[4];
  float* dW1 = (float*)a.p[5];
  float* dW2 = (float*)a.p[6];
  const int N = (int)a.i[0], C = (int)a.i[2], Ch = (int)a.i[3], rows = (int)a.i[4];
  const bool accumulate = a.i[5] != 0;
  const float* __restrict__ ds = part + (size_t)N * rows * C;
  const float* __restrict__ dh = part + (size_t)2 * N * rows * C;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  // the loads of 8 samples are issued together; the fma chain keeps the sample order
  if (e < C * Ch) {  // dW2[c][h] = sum_n ds[n][c] * (hid_a[n][h] + hid_m[n][h])
    const int c = e / Ch, h = e % Ch;
    float s = 0.0f;
    for (int n0 = 0; n0 < N; n0 += 8) {
      float x[8], y[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int n = n0 + u;
        const bool ok = n < N;
        x[u] = ok ? __ldg(ds + (size_t)n * C + c) : 0.0f;
        y[u] = ok ? __ldg(hidden + (size_t)n * 2 * Ch + h) + __ldg(hidden + (size_t)n * 2 * Ch + Ch + h) : 0.0f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (n0 + u < N) s = fmaf(x[u], y[u], s);
    }
    dW2[e] = accumulate ? dW2[e] + s : s;
  } else if (e < 2 * C * Ch) {  // dW1[h][c] = sum_n dha[n][h]*avg[n][c] + dhm[n][h]*max[n][c]
    const int e1 = e - C * Ch;
    const int h = e1 / C, c = e1 % C;
    float s = 0.0f;
    for (int n0 = 0; n0 < N; n0 += 8) {
      float da[8], dm[8], va[8], vm[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int n = n0 + u;
        const bool ok = n < N;
        const float* sv = saved + (size_t)n * 4 * C;
        da[u] = ok ? __ldg(dh + (size_t)n * 2 * Ch + h) : 0.0f;
        dm[u] = ok ? __ldg(dh + (size_t)n * 2 * Ch + Ch + h) : 0.0f;
        va[u] = ok ? __ldg(sv + c) : 0.0f;
        vm[u] = ok ? __ldg(sv + C + c) : 0.0f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (n0 + u < N) {
          s = fmaf(da[u], va[u], s);
          s = fmaf(dm[u], vm[u], s);
        }
    }
    dW1[e1] = accumulate ? dW1[e1] + s : s;
  }
}

__global__ void __launch_bounds__(kEwThreads) se_bwd_apply_kernel(const EwArgs a) {
  pdl_prologue();
  const bf16* dout = (const bf16*)a.p[0];
  const float* saved = (const float*)a.p[1];
  const float* dpool = (const float*)a.p[2];
  bf16* dx = (bf16*)a.p[3];
  const int N = (int)a.i[0], HW = (int)a.i[1], C = (int)a.i[2];
  const long long ldd = a.i[3], ldx = a.i[4];
  const int cv = C >> 3;
  const float inv_hw = 1.0f / (float)HW;
  const long long total = (long long)N * HW * cv;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    long long p, nn;
    int c8, pl;
    ew_divmod(idx, cv, p, c8);
    ew_divmod(p, HW, nn, pl);
    const int c0 = c8 * 8, n = (int)nn;
    const float* sv = saved + (size_t)n * 4 * C;
    F8 v = ld_bf8(dout + p * ldd + c0);
    const F8 g = ld_f8(sv + 2 * C + c0), am = ld_f8(sv + 3 * C + c0);
    const F8 da = ld_f8(dpool + ((size_t)n * 2) * C + c0), dm = ld_f8(dpool + ((size_t)n * 2 + 1) * C + c0);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      v.v[j] = fmaf(v.v[j], g.v[j], da.v[j] * inv_hw);
      if (__float_as_int(am.v[j]) == pl) v.v[j] += dm.v[j];
    }
    st_bf8(dx + p * ldx + c0, v);
  }
}

#define EW_LAUNCHER(name, kernel)                                                         \
  int name(const EwArgs& a, dim3 grid, dim3 block, size_t smem, cudaStream_t st) {        \
    B2_CHECK_CUDA(b2_launch(kernel, grid, block, smem, st, a));                                                 \
    B2_CHECK_CUDA(cudaGetLastError());                                                    \
    return B2SEG_OK;                                                                      \
  }
EW_LAUNCHER(launch_se_pool, se_pool_kernel)
EW_LAUNCHER(launch_se_mlp, se_mlp_kernel)
EW_LAUNCHER(launch_se_scale, se_scale_kernel)
EW_LAUNCHER(launch_se_bwd_reduce, se_bwd_reduce_kernel)
int launch_se_mlp_bwd(const EwArgs& a, dim3 grid, dim3 block, size_t smem, cudaStream_t st) {
  B2_CHECK_CUDA(b2_launch(se_mlp_bwd_sample_kernel, grid, block, smem, st, a));
  B2_CHECK_CUDA(cudaGetLastError());
  const int elems = 2 * (int)a.i[2] * (int)a.i[3];
  B2_CHECK_CUDA(b2_launch(se_mlp_bwd_weight_kernel, dim3((elems + kEwThreads - 1) / kEwThreads), dim3(kEwThreads), 0, st, a));
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}
EW_LAUNCHER(launch_se_bwd_apply, se_bwd_apply_kernel)

}  // namespace

int ew_se_rows(int HW, int C) {
  int r = ew_rows(HW, C);
  return r > 64 ? 64 : r;
}

int ew_plan_se(const b2seg_ew_desc* d, EwPlan* pl) {
  const int64_t* i = d->i;
  void* const* p = d->p;
  pl->smem = 0;
  pl->launches = 1;
  pl->block = dim3(kEwThreads);
  const int N = (int)i[0], HW = (int)i[1], C = (int)i[2];
  B2_REQUIRE(N >= 1 && N <= 65535 && HW >= 1 && C >= 16 && C % 16 == 0 && C <= 2048,
             "se: bad shape N=%d HW=%d C=%d", N, HW, C);
  const int rows = ew_se_rows(HW, C);
  switch (d->kind) {
    case B2SEG_EW_SE_POOL:
      B2_REQUIRE(p[0] && p[1] && i[3] % 8 == 0, "se_pool: bad arguments");
      pl->fn = launch_se_pool;
      pl->grid = dim3(rows, N);
      pl->smem = (size_t)3 * ew_pl(C) * C * sizeof(float);
      return B2SEG_OK;
    case B2SEG_EW_SE_MLP:
      B2_REQUIRE(p[0] && p[1] && p[2] && p[3] && p[4] && i[3] >= 1 && i[4] == rows, "se_mlp: bad arguments");
      pl->fn = launch_se_mlp;
      pl->grid = dim3(N);
      pl->smem = (size_t)(2 * C + 2 * i[3]) * sizeof(float);
      return B2SEG_OK;
    case B2SEG_EW_SE_SCALE:
      B2_REQUIRE(p[0] && p[1] && p[2] && i[3] % 8 == 0 && i[4] % 8 == 0, "se_scale: bad arguments");
      pl->fn = launch_se_scale;
      pl->grid = dim3(ew_resident_blocks(se_scale_kernel, (long long)N * HW * (C / 8)));
      return B2SEG_OK;
    case B2SEG_EW_SE_BWD_REDUCE:
      B2_REQUIRE(p[0] && p[1] && p[2] && i[3] % 8 == 0 && i[4] % 8 == 0, "se_bwd_reduce: bad arguments");
      pl->fn = launch_se_bwd_reduce;
      pl->grid = dim3(rows, N);
      pl->smem = (size_t)ew_pl(C) * C * sizeof(float);
      return B2SEG_OK;
    case B2SEG_EW_SE_MLP_BWD:
      B2_REQUIRE(p[0] && p[1] && p[2] && p[3] && p[4] && p[5] && p[6] && p[7] && i[3] >= 1 && i[4] == rows,
                 "se_mlp_bwd: bad arguments");
      pl->fn = launch_se_mlp_bwd;
      pl->grid = dim3(N);
      pl->launches = 2;
      pl->smem = (size_t)(C + 2 * i[3]) * sizeof(float);
      return B2SEG_OK;
    case B2SEG_EW_SE_BWD_APPLY:
      B2_REQUIRE(p[0] && p[1] && p[2] && p[3] && i[3] % 8 == 0 && i[4] % 8 == 0, "se_bwd_apply: bad arguments");
      pl->fn = launch_se_bwd_apply;
      pl->grid = dim3(ew_resident_blocks(se_bwd_apply_kernel, (long long)N * HW * (C / 8)));
      return B2SEG_OK;
  }
  b2seg_set_error("ew_plan_se: kind %d not handled here", d->kind);
  return B2SEG_EINVAL;
}
