// 1-output-channel 1x1 convolutions (conv_final advanced_models.py:160,350; AttentionGate.psi
// :20-24), the attention-gate tail (:34-40) and ChannelAttention squeeze-excite (:43-61),
// forward and backward.  All bandwidth-bound; a sub-warp group of g lanes owns one pixel so a
// 16-channel pixel (32 B) does not waste a warp.
#include "ew.cuh"

namespace {

__device__ __forceinline__ float group_sum(float v, int g) {
  for (int o = g >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int group_of(int C) {
  const int cv = C >> 3;
  int g = 1;
  while (g < cv && g < 32) g <<= 1;
  return g;
}
// block-wide sum of `n` per-thread values -> thread 0 (fixed order)
template <int N>
__device__ __forceinline__ void block_sum(float (&v)[N], float* sh /* [N][8] */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    v[k] = warp_sum(v[k]);
    if (lane == 0) sh[k * 8 + warp] = v[k];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < N; ++k) {
      float s = 0.0f;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += sh[k * 8 + w];
      v[k] = s;
    }
  }
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + __expf(-x)); }

// ------------------------------------------------------------------- DOT1_FWD ----
// KMAX = channel vectors per lane (1 for C <= 256: every use in the U-Net); the lane's weights
// stay in registers and UNR pixels are in flight per thread.
template <int KMAX, int UNR>
__device__ __forceinline__ void dot1_fwd_body(const EwArgs& a) {
  const bf16* __restrict__ v = (const bf16*)a.p[0];
  const float* __restrict__ w = (const float*)a.p[1];
  const float* __restrict__ b = (const float*)a.p[2];
  float* __restrict__ out = (float*)a.p[3];
  float* __restrict__ partial = (float*)a.p[4];
  const long long P = a.i[0], ld = a.i[2];
  const int C = (int)a.i[1], cv = C >> 3, g = group_of(C);
  const int lg = threadIdx.x % g, grp = threadIdx.x / g, gpb = blockDim.x / g;
  const float bias = b ? b[0] : 0.0f;
  __shared__ float sh[16];
  float st[2] = {0.0f, 0.0f};
  F8 wr[KMAX];
#pragma unroll
  for (int k = 0; k < KMAX; ++k)
    if (lg + k * g < cv) wr[k] = ld_f8(w + (lg + k * g) * 8);
  const long long stride = (long long)gridDim.x * gpb;
  for (long long base = (long long)blockIdx.x * gpb; base < P; base += UNR * stride) {
    uint4 raw[UNR][KMAX];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const long long p = base + u * stride + grp;
#pragma unroll
      for (int k = 0; k < KMAX; ++k)
        if (p < P && lg + k * g < cv)
          raw[u][k] = *reinterpret_cast<const uint4*>(v + p * ld + (lg + k * g) * 8);
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const long long p = base + u * stride + grp;
      float s = 0.0f;
#pragma unroll
      for (int k = 0; k < KMAX; ++k)
        if (p < P && lg + k * g < cv) {
          const F8 x = unpack_bf8(raw[u][k]);
#pragma unroll
          for (int j = 0; j < 8; ++j) s = fmaf(x.v[j], wr[k].v[j], s);
        }
      s = group_sum(s, g) + bias;  // every lane takes part (base + u*stride is block-uniform)
      if (p < P && lg == 0) {
        out[p] = s;
        st[0] += s;
        st[1] = fmaf(s, s, st[1]);
      }
    }
  }
  if (partial) {
    block_sum<2>(st, sh);
    if (threadIdx.x == 0) {
      partial[blockIdx.x * 2] = st[0];
      partial[blockIdx.x * 2 + 1] = st[1];
    }
  }
}
__global__ void __launch_bounds__(kEwThreads) dot1_fwd_kernel(const EwArgs a) {
  pdl_prologue();
  dot1_fwd_body<1, 4>(a);
}
__global__ void __launch_bounds__(kEwThreads) dot1_fwd_wide_kernel(const EwArgs a) {  // C > 256
  pdl_prologue();
  dot1_fwd_body<4, 1>(a);
}

// ------------------------------------------------------------------- DOT1_BWD ----
// ds[p] = dscalar[p]  or  BatchNorm2d(1)-backward of it;  dv = ds * w;  partial = (dW[C], db).
template <int KMAX, int UNR>
__device__ __forceinline__ void dot1_bwd_body(const EwArgs& a, float* red) {
  const float* __restrict__ dsc = (const float*)a.p[0];
  const bf16* __restrict__ v = (const bf16*)a.p[1];
  const float* __restrict__ w = (const float*)a.p[2];
  bf16* __restrict__ dv = (bf16*)a.p[3];
  float* __restrict__ partial = (float*)a.p[4];
  const float* __restrict__ coef = (const float*)a.p[5];
  const float* __restrict__ ypre = (const float*)a.p[6];
  const float* __restrict__ bnp = (const float*)a.p[7];
  const long long P = a.i[0], ldv = a.i[2], lddv = a.i[3];
  const int C = (int)a.i[1], cv = C >> 3, g = group_of(C);
  const int lg = threadIdx.x % g, grp = threadIdx.x / g, gpb = blockDim.x / g;
  float c1 = 1.0f, c2 = 0.0f, c3 = 0.0f, mean = 0.0f, inv = 0.0f;
  if (coef) { c1 = coef[0]; c2 = coef[1]; c3 = coef[2]; mean = bnp[2]; inv = bnp[3]; }
  float acc[KMAX][8];
  F8 wr[KMAX];
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[k][j] = 0.0f;
    if (lg + k * g < cv) wr[k] = ld_f8(w + (lg + k * g) * 8);
  }
  float db = 0.0f;
  const long long stride = (long long)gridDim.x * gpb;
  for (long long p0 = (long long)blockIdx.x * gpb + grp; p0 < P; p0 += UNR * stride) {
    uint4 raw[UNR][KMAX];
    float ds[UNR], yp[UNR];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const long long p = p0 + u * stride;
      ds[u] = 0.0f;
      yp[u] = 0.0f;
      if (p < P) {
        ds[u] = dsc[p];
        if (coef) yp[u] = ypre[p];
#pragma unroll
        for (int k = 0; k < KMAX; ++k)
          if (lg + k * g < cv)
            raw[u][k] = *reinterpret_cast<const uint4*>(v + p * ldv + (lg + k * g) * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const long long p = p0 + u * stride;
      if (p >= P) continue;
      float d = ds[u];
      if (coef) d = c1 * (d - c2 - (yp[u] - mean) * inv * c3);
      if (lg == 0) db += d;
#pragma unroll
      for (int k = 0; k < KMAX; ++k) {
        const int c8 = lg + k * g;
        if (c8 < cv) {
          const F8 x = unpack_bf8(raw[u][k]);
          F8 o;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            acc[k][j] = fmaf(d, x.v[j], acc[k][j]);
            o.v[j] = d * wr[k].v[j];
          }
          st_bf8(dv + p * lddv + c8 * 8, o);
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    const int c8 = lg + k * g;
    if (c8 < cv)
#pragma unroll
      for (int j = 0; j < 8; ++j) red[(size_t)grp * C + c8 * 8 + j] = acc[k][j];
  }
  __shared__ float sh[8];
  float dbv[1] = {db};
  block_sum<1>(dbv, sh);  // contains __syncthreads -> red is visible below
  float* row = partial + (size_t)blockIdx.x * (C + 1);
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float s = 0.0f;
    for (int l = 0; l < gpb; ++l) s += red[(size_t)l * C + c];
    row[c] = s;
  }
  if (threadIdx.x == 0) row[C] = dbv[0];
}
__global__ void __launch_bounds__(kEwThreads) dot1_bwd_kernel(const EwArgs a) {
  pdl_prologue();
  extern __shared__ float red[];  // [gpb][C]
  dot1_bwd_body<1, 4>(a, red);
}
__global__ void __launch_bounds__(kEwThreads) dot1_bwd_wide_kernel(const EwArgs a) {  // C > 256
  pdl_prologue();
  extern __shared__ float red[];
  dot1_bwd_body<4, 1>(a, red);
}

// ----------------------------------------------------------------- GATE_APPLY ----
__global__ void __launch_bounds__(kEwThreads) gate_apply_kernel(const EwArgs a) {
  pdl_prologue();
  const float* pre = (const float*)a.p[0];
  const float* bnp = (const float*)a.p[1];
  const bf16* x = (const bf16*)a.p[2];
  bf16* out = (bf16*)a.p[3];
  float* psi_out = (float*)a.p[4];
  const long long P = a.i[0], ldx = a.i[2], ldo = a.i[3];
  const int C = (int)a.i[1], cv = C >> 3;
  const float scale = bnp[0], shift = bnp[1];
  const long long total = P * cv;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    long long p;
    int c8;
    ew_divmod(idx, cv, p, c8);
    const float psi = sigmoidf_(fmaf(pre[p], scale, shift));
    if (c8 == 0) psi_out[p] = psi;
    F8 v = ld_bf8(x + p * ldx + c8 * 8);
#pragma unroll
    for (int j = 0; j < 8; ++j) v.v[j] *= psi;
    st_bf8(out + p * ldo + c8 * 8, v);
  }
}

// ------------------------------------------------------------------- GATE_BWD ----
__global__ void __launch_bounds__(kEwThreads) gate_bwd_kernel(const EwArgs a) {
  pdl_prologue();
  const bf16* dout = (const bf16*)a.p[0];
  const bf16* x = (const bf16*)a.p[1];
  const float* psi = (const float*)a.p[2];
  const float* pre = (const float*)a.p[3];
  const float* bnp = (const float*)a.p[4];
  float* dz = (float*)a.p[5];
  bf16* dx = (bf16*)a.p[6];
  float* partial = (float*)a.p[7];
  const long long P = a.i[0], lddo = a.i[2], ldx = a.i[3], lddx = a.i[4];
  const bool accumulate = a.i[5] != 0;
  const int C = (int)a.i[1], cv = C >> 3, g = group_of(C);
  const int lg = threadIdx.x % g, grp = threadIdx.x / g, gpb = blockDim.x / g;
  const float mean = bnp[2], inv = bnp[3];
  __shared__ float sh[16];
  float st[2] = {0.0f, 0.0f};
  for (long long base = (long long)blockIdx.x * gpb; base < P; base += (long long)gridDim.x * gpb) {
    const long long p = base + grp;
    float s = 0.0f, ps = 0.0f;
    if (p < P) {
      ps = psi[p];
      for (int c8 = lg; c8 < cv; c8 += g) {
        const F8 d = ld_bf8(dout + p * lddo + c8 * 8), xv = ld_bf8(x + p * ldx + c8 * 8);
        F8 o;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          s = fmaf(d.v[j], xv.v[j], s);
          o.v[j] = d.v[j] * ps;
        }
        bf16* dst = dx + p * lddx + c8 * 8;
        if (accumulate) {
          const F8 old = ld_bf8(dst);
#pragma unroll
          for (int j = 0; j < 8; ++j) o.v[j] += old.v[j];
        }
        st_bf8(dst, o);
      }
    }
    s = group_sum(s, g);
    if (p < P && lg == 0) {
      const float z = s * ps * (1.0f - ps);
      dz[p] = z;
      st[0] += z;
      st[1] = fmaf(z, (pre[p] - mean) * inv, st[1]);
    }
  }
  block_sum<2>(st, sh);
  if (threadIdx.x == 0) {
    partial[blockIdx.x * 2] = st[0];
    partial[blockIdx.x * 2 + 1] = st[1];
  }
}

#define EW_LAUNCHER(name, kernel)                                                         \
  int name(const EwArgs& a, dim3 grid, dim3 block, size_t smem, cudaStream_t st) {        \
    B2_CHECK_CUDA(b2_launch(kernel, grid, block, smem, st, a));                                                 \
    B2_CHECK_CUDA(cudaGetLastError());                                                    \
    return B2SEG_OK;                                                                      \
  }
#define EW_LAUNCHER_BY_WIDTH(name, kernel, wide)                                           \
  int name(const EwArgs& a, dim3 grid, dim3 block, size_t smem, cudaStream_t st) {        \
    B2_CHECK_CUDA(b2_launch(a.i[1] <= 256 ? kernel : wide, grid, block, smem, st, a));    \
    B2_CHECK_CUDA(cudaGetLastError());                                                    \
    return B2SEG_OK;                                                                      \
  }
EW_LAUNCHER_BY_WIDTH(launch_dot1_fwd, dot1_fwd_kernel, dot1_fwd_wide_kernel)
EW_LAUNCHER_BY_WIDTH(launch_dot1_bwd, dot1_bwd_kernel, dot1_bwd_wide_kernel)
EW_LAUNCHER(launch_gate_apply, gate_apply_kernel)
EW_LAUNCHER(launch_gate_bwd, gate_bwd_kernel)

bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace

// rows (= blocks) of the per-pixel-group kernels
int ew_dot_rows(long long P, int C) {
  const int gpb = kEwThreads / ew_group(C);
  long long r = (P + (long long)gpb * 4 - 1) / ((long long)gpb * 4);
  if (r < 1) r = 1;
  if (r > B2SEG_EW_MAX_PARTIAL_ROWS) r = B2SEG_EW_MAX_PARTIAL_ROWS;
  return (int)r;
}

int ew_plan_attn(const b2seg_ew_desc* d, EwPlan* pl) {
  const int64_t* i = d->i;
  void* const* p = d->p;
  pl->smem = 0;
  pl->launches = 1;
  pl->block = dim3(kEwThreads);
  const int C = (int)i[1];
  B2_REQUIRE(C >= 8 && C % 8 == 0 && C <= 1024, "attn: C=%d must be a multiple of 8 in [8,1024]", C);
  B2_REQUIRE(i[0] >= 1, "attn: empty pixel range");
  switch (d->kind) {
    case B2SEG_EW_DOT1_FWD:
      B2_REQUIRE(p[0] && p[1] && p[3] && i[2] % 8 == 0 && al16(p[0]) && al16(p[1]), "dot1_fwd: bad arguments");
      pl->fn = launch_dot1_fwd;
      pl->grid = dim3(ew_dot_rows(i[0], C));
      return B2SEG_OK;
    case B2SEG_EW_DOT1_BWD:
      B2_REQUIRE(p[0] && p[1] && p[2] && p[3] && p[4] && i[2] % 8 == 0 && i[3] % 8 == 0 &&
                     al16(p[1]) && al16(p[2]) && al16(p[3]),
                 "dot1_bwd: bad arguments");
      B2_REQUIRE(!p[5] || (p[6] && p[7]), "dot1_bwd: BatchNorm path needs coef, ypre and bnp");
      pl->fn = launch_dot1_bwd;
      pl->grid = dim3(ew_dot_rows(i[0], C));
      pl->smem = (size_t)(kEwThreads / ew_group(C)) * C * sizeof(float);
      return B2SEG_OK;
    case B2SEG_EW_GATE_APPLY:
      B2_REQUIRE(p[0] && p[1] && p[2] && p[3] && p[4] && i[2] % 8 == 0 && i[3] % 8 == 0 &&
                     al16(p[2]) && al16(p[3]),
                 "gate_apply: bad arguments");
      pl->fn = launch_gate_apply;
      pl->grid = dim3(ew_resident_blocks(gate_apply_kernel, i[0] * (C / 8)));
      return B2SEG_OK;
    case B2SEG_EW_GATE_BWD:
      B2_REQUIRE(p[0] && p[1] && p[2] && p[3] && p[4] && p[5] && p[6] && p[7] && i[2] % 8 == 0 &&
                     i[3] % 8 == 0 && i[4] % 8 == 0 && al16(p[0]) && al16(p[1]) && al16(p[6]),
                 "gate_bwd: bad arguments");
      pl->fn = launch_gate_bwd;
      pl->grid = dim3(ew_dot_rows(i[0], C));
      return B2SEG_OK;
  }
  b2seg_set_error("ew_plan_attn: kind %d not handled here", d->kind);
  return B2SEG_EINVAL;
}
