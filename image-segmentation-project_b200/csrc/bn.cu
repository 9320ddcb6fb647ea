// BatchNorm2d forward/backward, per-channel column sums, weight packing, fill.
//
// Replaces torch native_batch_norm / native_batch_norm_backward / threshold_backward reached
// from every nn.BatchNorm2d + ReLU of the reference model (advanced_models.py:12,17,22,200-204
// and the torchvision BasicBlock/Bottleneck the encoder is built from, :81-87,111-117).
// All kernels are HBM-bound: 128-bit accesses, 8 channels per thread, no atomics.
#include "ew.cuh"

namespace {

// ------------------------------------------------------------------ BN_FINALIZE ----
// grid (C/32, slices), block (32 channels, 32 row lanes): see ew_rows_reduce.
__global__ void __launch_bounds__(1024) bn_finalize_kernel(const EwArgs a) {
  pdl_prologue();
  const float* partial = (const float*)a.p[0];
  const float* gamma = (const float*)a.p[1];
  const float* beta = (const float*)a.p[2];
  float* rmean = (float*)a.p[3];
  float* rvar = (float*)a.p[4];
  long long* nbt = (long long*)a.p[5];
  float* bnp = (float*)a.p[6];
  const int rows = (int)a.i[0], C = (int)a.i[1];
  const double count = (double)a.i[2];
  const bool training = a.i[3] != 0;
  const float eps = a.f[0], mom = a.f[1];
  ew_rows_reduce<2>(training ? partial : nullptr, rows, 2 * (long long)C, C, C, [&](int c, const double* v) {
    float mean, invstd;
    if (training) {
      const double m = v[0] / count;
      double var = v[1] / count - m * m;
      if (var < 0.0) var = 0.0;
      mean = (float)m;
      invstd = (float)(1.0 / sqrt(var + (double)eps));
      if (rmean) {
        const double unb = count > 1.0 ? var * count / (count - 1.0) : var;
        rmean[c] = (1.0f - mom) * rmean[c] + mom * mean;
        rvar[c] = (1.0f - mom) * rvar[c] + mom * (float)unb;
      }
      if (nbt && c == 0) *nbt += 1;
    } else {
      mean = rmean[c];
      invstd = 1.0f / sqrtf(rvar[c] + eps);
    }
    const float g = gamma ? gamma[c] : 1.0f, b = beta ? beta[c] : 0.0f;
    const float scale = g * invstd;
    bnp[c] = scale;
    bnp[C + c] = b - mean * scale;
    bnp[2 * C + c] = mean;
    bnp[3 * C + c] = invstd;
  });
}

// -------------------------------------------------------------------- BN_APPLY ----
// Per-channel constants live in registers: with C a power of two the 8-channel vector a thread
// owns never changes along its grid-stride walk (256 % (C/8) == 0), so there is no per-element
// index division and no per-element reload of scale/shift.
struct BnApplyCh {
  F8 sc, sf, rs, rf;
};
__device__ __forceinline__ uint4 bn_ld_raw8(const bf16* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ F8 bn_unpack8(const uint4& u) {
  F8 r;
  float2 t;
  t = bf2_to_f2(u.x); r.v[0] = t.x; r.v[1] = t.y;
  t = bf2_to_f2(u.y); r.v[2] = t.x; r.v[3] = t.y;
  t = bf2_to_f2(u.z); r.v[4] = t.x; r.v[5] = t.y;
  t = bf2_to_f2(u.w); r.v[6] = t.x; r.v[7] = t.y;
  return r;
}
__device__ __forceinline__ F8 bn_apply_one(const uint4& yraw, const uint4& rraw, bool has_res,
                                           const BnApplyCh& ch, bool rnorm, bool relu) {
  const F8 yv = bn_unpack8(yraw);
  F8 v;
#pragma unroll
  for (int j = 0; j < 8; ++j) v.v[j] = fmaf(yv.v[j], ch.sc.v[j], ch.sf.v[j]);
  if (has_res) {
    const F8 rv = bn_unpack8(rraw);
#pragma unroll
    for (int j = 0; j < 8; ++j)
      v.v[j] += rnorm ? fmaf(rv.v[j], ch.rs.v[j], ch.rf.v[j]) : rv.v[j];
  }
  if (relu) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v.v[j] = fmaxf(v.v[j], 0.0f);
  }
  return v;
}

__global__ void __launch_bounds__(kEwThreads) bn_apply_kernel(const EwArgs a) {
  pdl_prologue();
  const bf16* y = (const bf16*)a.p[0];
  const float* bnp = (const float*)a.p[1];
  bf16* out = (bf16*)a.p[2];
  const bf16* res = (const bf16*)a.p[3];
  const float* rbnp = (const float*)a.p[4];
  const long long P = a.i[0];
  const int C = (int)a.i[1];
  const long long ldy = a.i[2], ldo = a.i[3], ldr = a.i[4];
  const bool relu = a.i[5] != 0;
  const int cv = C >> 3;
  BnApplyCh ch;
  if (kEwThreads % cv == 0) {
    const int ppb = kEwThreads / cv;  // pixels per block per step
    const int c0 = (threadIdx.x % cv) * 8;
    const long long pstep = (long long)gridDim.x * ppb;
    ch.sc = ld_f8(bnp + c0);
    ch.sf = ld_f8(bnp + C + c0);
    if (res && rbnp) {
      ch.rs = ld_f8(rbnp + c0);
      ch.rf = ld_f8(rbnp + C + c0);
    }
    long long p = (long long)blockIdx.x * ppb + threadIdx.x / cv;
    // two pixels in flight per thread: all loads of a group are issued before the first store (four cost
    // registers and occupancy: forward graph 2.66 -> 2.77 ms)
    constexpr int NB = 2;
    for (; p + (NB - 1) * pstep < P; p += NB * pstep) {
      uint4 yv[NB], rv[NB];
#pragma unroll
      for (int b = 0; b < NB; ++b) {
        yv[b] = bn_ld_raw8(y + (p + b * pstep) * ldy + c0);
        rv[b] = make_uint4(0, 0, 0, 0);
        if (res) rv[b] = bn_ld_raw8(res + (p + b * pstep) * ldr + c0);
      }
#pragma unroll
      for (int b = 0; b < NB; ++b)
        st_bf8(out + (p + b * pstep) * ldo + c0, bn_apply_one(yv[b], rv[b], res != nullptr, ch, rbnp != nullptr, relu));
    }
    for (; p < P; p += pstep) {
      const uint4 y0 = bn_ld_raw8(y + p * ldy + c0);
      uint4 r0 = make_uint4(0, 0, 0, 0);
      if (res) r0 = bn_ld_raw8(res + p * ldr + c0);
      st_bf8(out + p * ldo + c0, bn_apply_one(y0, r0, res != nullptr, ch, rbnp != nullptr, relu));
    }
    return;
  }
  const long long total = P * cv;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long p = idx / cv;
    const int c0 = (int)(idx % cv) * 8;
    ch.sc = ld_f8(bnp + c0);
    ch.sf = ld_f8(bnp + C + c0);
    const uint4 y0 = bn_ld_raw8(y + p * ldy + c0);
    uint4 r0 = make_uint4(0, 0, 0, 0);
    if (res) {
      r0 = bn_ld_raw8(res + p * ldr + c0);
      if (rbnp) {
        ch.rs = ld_f8(rbnp + c0);
        ch.rf = ld_f8(rbnp + C + c0);
      }
    }
    st_bf8(out + p * ldo + c0, bn_apply_one(y0, r0, res != nullptr, ch, rbnp != nullptr, relu));
  }
}

// ------------------------------------------------------------------ BN backward ----
struct BnBwdIn {
  const bf16 *da1, *da2, *act, *y;
  const float* bnp;
  long long ld1, ld2, lda, ldy;
  int C;
  int self_mask;  // ReLU mask recomputed from y: [bf16(y*scale+shift) > 0] (no activation read)
};
__device__ __forceinline__ BnBwdIn bn_bwd_in(const EwArgs& a) {
  BnBwdIn s;
  s.da1 = (const bf16*)a.p[0]; s.da2 = (const bf16*)a.p[1];
  s.act = (const bf16*)a.p[2]; s.y = (const bf16*)a.p[3];
  s.bnp = (const float*)a.p[4];
  s.C = (int)a.i[1];
  s.ld1 = a.i[2]; s.ld2 = a.i[3]; s.lda = a.i[4]; s.ldy = a.i[5];
  s.self_mask = (int)a.i[9];
  return s;
}
// one pixel's operands as loaded (packed bf16: 4 registers per 8 channels until they are used)
struct BnBwdRaw {
  uint4 d1, d2, m, y, old;
};
__device__ __forceinline__ BnBwdRaw bn_bwd_fetch(const BnBwdIn& s, long long p, int c0) {
  BnBwdRaw r;
  r.d1 = bn_ld_raw8(s.da1 + p * s.ld1 + c0);
  if (s.da2) r.d2 = bn_ld_raw8(s.da2 + p * s.ld2 + c0);
  if (s.act) r.m = bn_ld_raw8(s.act + p * s.lda + c0);
  r.y = bn_ld_raw8(s.y + p * s.ldy + c0);
  return r;
}
// dz = (da1 [+ da2]) * [a > 0] and the raw pre-BN value y of one pixel's 8 channels.
__device__ __forceinline__ void bn_bwd_dz(const BnBwdIn& s, const F8& sc, const F8& sf, const BnBwdRaw& r,
                                          F8& dz, F8& y) {
  dz = bn_unpack8(r.d1);
  if (s.da2) {
    const F8 d2 = bn_unpack8(r.d2);
#pragma unroll
    for (int j = 0; j < 8; ++j) dz.v[j] += d2.v[j];
  }
  if (s.act) {
    const F8 m = bn_unpack8(r.m);
#pragma unroll
    for (int j = 0; j < 8; ++j) dz.v[j] = m.v[j] > 0.0f ? dz.v[j] : 0.0f;
  }
  y = bn_unpack8(r.y);
  if (s.self_mask) {  // the forward pass stored bf16(relu(fma(y, scale, shift))): same arithmetic here
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float act = __bfloat162float(__float2bfloat16_rn(fmaf(y.v[j], sc.v[j], sf.v[j])));
      dz.v[j] = act > 0.0f ? dz.v[j] : 0.0f;
    }
  }
}

// Block partials of (sum dz, sum dz*y).  The normalised form sum dz*yhat = invstd * (sum dz*y -
// mean * sum dz) is taken in fp64 by BN_BWD_FINALIZE (i5 = 1): the streaming loop then carries no
// per-channel mean / invstd (16 registers less, more loads in flight).
__global__ void __launch_bounds__(kEwThreads) bn_bwd_reduce_kernel(const EwArgs a) {
  pdl_prologue();
  extern __shared__ float red[];
  const BnBwdIn s = bn_bwd_in(a);
  // ew_channel_reduce gives every thread one fixed channel vector: constants loaded once
  const int c0t = (int)(threadIdx.x % (s.C >> 3)) * 8;
  F8 sc, sf;
  if (s.self_mask) {
    sc = ld_f8(s.bnp + c0t);
    sf = ld_f8(s.bnp + s.C + c0t);
  }
  ew_channel_reduce_batched<2, 4>(
      0, a.i[0], s.C, (float*)a.p[5], red, [&](long long p, int c0) { return bn_bwd_fetch(s, p, c0); },
      [&](const BnBwdRaw& r, int, float (&acc)[2][8]) {
        F8 dz, y;
        bn_bwd_dz(s, sc, sf, r, dz, y);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc[0][j] += dz.v[j];
          acc[1][j] = fmaf(dz.v[j], y.v[j], acc[1][j]);
        }
      },
      a.i[6] != 0);  // i6: fixed-point accumulators instead of block-partial rows
}

__global__ void __launch_bounds__(1024) bn_bwd_finalize_kernel(const EwArgs a) {
  pdl_prologue();
  const float* partial = (const float*)a.p[0];
  const float* gamma = (const float*)a.p[1];
  const float* bnp = (const float*)a.p[2];
  float* dgamma = (float*)a.p[3];
  float* dbeta = (float*)a.p[4];
  float* coef = (float*)a.p[5];
  float* dbias = (float*)a.p[6];
  const int rows = (int)a.i[0], C = (int)a.i[1];
  const double count = (double)a.i[2];
  const bool accumulate = a.i[3] != 0, training = a.i[4] != 0, raw_q = a.i[5] != 0;
  ew_rows_reduce<2>(partial, rows, 2 * (long long)C, C, C, [&](int c, const double* v) {
    const double s = v[0];
    // raw_q: the second accumulator is sum dz*y; sum dz*yhat = invstd * (sum dz*y - mean * sum dz)
    const double q = raw_q ? (double)bnp[3 * C + c] * (v[1] - (double)bnp[2 * C + c] * s) : v[1];
    const float g = gamma ? gamma[c] : 1.0f;
    const float c1 = g * bnp[3 * C + c];
    if (dgamma) dgamma[c] = accumulate ? dgamma[c] + (float)q : (float)q;
    if (dbeta) dbeta[c] = accumulate ? dbeta[c] + (float)s : (float)s;
    coef[c] = c1;
    coef[C + c] = training ? (float)(s / count) : 0.0f;
    coef[2 * C + c] = training ? (float)(q / count) : 0.0f;
    if (dbias) {  // gradient of a conv bias feeding this BN: sum of dy (0 with batch statistics)
      const float vb = training ? 0.0f : c1 * (float)s;
      dbias[c] = accumulate ? dbias[c] + vb : vb;
    }
  });
}

// dy = c1*(dz - c2 - yhat*c3) with yhat = (y - mean)*invstd, folded per channel into
// dy = A*dz + B*y + D:  A = c1,  B = -c1*c3*invstd,  D = c1*(c3*invstd*mean - c2).
struct BnBwdLin {
  F8 A, B, D, sc, sf;
};
__device__ __forceinline__ BnBwdLin bn_bwd_lin(const BnBwdIn& s, const float* coef, int c0) {
  BnBwdLin k;
  const int C = s.C;
  const F8 c1 = ld_f8(coef + c0), c2 = ld_f8(coef + C + c0), c3 = ld_f8(coef + 2 * C + c0);
  const F8 mean = ld_f8(s.bnp + 2 * C + c0), inv = ld_f8(s.bnp + 3 * C + c0);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float t = c3.v[j] * inv.v[j];
    k.A.v[j] = c1.v[j];
    k.B.v[j] = -c1.v[j] * t;
    k.D.v[j] = c1.v[j] * (t * mean.v[j] - c2.v[j]);
  }
  if (s.self_mask) {
    k.sc = ld_f8(s.bnp + c0);
    k.sf = ld_f8(s.bnp + C + c0);
  }
  return k;
}
// The same constants straight from the fixed-point sums (a.i[10] bit 0): the arithmetic of bn_bwd_finalize_kernel
// per channel, no finalize launch.  Block 0's first C/8 threads also write dgamma / dbeta / dbias.
//   p5 int64 [2][C] (sum dz, sum dz*y) x 2^40   p8 gamma|NULL  p9 dgamma|NULL  p10 dbeta|NULL  p11 dbias|NULL
//   i10: bit 0 fixed, bit 1 training, bit 2 accumulate into dgamma / dbeta / dbias
__device__ __forceinline__ BnBwdLin bn_bwd_lin_fixed(const BnBwdIn& s, const EwArgs& a, int c0, bool publish) {
  BnBwdLin k;
  const int C = s.C;
  const long long* acc = (const long long*)a.p[5];
  const float* gamma = (const float*)a.p[8];
  float* dgamma = (float*)a.p[9];
  float* dbeta = (float*)a.p[10];
  float* dbias = (float*)a.p[11];
  const double count = (double)a.i[0];
  const bool training = (a.i[10] & 2) != 0, accumulate = (a.i[10] & 4) != 0;
  long long r0[8], r1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    r0[j] = __ldcg(acc + c0 + j);
    r1[j] = __ldcg(acc + C + c0 + j);
  }
  const F8 mean = ld_f8(s.bnp + 2 * C + c0), inv = ld_f8(s.bnp + 3 * C + c0);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const double sd = (double)r0[j] * (1.0 / B2SEG_BWD_FIX_SCALE);
    const double q = (double)inv.v[j] * ((double)r1[j] * (1.0 / B2SEG_BWD_FIX_SCALE) - (double)mean.v[j] * sd);
    const float g = gamma ? gamma[c0 + j] : 1.0f;
    const float c1 = g * inv.v[j];
    const float c2 = training ? (float)(sd / count) : 0.0f;
    const float c3 = training ? (float)(q / count) : 0.0f;
    const float t = c3 * inv.v[j];
    k.A.v[j] = c1;
    k.B.v[j] = -c1 * t;
    k.D.v[j] = c1 * (t * mean.v[j] - c2);
    if (publish) {
      const int c = c0 + j;
      if (dgamma) dgamma[c] = accumulate ? dgamma[c] + (float)q : (float)q;
      if (dbeta) dbeta[c] = accumulate ? dbeta[c] + (float)sd : (float)sd;
      if (dbias) {
        const float vb = training ? 0.0f : c1 * (float)sd;
        dbias[c] = accumulate ? dbias[c] + vb : vb;
      }
    }
  }
  if (s.self_mask) {
    k.sc = ld_f8(s.bnp + c0);
    k.sf = ld_f8(s.bnp + C + c0);
  }
  return k;
}
__device__ __forceinline__ void bn_bwd_apply_store(const BnBwdIn& s, const BnBwdLin& k, const BnBwdRaw& r,
                                                   bool dz_acc, long long p, int c0, bf16* dy,
                                                   long long lddy, bf16* dzo, long long lddz) {
  F8 dz, y, o;
  bn_bwd_dz(s, k.sc, k.sf, r, dz, y);
#pragma unroll
  for (int j = 0; j < 8; ++j) o.v[j] = fmaf(k.A.v[j], dz.v[j], fmaf(k.B.v[j], y.v[j], k.D.v[j]));
  st_bf8(dy + p * lddy + c0, o);
  if (dzo) {
    if (dz_acc) {
      const F8 old = bn_unpack8(r.old);
#pragma unroll
      for (int j = 0; j < 8; ++j) dz.v[j] += old.v[j];
    }
    st_bf8(dzo + p * lddz + c0, dz);
  }
}

__global__ void __launch_bounds__(kEwThreads) bn_bwd_apply_kernel(const EwArgs a) {
  pdl_prologue();
  const BnBwdIn s = bn_bwd_in(a);
  const float* coef = (const float*)a.p[5];
  bf16* dy = (bf16*)a.p[6];
  bf16* dzo = (bf16*)a.p[7];
  const long long P = a.i[0], lddy = a.i[6], lddz = a.i[7];
  const bool dz_acc = dzo && (int)a.i[8] == 2;  // 1 = store, 2 = accumulate
  const int C = s.C, cv = C >> 3;
  if (kEwThreads % cv == 0) {
    const int ppb = kEwThreads / cv;
    const int c0 = (threadIdx.x % cv) * 8;
    const long long pstep = (long long)gridDim.x * ppb;
    const BnBwdLin k = (a.i[10] & 1) ? bn_bwd_lin_fixed(s, a, c0, blockIdx.x == 0 && threadIdx.x < cv)
                                     : bn_bwd_lin(s, coef, c0);
    long long p = (long long)blockIdx.x * ppb + threadIdx.x / cv;
    constexpr int U = 3;  // pixels in flight per thread
    for (; p + (U - 1) * pstep < P; p += U * pstep) {
      BnBwdRaw r[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        r[u] = bn_bwd_fetch(s, p + u * pstep, c0);
        if (dz_acc) r[u].old = bn_ld_raw8(dzo + (p + u * pstep) * lddz + c0);
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
        bn_bwd_apply_store(s, k, r[u], dz_acc, p + u * pstep, c0, dy, lddy, dzo, lddz);
    }
    for (; p < P; p += pstep) {
      BnBwdRaw r0 = bn_bwd_fetch(s, p, c0);
      if (dz_acc) r0.old = bn_ld_raw8(dzo + p * lddz + c0);
      bn_bwd_apply_store(s, k, r0, dz_acc, p, c0, dy, lddy, dzo, lddz);
    }
    return;
  }
  const long long total = P * cv;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    long long p;
    int c8;
    ew_divmod(idx, cv, p, c8);
    const int c0 = c8 * 8;
    const BnBwdLin k = bn_bwd_lin(s, coef, c0);
    BnBwdRaw r0 = bn_bwd_fetch(s, p, c0);
    if (dz_acc) r0.old = bn_ld_raw8(dzo + p * lddz + c0);
    bn_bwd_apply_store(s, k, r0, dz_acc, p, c0, dy, lddy, dzo, lddz);
  }
}

// ------------------------------------------------- fused (single-launch) BatchNorm ----
// Both fused kernels run 512-thread blocks, at most one (backward) or two (forward) per SM.
constexpr int kBnT = 512;

// Fold the rows of a block-partial buffer [rows][NACC][C] into sm_sum[NACC*C] (fp64), by the whole
// block, in an order fixed by (rows, C, blockDim) alone -> every block of every run computes
// bit-identical sums.  Column q = k*C + c; with fewer columns than threads the rows are dealt to
// blockDim/cols lanes and the lane sums are added in lane order.  Ends with __syncthreads().
template <int NACC>
__device__ __forceinline__ void bn_fold_rows(const float* partial, int rows, int C, double* sm_sum,
                                             double* sm_tmp /* blockDim doubles */) {
  constexpr int U = 32;  // loads in flight per thread and step: the fold is L2-latency bound
  const int cols = NACC * C, T = blockDim.x, t = threadIdx.x;
  if (cols >= T) {
    for (int q = t; q < cols; q += T) {
      double acc = 0.0;
      for (int r = 0; r < rows; r += U) {
        float v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = r + u < rows ? __ldcg(partial + (size_t)(r + u) * cols + q) : 0.0f;
#pragma unroll
        for (int u = 0; u < U; ++u) acc += (double)v[u];
      }
      sm_sum[q] = acc;
    }
    __syncthreads();
    return;
  }
  const int nl = T / cols, lane = t / cols, q = t - lane * cols;
  double acc = 0.0;
  if (lane < nl) {
    for (int r = lane; r < rows; r += U * nl) {
      float v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) v[u] = r + u * nl < rows ? __ldcg(partial + (size_t)(r + u * nl) * cols + q) : 0.0f;
#pragma unroll
      for (int u = 0; u < U; ++u) acc += (double)v[u];
    }
    sm_tmp[t] = acc;
  }
  __syncthreads();
  if (t < cols) {
    double s = 0.0;
    for (int l = 0; l < nl; ++l) s += sm_tmp[l * cols + t];
    sm_sum[t] = s;
  }
  __syncthreads();
}

// rows < 0: `partial` is the int64 [2][C] fixed-point accumulator the conv epilogues added to with atomics
// (b2seg_tapgemm_desc::stats_fixed): nothing to fold.
__device__ __forceinline__ void bn_fold_stats(const float* partial, int rows, int C, double* sm_sum, double* sm_tmp) {
  if (rows >= 0) {
    bn_fold_rows<2>(partial, rows, C, sm_sum, sm_tmp);
    return;
  }
  const long long* acc = reinterpret_cast<const long long*>(partial);
  for (int q = threadIdx.x; q < 2 * C; q += blockDim.x)
    sm_sum[q] = (double)__ldcg(acc + q) * (q < C ? 1.0 / (double)B2SEG_STATS_FIX_SUM : 1.0 / (double)B2SEG_STATS_FIX_SQ);
  __syncthreads();
}

// (sum, sumsq) -> scale / shift in shared memory; block 0 also publishes bnp = scale, shift, mean,
// invstd and updates the running statistics: the arithmetic of bn_finalize_kernel, verbatim.
__device__ __forceinline__ void bn_affine_channel(int c, double sum, double sumsq, int C, double count, float eps,
                                                  float mom, const float* gamma, const float* beta, float* rmean,
                                                  float* rvar, long long* nbt, float* bnp, float* sm_scale,
                                                  float* sm_shift) {
  const double m = sum / count;
  double var = sumsq / count - m * m;
  if (var < 0.0) var = 0.0;
  const float mean = (float)m;
  const float invstd = (float)(1.0 / sqrt(var + (double)eps));
  const float g = gamma ? gamma[c] : 1.0f, b = beta ? beta[c] : 0.0f;
  const float scale = g * invstd;
  const float shift = b - mean * scale;
  sm_scale[c] = scale;
  sm_shift[c] = shift;
  if (blockIdx.x == 0) {
    if (rmean) {
      const double unb = count > 1.0 ? var * count / (count - 1.0) : var;
      rmean[c] = (1.0f - mom) * rmean[c] + mom * mean;
      rvar[c] = (1.0f - mom) * rvar[c] + mom * (float)unb;
    }
    if (nbt && c == 0) *nbt += 1;
    bnp[c] = scale;
    bnp[C + c] = shift;
    bnp[2 * C + c] = mean;
    bnp[3 * C + c] = invstd;
  }
}
__device__ __forceinline__ void bn_affine_from_sums(const double* sm_sum, int C, double count, float eps,
                                                    float mom, const float* gamma, const float* beta,
                                                    float* rmean, float* rvar, long long* nbt, float* bnp,
                                                    float* sm_scale, float* sm_shift) {
  for (int c = threadIdx.x; c < C; c += blockDim.x)
    bn_affine_channel(c, sm_sum[c], sm_sum[C + c], C, count, eps, mom, gamma, beta, rmean, rvar, nbt, bnp,
                      sm_scale, sm_shift);
}
// ... straight from the int64 fixed-point accumulators (no fold, no fp64 staging)
__device__ __forceinline__ void bn_affine_from_fixed(const long long* acc, int C, double count, float eps,
                                                     float mom, const float* gamma, const float* beta,
                                                     float* rmean, float* rvar, long long* nbt, float* bnp,
                                                     float* sm_scale, float* sm_shift) {
  for (int c = threadIdx.x; c < C; c += blockDim.x)
    bn_affine_channel(c, (double)__ldcg(acc + c) * (1.0 / (double)B2SEG_STATS_FIX_SUM),
                      (double)__ldcg(acc + C + c) * (1.0 / (double)B2SEG_STATS_FIX_SQ), C, count, eps, mom, gamma,
                      beta, rmean, rvar, nbt, bnp, sm_scale, sm_shift);
}

// EW_BN_APPLY_FUSED with fixed-point statistics (rows < 0 for every BatchNorm involved): bn_apply_kernel's
// blocks, grid (one resident wave) and streaming loop; the first pixels of a thread are requested before the
// statistics are turned into scale / shift, so the accumulators' L2 round trip and the fp64 chain hide behind them.
// dynamic shared memory: [4C] floats (sc, sf, rs, rf)
__global__ void __launch_bounds__(kEwThreads, 4) bn_apply_fixed_kernel(const EwArgs a) {
  pdl_prologue();
  extern __shared__ __align__(16) float bn_aff_sm[];
  const bf16* y = (const bf16*)a.p[0];
  bf16* out = (bf16*)a.p[1];
  const bf16* res = (const bf16*)a.p[2];
  const long long P = a.i[0];
  const int C = (int)a.i[1];
  const long long ldy = a.i[2], ldo = a.i[3], ldr = a.i[4];
  const bool relu = a.i[5] != 0;
  const bool rnorm = a.p[10] != nullptr;
  const float eps = a.f[0], mom = a.f[1];
  const int cv = C >> 3;
  const bool fast = kEwThreads % cv == 0;
  constexpr int NB = 2;
  const uint4 zero4 = make_uint4(0, 0, 0, 0);
  long long p = 0, pstep = 0;
  int c0 = 0;
  uint4 yv[NB], rv[NB];
#pragma unroll
  for (int b = 0; b < NB; ++b) yv[b] = rv[b] = zero4;
  if (fast) {
    const int ppb = kEwThreads / cv;
    c0 = (threadIdx.x % cv) * 8;
    pstep = (long long)gridDim.x * ppb;
    p = (long long)blockIdx.x * ppb + threadIdx.x / cv;
#pragma unroll
    for (int b = 0; b < NB; ++b)
      if (p + b * pstep < P) {
        yv[b] = bn_ld_raw8(y + (p + b * pstep) * ldy + c0);
        if (res) rv[b] = bn_ld_raw8(res + (p + b * pstep) * ldr + c0);
      }
  }
  bn_affine_from_fixed((const long long*)a.p[3], C, (double)a.i[7], eps, mom, (const float*)a.p[4],
                       (const float*)a.p[5], (float*)a.p[6], (float*)a.p[7], (long long*)a.p[8], (float*)a.p[9],
                       bn_aff_sm, bn_aff_sm + C);
  if (rnorm)
    bn_affine_from_fixed((const long long*)a.p[10], C, (double)a.i[7], eps, mom, (const float*)a.p[11],
                         (const float*)a.p[12], (float*)a.p[13], (float*)a.p[14], (long long*)a.p[15],
                         (float*)a.p[16], bn_aff_sm + 2 * C, bn_aff_sm + 3 * C);
  __syncthreads();
  BnApplyCh ch;
  if (fast) {
    ch.sc = ld_f8(bn_aff_sm + c0);
    ch.sf = ld_f8(bn_aff_sm + C + c0);
    if (rnorm) {
      ch.rs = ld_f8(bn_aff_sm + 2 * C + c0);
      ch.rf = ld_f8(bn_aff_sm + 3 * C + c0);
    }
    // the group requested before the prologue
#pragma unroll
    for (int b = 0; b < NB; ++b)
      if (p + b * pstep < P)
        st_bf8(out + (p + b * pstep) * ldo + c0, bn_apply_one(yv[b], rv[b], res != nullptr, ch, rnorm, relu));
    p += NB * pstep;
    for (; p + (NB - 1) * pstep < P; p += NB * pstep) {
#pragma unroll
      for (int b = 0; b < NB; ++b) {
        yv[b] = bn_ld_raw8(y + (p + b * pstep) * ldy + c0);
        if (res) rv[b] = bn_ld_raw8(res + (p + b * pstep) * ldr + c0);
      }
#pragma unroll
      for (int b = 0; b < NB; ++b)
        st_bf8(out + (p + b * pstep) * ldo + c0, bn_apply_one(yv[b], rv[b], res != nullptr, ch, rnorm, relu));
    }
    for (; p < P; p += pstep) {
      const uint4 y0 = bn_ld_raw8(y + p * ldy + c0);
      uint4 r0 = zero4;
      if (res) r0 = bn_ld_raw8(res + p * ldr + c0);
      st_bf8(out + p * ldo + c0, bn_apply_one(y0, r0, res != nullptr, ch, rnorm, relu));
    }
    return;
  }
  const long long total = P * cv;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long q = idx / cv;
    const int d0 = (int)(idx % cv) * 8;
    ch.sc = ld_f8(bn_aff_sm + d0);
    ch.sf = ld_f8(bn_aff_sm + C + d0);
    const uint4 y0 = bn_ld_raw8(y + q * ldy + d0);
    uint4 r0 = zero4;
    if (res) {
      r0 = bn_ld_raw8(res + q * ldr + d0);
      if (rnorm) {
        ch.rs = ld_f8(bn_aff_sm + 2 * C + d0);
        ch.rf = ld_f8(bn_aff_sm + 3 * C + d0);
      }
    }
    st_bf8(out + q * ldo + d0, bn_apply_one(y0, r0, res != nullptr, ch, rnorm, relu));
  }
}

// dynamic shared memory: [2C] doubles (sums) | [blockDim] doubles (lanes) | [4C] floats (sc, sf, rs, rf)
__global__ void __launch_bounds__(kBnT, 2) bn_apply_fused_kernel(const EwArgs a) {
  pdl_prologue();
  extern __shared__ __align__(16) unsigned char bn_sm[];
  const bf16* y = (const bf16*)a.p[0];
  bf16* out = (bf16*)a.p[1];
  const bf16* res = (const bf16*)a.p[2];
  const long long P = a.i[0];
  const int C = (int)a.i[1];
  const long long ldy = a.i[2], ldo = a.i[3], ldr = a.i[4];
  const bool relu = a.i[5] != 0;
  const bool rnorm = a.p[10] != nullptr;
  double* sm_sum = reinterpret_cast<double*>(bn_sm);
  double* sm_tmp = sm_sum + 2 * C;
  float* sm_aff = reinterpret_cast<float*>(sm_tmp + kBnT);
  const float eps = a.f[0], mom = a.f[1];
  bn_fold_stats((const float*)a.p[3], (int)a.i[6], C, sm_sum, sm_tmp);
  bn_affine_from_sums(sm_sum, C, (double)a.i[7], eps, mom, (const float*)a.p[4], (const float*)a.p[5],
                      (float*)a.p[6], (float*)a.p[7], (long long*)a.p[8], (float*)a.p[9], sm_aff, sm_aff + C);
  __syncthreads();
  if (rnorm) {
    bn_fold_stats((const float*)a.p[10], (int)a.i[8], C, sm_sum, sm_tmp);
    bn_affine_from_sums(sm_sum, C, (double)a.i[7], eps, mom, (const float*)a.p[11], (const float*)a.p[12],
                        (float*)a.p[13], (float*)a.p[14], (long long*)a.p[15], (float*)a.p[16],
                        sm_aff + 2 * C, sm_aff + 3 * C);
    __syncthreads();
  }
  const int cv = C >> 3;
  BnApplyCh ch;
  if (kBnT % cv == 0) {
    const int ppb = kBnT / cv;
    const int c0 = (threadIdx.x % cv) * 8;
    const long long pstep = (long long)gridDim.x * ppb;
    ch.sc = ld_f8(sm_aff + c0);
    ch.sf = ld_f8(sm_aff + C + c0);
    if (rnorm) {
      ch.rs = ld_f8(sm_aff + 2 * C + c0);
      ch.rf = ld_f8(sm_aff + 3 * C + c0);
    }
    long long p = (long long)blockIdx.x * ppb + threadIdx.x / cv;
    for (; p + pstep < P; p += 2 * pstep) {
      const long long q = p + pstep;
      const uint4 y0 = bn_ld_raw8(y + p * ldy + c0), y1 = bn_ld_raw8(y + q * ldy + c0);
      uint4 r0 = make_uint4(0, 0, 0, 0), r1 = r0;
      if (res) {
        r0 = bn_ld_raw8(res + p * ldr + c0);
        r1 = bn_ld_raw8(res + q * ldr + c0);
      }
      st_bf8(out + p * ldo + c0, bn_apply_one(y0, r0, res != nullptr, ch, rnorm, relu));
      st_bf8(out + q * ldo + c0, bn_apply_one(y1, r1, res != nullptr, ch, rnorm, relu));
    }
    if (p < P) {
      const uint4 y0 = bn_ld_raw8(y + p * ldy + c0);
      uint4 r0 = make_uint4(0, 0, 0, 0);
      if (res) r0 = bn_ld_raw8(res + p * ldr + c0);
      st_bf8(out + p * ldo + c0, bn_apply_one(y0, r0, res != nullptr, ch, rnorm, relu));
    }
    return;
  }
  const long long total = P * cv;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    long long p;
    int c8;
    ew_divmod(idx, cv, p, c8);
    const int c0 = c8 * 8;
    ch.sc = ld_f8(sm_aff + c0);
    ch.sf = ld_f8(sm_aff + C + c0);
    const uint4 y0 = bn_ld_raw8(y + p * ldy + c0);
    uint4 r0 = make_uint4(0, 0, 0, 0);
    if (res) {
      r0 = bn_ld_raw8(res + p * ldr + c0);
      if (rnorm) {
        ch.rs = ld_f8(sm_aff + 2 * C + c0);
        ch.rf = ld_f8(sm_aff + 3 * C + c0);
      }
    }
    st_bf8(out + p * ldo + c0, bn_apply_one(y0, r0, res != nullptr, ch, rnorm, relu));
  }
}

// Device-wide barrier of a grid whose blocks are all co-resident (<= one per SM).  Sense-reversing
// on a generation word, so it needs no host reset between launches or graph replays.  Bounded: a
// scheduling surprise traps instead of hanging the GPU.
static __device__ unsigned g_bn_bar_count = 0;
static __device__ unsigned g_bn_bar_gen = 0;
__device__ __forceinline__ void bn_grid_barrier() {
  __syncthreads();
  if (threadIdx.x == 0) {
    volatile unsigned* genp = &g_bn_bar_gen;
    const unsigned gen = *genp;  // read BEFORE arriving: the generation cannot advance without us
    __threadfence();
    if (atomicAdd(&g_bn_bar_count, 1u) == gridDim.x - 1) {
      g_bn_bar_count = 0;
      __threadfence();
      atomicAdd(&g_bn_bar_gen, 1u);
    } else {
      const long long t0 = clock64();
      while (*genp == gen) {
        __nanosleep(64);
        if (clock64() - t0 > 8000000000LL) {  // ~4 s
          printf("b2seg: bn grid barrier timeout (block %d of %d)\n", blockIdx.x, gridDim.x);
          __trap();
        }
      }
    }
    __threadfence();
  }
  __syncthreads();
}

// dynamic shared memory: [2][pl][C] floats (block reduction; reused as [2C] doubles + [blockDim]
// doubles for the fold) | [3C] floats (A, B, D).
// 256 threads, <= 128 registers, <= ~22 KB of shared memory and ONE block per SM: the block fits
// beside a weight-gradient CTA of the side branch (192 threads, ~50 registers, 197 KB), so the
// barrier never has to wait for the side branch to drain.
constexpr int kBnB = 256;
struct BnBwdLean {
  uint4 d1, m, y, old;
};
template <bool HAS_ACT>
__device__ __forceinline__ BnBwdLean bn_lean_fetch(const BnBwdIn& s, long long p, int c0) {
  BnBwdLean r;
  r.d1 = bn_ld_raw8(s.da1 + p * s.ld1 + c0);
  if (HAS_ACT) r.m = bn_ld_raw8(s.act + p * s.lda + c0);
  r.y = bn_ld_raw8(s.y + p * s.ldy + c0);
  return r;
}
template <bool HAS_ACT>
__device__ __forceinline__ void bn_lean_dz(const BnBwdIn& s, const F8& sc, const F8& sf, const BnBwdLean& r,
                                           F8& dz, F8& y) {
  dz = bn_unpack8(r.d1);
  if (HAS_ACT) {
    const F8 m = bn_unpack8(r.m);
#pragma unroll
    for (int j = 0; j < 8; ++j) dz.v[j] = m.v[j] > 0.0f ? dz.v[j] : 0.0f;
  }
  y = bn_unpack8(r.y);
  if (!HAS_ACT && s.self_mask) {  // the forward pass stored bf16(relu(fma(y, scale, shift))): same arithmetic
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float act = __bfloat162float(__float2bfloat16_rn(fmaf(y.v[j], sc.v[j], sf.v[j])));
      dz.v[j] = act > 0.0f ? dz.v[j] : 0.0f;
    }
  }
}

template <bool HAS_ACT>
__global__ void __launch_bounds__(kBnB, 1) bn_bwd_fused_kernel(const EwArgs a) {
  pdl_prologue();
  extern __shared__ __align__(16) unsigned char bn_sm[];
  BnBwdIn s = bn_bwd_in(a);
  s.self_mask = (int)a.i[9];
  float* partial = (float*)a.p[5];
  bf16* dy = (bf16*)a.p[6];
  bf16* dzo = (bf16*)a.p[7];
  const long long P = a.i[0], lddy = a.i[6], lddz = a.i[7];
  const bool dz_acc = dzo && (int)a.i[8] == 2;
  const bool training = a.i[10] != 0, accumulate = a.i[11] != 0;
  const int C = s.C, cv = C >> 3;
  const int pl = kBnB / cv > 0 ? kBnB / cv : 1;
  const int lc = threadIdx.x % cv, lp = threadIdx.x / cv;
  const int c0 = lc * 8;
  const bool active = lp < pl;
  const long long pstep = (long long)gridDim.x * pl;
  const long long pfirst = (long long)blockIdx.x * pl + lp;
  float* red = reinterpret_cast<float*>(bn_sm);
  const size_t red_bytes = (size_t)2 * pl * C * sizeof(float);
  const size_t fold_bytes = (size_t)(2 * C + kBnB) * sizeof(double);
  float* sm_k = reinterpret_cast<float*>(bn_sm + (red_bytes > fold_bytes ? red_bytes : fold_bytes));

  // ---- phase 1: block partials of (sum dz, sum dz*y), four pixels in flight
  F8 sc, sf;
  if (!HAS_ACT && s.self_mask) {
    sc = ld_f8(s.bnp + c0);
    sf = ld_f8(s.bnp + C + c0);
  }
  float acc[2][8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[0][j] = acc[1][j] = 0.0f;
  long long nmine = 0;  // pixels this thread owns: pfirst + k*pstep, k < nmine
  if (active && pfirst < P) nmine = (P - pfirst + pstep - 1) / pstep;
  {
    constexpr int U1 = 4;
    long long k = 0;
    for (; k + U1 <= nmine; k += U1) {
      BnBwdLean r[U1];
#pragma unroll
      for (int u = 0; u < U1; ++u) r[u] = bn_lean_fetch<HAS_ACT>(s, pfirst + (k + u) * pstep, c0);
#pragma unroll
      for (int u = 0; u < U1; ++u) {
        F8 dz, yv;
        bn_lean_dz<HAS_ACT>(s, sc, sf, r[u], dz, yv);
#pragma unroll
        for (int j = 0; j < 8; ++j) { acc[0][j] += dz.v[j]; acc[1][j] = fmaf(dz.v[j], yv.v[j], acc[1][j]); }
      }
    }
    for (; k < nmine; ++k) {
      F8 dz, yv;
      bn_lean_dz<HAS_ACT>(s, sc, sf, bn_lean_fetch<HAS_ACT>(s, pfirst + k * pstep, c0), dz, yv);
#pragma unroll
      for (int j = 0; j < 8; ++j) { acc[0][j] += dz.v[j]; acc[1][j] = fmaf(dz.v[j], yv.v[j], acc[1][j]); }
    }
  }
  if (active) {
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
      for (int j = 0; j < 8; ++j) red[((size_t)k * pl + lp) * C + c0 + j] = acc[k][j];
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < 2 * C; idx += blockDim.x) {
    const int k = idx / C, c = idx % C;
    float v = 0.0f;
    for (int l = 0; l < pl; ++l) v += red[((size_t)k * pl + l) * C + c];
    partial[((size_t)blockIdx.x * 2 + k) * C + c] = v;
  }
  bn_grid_barrier();

  // ---- distributed fold: block b owns channels [b*nch, (b+1)*nch); one warp per (channel, sum)
  // pair adds that column over the gridDim.x rows (lane l takes rows l, l+32, ...; lanes combined by
  // a fixed shuffle tree -> bit-reproducible), then the per-channel constants of BN_BWD_FINALIZE +
  // bn_bwd_lin go to a [3][C] table behind the partial rows; a second barrier publishes the table.
  {
    const int G = (int)gridDim.x;
    const int nch = (C + G - 1) / G;
    const int cbase = (int)blockIdx.x * nch;
    double* sm_pair = reinterpret_cast<double*>(bn_sm);  // [nch][2]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int pair = warp; pair < 2 * nch; pair += kBnB / 32) {
      const int c = cbase + (pair >> 1), k = pair & 1;
      double acc = 0.0;
      if (c < C) {
        float v[5];
#pragma unroll
        for (int u = 0; u < 5; ++u) {
          const int r = lane + 32 * u;
          v[u] = r < G ? __ldcg(partial + ((size_t)r * 2 + k) * C + c) : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < 5; ++u) acc += (double)v[u];
        for (int r = lane + 160; r < G; r += 32) acc += (double)__ldcg(partial + ((size_t)r * 2 + k) * C + c);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) sm_pair[pair] = acc;
    }
    __syncthreads();
    float* coef = partial + (size_t)G * 2 * C;
    const float* gamma = (const float*)a.p[8];
    float* dgamma = (float*)a.p[9];
    float* dbeta = (float*)a.p[10];
    float* dbias = (float*)a.p[11];
    const double count = (double)P;
    for (int j = threadIdx.x; j < nch; j += blockDim.x) {
      const int c = cbase + j;
      if (c >= C) break;
      const double sd = sm_pair[2 * j], sq = sm_pair[2 * j + 1];
      const float mean = s.bnp[2 * C + c], inv = s.bnp[3 * C + c];
      const double q = (double)inv * (sq - (double)mean * sd);
      const float g = gamma ? gamma[c] : 1.0f;
      const float c1 = g * inv;
      const float c2 = training ? (float)(sd / count) : 0.0f;
      const float c3 = training ? (float)(q / count) : 0.0f;
      const float t = c3 * inv;
      coef[c] = c1;
      coef[C + c] = -c1 * t;
      coef[2 * C + c] = c1 * (t * mean - c2);
      if (dgamma) dgamma[c] = accumulate ? dgamma[c] + (float)q : (float)q;
      if (dbeta) dbeta[c] = accumulate ? dbeta[c] + (float)sd : (float)sd;
      if (dbias) {
        const float vb = training ? 0.0f : c1 * (float)sd;
        dbias[c] = accumulate ? dbias[c] + vb : vb;
      }
    }
    bn_grid_barrier();
    for (int e = threadIdx.x; e < 3 * C; e += blockDim.x) sm_k[e] = __ldcg(coef + e);
  }
  __syncthreads();

  // ---- phase 2: dy = A*dz + B*y + D over the same pixels, last read first (L2-hot)
  if (!active) return;
  const F8 kA = ld_f8(sm_k + c0), kB = ld_f8(sm_k + C + c0), kD = ld_f8(sm_k + 2 * C + c0);
  auto emit = [&](const BnBwdLean& r, long long p) {
    F8 dz, yv, o;
    bn_lean_dz<HAS_ACT>(s, sc, sf, r, dz, yv);
#pragma unroll
    for (int j = 0; j < 8; ++j) o.v[j] = fmaf(kA.v[j], dz.v[j], fmaf(kB.v[j], yv.v[j], kD.v[j]));
    st_bf8(dy + p * lddy + c0, o);
    if (dzo) {
      if (dz_acc) {
        const F8 old = bn_unpack8(r.old);
#pragma unroll
        for (int j = 0; j < 8; ++j) dz.v[j] += old.v[j];
      }
      st_bf8(dzo + p * lddz + c0, dz);
    }
  };
  constexpr int U = 4;
  long long k = nmine;
  for (; k >= U; k -= U) {
    BnBwdLean r[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long p = pfirst + (k - 1 - u) * pstep;
      r[u] = bn_lean_fetch<HAS_ACT>(s, p, c0);
      if (dz_acc) r[u].old = bn_ld_raw8(dzo + p * lddz + c0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) emit(r[u], pfirst + (k - 1 - u) * pstep);
  }
  for (; k >= 1; --k) {
    const long long p = pfirst + (k - 1) * pstep;
    BnBwdLean r0 = bn_lean_fetch<HAS_ACT>(s, p, c0);
    if (dz_acc) r0.old = bn_ld_raw8(dzo + p * lddz + c0);
    emit(r0, p);
  }
}

// ---------------------------------------------------------------------- COLSUM ----
__global__ void __launch_bounds__(kEwThreads) colsum_kernel(const EwArgs a) {
  pdl_prologue();
  extern __shared__ float red[];
  const bf16* x = (const bf16*)a.p[0];
  const long long ld = a.i[2];
  ew_channel_reduce<1>(0, a.i[0], (int)a.i[1], (float*)a.p[1], red,
                       [&](long long p, int c0, float (&acc)[1][8]) {
                         const F8 v = ld_bf8(x + p * ld + c0);
#pragma unroll
                         for (int j = 0; j < 8; ++j) acc[0][j] += v.v[j];
                       });
}

// out[k] (+)= sum_r partial[r*stride + k]; grid (K/32, slices), block (32 k, 32 row lanes)
__global__ void __launch_bounds__(1024) colsum_finalize_kernel(const EwArgs a) {
  pdl_prologue();
  const float* partial = (const float*)a.p[0];
  float* out = (float*)a.p[1];
  const int rows = (int)a.i[0], K = (int)a.i[1];
  const bool accumulate = a.i[2] != 0;
  ew_rows_reduce<1>(partial, rows, a.i[3], 0, K, [&](int k, const double* v) {
    out[k] = accumulate ? out[k] + (float)v[0] : (float)v[0];
  });
}

// ------------------------------------------------------------------------ PACK ----
// One launch repacks every fp32 parameter into its bf16 GEMM operand layouts:
// out[t][a][b] = bf16(in[a*sa + b*sb + t*st]).  One block per 32(a) x 32(b) x T tile (job.start =
// first tile of the job): the tile is read along whichever of a / b is contiguous with the taps
// in the parameter (coalesced), transposed through shared memory and written along b.
// Every layout the plan asks for has st == 1 and the taps innermost (inner stride == T), so a
// tile is 32 contiguous runs of 32*T floats; for T = 1, 4, 9 (1x1, 2x2, 3x3 kernels) the index
// arithmetic is compile-time and the store packs two bf16 per thread.
constexpr int kPackT = 9;
constexpr int kPackTS = 32 * 33 + 4;  // tap stride in shared memory: bank = 4t + a + b

template <int T>
__device__ __forceinline__ void pack_tile_fast(const b2seg_pack_job& j, int a0, int b0, float* sm) {
  constexpr int RUN = 32 * T;
  const bool inner_b = j.sb < j.sa;
  const long long so = inner_b ? j.sa : j.sb;
  const float* __restrict__ in = j.in + (long long)a0 * j.sa + (long long)b0 * j.sb;
  const int outer_n = inner_b ? j.A - a0 : j.B - b0;
  const int run_n = (inner_b ? j.B - b0 : j.A - a0) * T;  // valid floats of a run (may exceed RUN)
#pragma unroll 4
  for (int e = threadIdx.x; e < 32 * RUN; e += kEwThreads) {
    const int u = e / RUN, r = e % RUN, v = r / T, t = r % T;
    const float val = (u < outer_n && r < run_n) ? __ldg(in + (long long)u * so + r) : 0.0f;
    const int la = inner_b ? u : v, lb = inner_b ? v : u;
    sm[t * kPackTS + la * 33 + lb] = val;
  }
  __syncthreads();
  bf16* __restrict__ out = (bf16*)j.out;
  const bool pairs = (j.B & 1) == 0 && (reinterpret_cast<uintptr_t>(out) & 3) == 0;
#pragma unroll 2
  for (int e = threadIdx.x; e < T * 512; e += kEwThreads) {
    const int lb = (e & 15) * 2, la = (e >> 4) & 31, t = e >> 9;
    if (a0 + la >= j.A || b0 + lb >= j.B) continue;
    const float* src = sm + t * kPackTS + la * 33 + lb;
    bf16* dst = out + ((long long)t * j.A + a0 + la) * j.B + b0 + lb;
    if (pairs) {  // B even and b0 + lb even: lb + 1 is in range and dst is 4-byte aligned
      *reinterpret_cast<uint32_t*>(dst) = f2_to_bf2(src[0], src[1]);
    } else {
      dst[0] = __float2bfloat16_rn(src[0]);
      if (b0 + lb + 1 < j.B) dst[1] = __float2bfloat16_rn(src[1]);
    }
  }
}

__device__ __forceinline__ void pack_tile_generic(const b2seg_pack_job& j, int a0, int b0, float* sm) {
  const bool inner_b = j.sb < j.sa;
  const int T = j.T, runlen = 32 * T;
  for (int e = threadIdx.x; e < 32 * runlen; e += blockDim.x) {
    const int u = e / runlen, r = e % runlen, v = r / T, t = r % T;
    const int la = inner_b ? u : v, lb = inner_b ? v : u;
    float val = 0.0f;
    if (a0 + la < j.A && b0 + lb < j.B)
      val = j.in[(long long)(a0 + la) * j.sa + (long long)(b0 + lb) * j.sb + (long long)t * j.st];
    sm[t * kPackTS + la * 33 + lb] = val;
  }
  __syncthreads();
  bf16* out = (bf16*)j.out;
  for (int e = threadIdx.x; e < T * 1024; e += blockDim.x) {
    const int lb = e & 31, la = (e >> 5) & 31, t = e >> 10;
    if (a0 + la < j.A && b0 + lb < j.B)
      out[((long long)t * j.A + a0 + la) * j.B + b0 + lb] = __float2bfloat16_rn(sm[t * kPackTS + la * 33 + lb]);
  }
}

__global__ void __launch_bounds__(kEwThreads) pack_kernel(const EwArgs a) {
  pdl_prologue();
  __shared__ float sm[kPackT * kPackTS];
  const b2seg_pack_job* jobs = (const b2seg_pack_job*)a.p[0];
  const int njobs = (int)a.i[0];
  const long long tile = blockIdx.x;
  int lo = 0, hi = njobs - 1;
  while (lo < hi) {  // last job with start <= tile
    const int mid = (lo + hi + 1) >> 1;
    if (jobs[mid].start <= tile) lo = mid; else hi = mid - 1;
  }
  const b2seg_pack_job j = jobs[lo];
  const int tilesB = (j.B + 31) / 32;
  const int lt = (int)(tile - j.start);
  const int a0 = (lt / tilesB) * 32, b0 = (lt % tilesB) * 32;
  const bool fast = j.st == 1 && (j.sb < j.sa ? j.sb : j.sa) == j.T;
  if (fast && j.T == 9) pack_tile_fast<9>(j, a0, b0, sm);
  else if (fast && j.T == 4) pack_tile_fast<4>(j, a0, b0, sm);
  else if (fast && j.T == 1) pack_tile_fast<1>(j, a0, b0, sm);
  else pack_tile_generic(j, a0, b0, sm);
}

__global__ void __launch_bounds__(kEwThreads) fill_kernel(const EwArgs a) {
  pdl_prologue();
  uint32_t* dst = (uint32_t*)a.p[0];
  const long long n = a.i[0] / 4;
  const uint32_t pat = (uint32_t)a.i[1];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    dst[i] = pat;
}

#define EW_LAUNCHER(name, kernel)                                                         \
  int name(const EwArgs& a, dim3 grid, dim3 block, size_t smem, cudaStream_t st) {        \
    B2_CHECK_CUDA(b2_launch(kernel, grid, block, smem, st, a));                                                 \
    B2_CHECK_CUDA(cudaGetLastError());                                                    \
    return B2SEG_OK;                                                                      \
  }
EW_LAUNCHER(launch_bn_finalize, bn_finalize_kernel)
EW_LAUNCHER(launch_bn_apply, bn_apply_kernel)
EW_LAUNCHER(launch_bn_bwd_reduce, bn_bwd_reduce_kernel)
EW_LAUNCHER(launch_bn_bwd_finalize, bn_bwd_finalize_kernel)
EW_LAUNCHER(launch_bn_bwd_apply, bn_bwd_apply_kernel)
EW_LAUNCHER(launch_colsum, colsum_kernel)
EW_LAUNCHER(launch_colsum_finalize, colsum_finalize_kernel)
EW_LAUNCHER(launch_pack, pack_kernel)
EW_LAUNCHER(launch_fill, fill_kernel)
EW_LAUNCHER(launch_bn_apply_fused, bn_apply_fused_kernel)
EW_LAUNCHER(launch_bn_apply_fixed, bn_apply_fixed_kernel)
EW_LAUNCHER(launch_bn_bwd_fused_act, bn_bwd_fused_kernel<true>)
EW_LAUNCHER(launch_bn_bwd_fused_noact, bn_bwd_fused_kernel<false>)

constexpr size_t kBnFusedSmemMax = 96 * 1024;
int bn_fused_attrs() {
  static bool done = false;
  if (!done) {
    B2_CHECK_CUDA(cudaFuncSetAttribute(bn_apply_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kBnFusedSmemMax));
    B2_CHECK_CUDA(cudaFuncSetAttribute(bn_bwd_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kBnFusedSmemMax));
    B2_CHECK_CUDA(cudaFuncSetAttribute(bn_bwd_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kBnFusedSmemMax));
    done = true;
  }
  return B2SEG_OK;
}

size_t bn_bwd_fused_smem(int C) {
  const int cv = C / 8;
  const int pl = kBnB / cv > 0 ? kBnB / cv : 1;
  const size_t red = (size_t)2 * pl * C * sizeof(float), fold = (size_t)(2 * C + kBnB) * sizeof(double);
  return (red > fold ? red : fold) + (size_t)3 * C * sizeof(float);
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace

#define EW_REQ_VEC(C, ...)                                                                \
  B2_REQUIRE((C) >= 8 && (C) % 8 == 0 && (C) <= 2048, "ew: C=%d must be a multiple of 8 in [8,2048]", (int)(C))

int ew_plan_bn(const b2seg_ew_desc* d, EwPlan* pl) {
  const int64_t* i = d->i;
  void* const* p = d->p;
  pl->smem = 0;
  pl->launches = 1;
  switch (d->kind) {
    case B2SEG_EW_BN_FINALIZE: {
      B2_REQUIRE(p[6] && i[1] >= 1 && i[2] >= 1, "bn_finalize: bad arguments");
      B2_REQUIRE(!i[3] || (p[0] && i[0] >= 1), "bn_finalize: training needs partial sums");
      B2_REQUIRE(i[3] || (p[3] && p[4]), "bn_finalize: eval needs running statistics");
      B2_REQUIRE((i[1] + 31) / 32 <= kFinMaxColBlocks, "bn_finalize: too many channels");
      pl->fn = launch_bn_finalize;
      pl->grid = dim3((unsigned)((i[1] + 31) / 32), i[3] ? ew_fin_slices(i[0], (int)((i[1] + 31) / 32)) : 1);
      pl->block = dim3(32, 32);
      return B2SEG_OK;
    }
    case B2SEG_EW_BN_APPLY: {
      EW_REQ_VEC(i[1]);
      B2_REQUIRE(p[0] && p[1] && p[2] && i[0] >= 1, "bn_apply: null argument");
      B2_REQUIRE(i[2] % 8 == 0 && i[3] % 8 == 0 && i[4] % 8 == 0 && aligned16(p[0]) &&
                     aligned16(p[2]) && aligned16(p[3]),
                 "bn_apply: leading dimensions / pointers must be 16-byte aligned");
      pl->fn = launch_bn_apply;
      pl->grid = dim3(ew_resident_blocks(bn_apply_kernel, i[0] * (i[1] / 8)));
      pl->block = dim3(kEwThreads);
      return B2SEG_OK;
    }
    case B2SEG_EW_BN_BWD_REDUCE:
    case B2SEG_EW_BN_BWD_APPLY: {
      EW_REQ_VEC(i[1]);
      B2_REQUIRE(p[0] && p[3] && p[4] && p[5] && i[0] >= 1, "bn_bwd: null argument");
      B2_REQUIRE(i[2] % 8 == 0 && i[3] % 8 == 0 && i[4] % 8 == 0 && i[5] % 8 == 0 &&
                     aligned16(p[0]) && aligned16(p[1]) && aligned16(p[2]) && aligned16(p[3]),
                 "bn_bwd: leading dimensions / pointers must be 16-byte aligned");
      if (d->kind == B2SEG_EW_BN_BWD_REDUCE) {
        pl->fn = launch_bn_bwd_reduce;
        pl->grid = dim3(ew_rows(i[0], (int)i[1], kBnBwdReduceBlocksPerSm));
        pl->block = dim3(kEwThreads);
        pl->smem = (size_t)2 * ew_pl((int)i[1]) * i[1] * sizeof(float);
      } else {
        B2_REQUIRE(p[6] && i[6] % 8 == 0 && i[7] % 8 == 0 && aligned16(p[6]) && aligned16(p[7]),
                   "bn_bwd_apply: dy/dz must be 16-byte aligned");
        B2_REQUIRE(!p[7] || i[8] == 1 || i[8] == 2, "bn_bwd_apply: dz mode must be 1 or 2");
        B2_REQUIRE(!(i[10] & 1) || kEwThreads % (i[1] / 8) == 0,
                   "bn_bwd_apply: fixed-point sums need C/8 to divide %d (C = %lld)", kEwThreads, i[1]);
        pl->fn = launch_bn_bwd_apply;
        pl->grid = dim3(ew_resident_blocks(bn_bwd_apply_kernel, i[0] * (i[1] / 8)));
        pl->block = dim3(kEwThreads);
      }
      return B2SEG_OK;
    }
    case B2SEG_EW_BN_BWD_FINALIZE: {
      B2_REQUIRE(p[0] && p[2] && p[5] && i[0] >= 1 && i[1] >= 1 && i[2] >= 1,
                 "bn_bwd_finalize: bad arguments");
      B2_REQUIRE((i[1] + 31) / 32 <= kFinMaxColBlocks, "bn_bwd_finalize: too many channels");
      pl->fn = launch_bn_bwd_finalize;
      pl->grid = dim3((unsigned)((i[1] + 31) / 32), ew_fin_slices(i[0], (int)((i[1] + 31) / 32)));
      pl->block = dim3(32, 32);
      return B2SEG_OK;
    }
    case B2SEG_EW_COLSUM: {
      EW_REQ_VEC(i[1]);
      B2_REQUIRE(p[0] && p[1] && i[0] >= 1 && i[2] % 8 == 0 && aligned16(p[0]), "colsum: bad arguments");
      pl->fn = launch_colsum;
      pl->grid = dim3(ew_rows(i[0], (int)i[1]));
      pl->block = dim3(kEwThreads);
      pl->smem = (size_t)ew_pl((int)i[1]) * i[1] * sizeof(float);
      return B2SEG_OK;
    }
    case B2SEG_EW_COLSUM_FINALIZE: {
      B2_REQUIRE(p[0] && p[1] && i[0] >= 1 && i[1] >= 1 && i[3] >= i[1], "colsum_finalize: bad arguments");
      B2_REQUIRE((i[1] + 31) / 32 <= kFinMaxColBlocks, "colsum_finalize: too many columns");
      pl->fn = launch_colsum_finalize;
      pl->grid = dim3((unsigned)((i[1] + 31) / 32), ew_fin_slices(i[0], (int)((i[1] + 31) / 32)));
      pl->block = dim3(32, 32);
      return B2SEG_OK;
    }
    case B2SEG_EW_PACK: {
      B2_REQUIRE(p[0] && i[0] >= 1 && i[1] >= 1, "pack: bad arguments");
      pl->fn = launch_pack;
      pl->grid = dim3((unsigned)i[1]);  /* one block per 32 x 32 x T tile */
      pl->block = dim3(kEwThreads);
      return B2SEG_OK;
    }
    case B2SEG_EW_BN_APPLY_FUSED: {
      EW_REQ_VEC(i[1]);
      B2_REQUIRE(p[0] && p[1] && p[3] && p[9] && i[0] >= 1 && i[6] != 0 && i[7] >= 1, "bn_apply_fused: null argument");
      B2_REQUIRE(!p[10] || (p[2] && p[16] && i[8] != 0), "bn_apply_fused: residual BatchNorm needs res, bnp and rows");
      B2_REQUIRE(i[2] % 8 == 0 && i[3] % 8 == 0 && i[4] % 8 == 0 && aligned16(p[0]) &&
                     aligned16(p[1]) && aligned16(p[2]),
                 "bn_apply_fused: leading dimensions / pointers must be 16-byte aligned");
      if (i[6] < 0 && (!p[10] || i[8] < 0) && !getenv("B2SEG_BN_FIXED_WIDE")) {  // fixed-point statistics only
        pl->fn = launch_bn_apply_fixed;
        pl->grid = dim3(ew_resident_blocks(bn_apply_fixed_kernel, i[0] * (i[1] / 8)));
        pl->block = dim3(kEwThreads);
        pl->smem = (size_t)4 * i[1] * sizeof(float);
        pl->smem_limit = 48 * 1024;
        B2_REQUIRE(pl->smem <= pl->smem_limit, "bn_apply_fused: %lld channels exceed the shared-memory table", i[1]);
        return B2SEG_OK;
      }
      pl->fn = launch_bn_apply_fused;
      long long b = (i[0] * (i[1] / 8) + kBnT * 4 - 1) / (kBnT * 4);  // >= 4 vectors per thread
      const long long cap = 2ll * b2_num_sms();
      pl->grid = dim3((unsigned)(b > cap ? cap : (b < 1 ? 1 : b)));
      pl->block = dim3(kBnT);
      pl->smem = (size_t)(2 * i[1] + kBnT) * sizeof(double) + (size_t)4 * i[1] * sizeof(float);
      pl->smem_limit = kBnFusedSmemMax;
      return bn_fused_attrs();
    }
    case B2SEG_EW_BN_BWD_FUSED: {
      EW_REQ_VEC(i[1]);
      B2_REQUIRE(p[0] && p[3] && p[4] && p[5] && p[6] && i[0] >= 1, "bn_bwd_fused: null argument");
      B2_REQUIRE(!p[1], "bn_bwd_fused: a second upstream gradient (da2) takes the three-kernel sequence");
      B2_REQUIRE(i[2] % 8 == 0 && i[3] % 8 == 0 && i[4] % 8 == 0 && i[5] % 8 == 0 && i[6] % 8 == 0 &&
                     i[7] % 8 == 0 && aligned16(p[0]) && aligned16(p[1]) && aligned16(p[2]) &&
                     aligned16(p[3]) && aligned16(p[6]) && aligned16(p[7]),
                 "bn_bwd_fused: leading dimensions / pointers must be 16-byte aligned");
      B2_REQUIRE(!p[7] || i[8] == 1 || i[8] == 2, "bn_bwd_fused: dz mode must be 1 or 2");
      pl->fn = p[2] ? launch_bn_bwd_fused_act : launch_bn_bwd_fused_noact;
      const int cvv = (int)i[1] / 8;
      const int plv = kBnB / cvv > 0 ? kBnB / cvv : 1;
      long long b = (i[0] + plv - 1) / plv;  // at least one pixel per lane
      const long long cap = b2_num_sms();     // ONE block per SM: the device-wide barrier needs co-residency
      pl->grid = dim3((unsigned)(b > cap ? cap : (b < 1 ? 1 : b)));
      pl->block = dim3(kBnB);
      pl->smem = bn_bwd_fused_smem((int)i[1]);
      pl->smem_limit = kBnFusedSmemMax;
      return bn_fused_attrs();
    }
    case B2SEG_EW_FILL: {
      B2_REQUIRE(p[0] && i[0] >= 4 && i[0] % 4 == 0, "fill: bytes must be a positive multiple of 4");
      pl->fn = launch_fill;
      pl->grid = dim3(ew_blocks(i[0] / 4));
      pl->block = dim3(kEwThreads);
      return B2SEG_OK;
    }
  }
  b2seg_set_error("ew_plan_bn: kind %d not handled here", d->kind);
  return B2SEG_EINVAL;
}
