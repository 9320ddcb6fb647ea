// Fused per-pixel segmentation losses + metrics (HBM-bound, single pass).
//
// Every closed-form loss of the reference (losses.py:13-171, 322-342) is a
// function of six batch-global sums; calculate_metrics (utils.py:120-151) adds
// two thresholded counts.  The forward kernel reads logits and targets exactly
// once (8 B/px), reduces 8 accumulators warp -> block -> grid (last block done
// finishes in fixed order, so results are run-to-run deterministic), evaluates
// the weighted loss and leaves the gradient coefficients on the device.  The
// backward kernel is a second streaming pass (12 B/px).
#include "common.cuh"

#include <cstddef>
#include <mutex>

namespace {

constexpr int kThreads = 256;
constexpr int kAcc = 8;
constexpr int kMaxBlocks = 148 * 8;

struct LossSpecDev {
  b2seg_loss_spec s;
};

// ---- per-element math ------------------------------------------------------
struct PixelTerms {
  float p, bce, one_m_pt;
};

// Single-instruction special functions (MUFU): the range fix-ups and the Newton step the CUDA
// intrinsics wrap around them cost ~15 issue slots per pixel and a divergent slow-path call, which
// made these streaming kernels issue-bound instead of HBM-bound.  Inputs here are always in the
// units' accurate range (e in (0,1], 1+e in [1,2]); relative error <= 2^-22.
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2_approx(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// softplus(-|x|) = log1p(exp(-|x|)) without the 1+e rounding loss for small e.
__device__ __forceinline__ float log1p_exp_neg(float e, float one_plus_e) {
  // e in (0, 1].  Series below 2^-6: error < e^4/4 relative 6e-8.
  const float series = e * (1.0f - e * (0.5f - e * (1.0f / 3.0f)));
  const float lg = lg2_approx(one_plus_e) * 0.69314718f;
  return e < 0.015625f ? series : lg;
}

// SOFT = false is the hard-target shortcut of the focal term: pt = exp(-bce) (losses.py:78-80) is
// p for t = 1 and 1-p for t = 0, and both are already in registers (cancellation-free), so the
// second exp is skipped.  Callers pick SOFT per group of 8 pixels (see is_hard4).
template <bool NEED_FOCAL, bool SOFT = true>
__device__ __forceinline__ PixelTerms pixel_terms(float x, float t) {
  PixelTerms r;
  const float e = ex2_approx(fabsf(x) * -1.44269504f);
  const float s1 = 1.0f + e;
  const float inv = rcp_approx(s1);
  const float small = e * inv;  // min(p, 1-p)
  r.p = x >= 0.0f ? inv : small;
  r.bce = fmaxf(x, 0.0f) - x * t + log1p_exp_neg(e, s1);
  if (NEED_FOCAL) {
    if (SOFT) {
      r.one_m_pt = 1.0f - ex2_approx(r.bce * -1.44269504f);
    } else {
      // 1 - pt = sigmoid(-z) with z = x for t = 1 and -x for t = 0: min(p, 1-p) when z >= 0
      r.one_m_pt = ((x >= 0.0f) == (t == 1.0f)) ? small : inv;
    }
  } else {
    r.one_m_pt = 0.0f;
  }
  return r;
}
__device__ __forceinline__ bool is_hard4(const float4& t) {
  return (t.x == 0.0f || t.x == 1.0f) && (t.y == 0.0f || t.y == 1.0f) &&
         (t.z == 0.0f || t.z == 1.0f) && (t.w == 0.0f || t.w == 1.0f);
}
// Eight targets are all hard iff sum(t^2 - t) == 0: every term is <= 0 on [0, 1] and zero only at 0 and 1
// (exactly, in floating point: t*t - t rounds to 0 only for t in {0, 1}).  8 FFMA + 7 FADD instead of 16
// compares and their predicate logic -- the focal forward pass is issue bound, every slot per pixel counts.
__device__ __forceinline__ bool is_hard8(const float4& a, const float4& b) {
  const float s = ((fmaf(a.x, a.x, -a.x) + fmaf(a.y, a.y, -a.y)) + (fmaf(a.z, a.z, -a.z) + fmaf(a.w, a.w, -a.w))) +
                  ((fmaf(b.x, b.x, -b.x) + fmaf(b.y, b.y, -b.y)) + (fmaf(b.z, b.z, -b.z) + fmaf(b.w, b.w, -b.w)));
  return s == 0.0f;
}

__device__ __forceinline__ float pow_pos(float base, float g) {
  // base in [0,1]; g > 0.  lg2(0) = -inf -> ex2(-inf) = 0.
  return ex2_approx(g * lg2_approx(base));
}

// Per-pixel term of ExponentialLogarithmicLoss (losses.py:190-192), in the reference's own float32
// formulation -- through probabilities, F.binary_cross_entropy's logs clamped at -100, then
//   f = (-log(clamp(1 - ce, 1e-7)))^g,   df/dx = g u^(g-1) / (1 - ce) * (p - t),  u = -log(1 - ce).
// For confident pixels `1 - ce` sits a few float32 steps below 1 and f is steep there, so that
// rounding is part of the reference's answer and is kept (accurate logf / log1pf, rn subtract).
// 1 - ce < 1e-7 (a badly wrong pixel) is the clamped constant with zero slope, as autograd sees
// it; u == 0 (p rounds to the target) has an infinite slope that the reference turns into NaN
// through inf * 0 -- it contributes 0 here.
__device__ __forceinline__ float explog_om(float p, float t) {
  const float l1 = fmaxf(logf(p), -100.0f), l0 = fmaxf(log1pf(-p), -100.0f);
  const float ce = __fsub_rn(__fmul_rn(t - 1.0f, l0), __fmul_rn(t, l1));
  return __fsub_rn(1.0f, ce);
}
__device__ __forceinline__ float explog_f(float om, float g) {
  const float u = -logf(fmaxf(om, 1e-7f));
  return u > 0.0f ? powf(u, g) : 0.0f;
}
__device__ __forceinline__ float explog_df(float om, float g) {
  if (!(om >= 1e-7f)) return 0.0f;
  const float u = -logf(om);
  return u > 0.0f ? g * powf(u, g - 1.0f) / om : 0.0f;
}

// MODE selects what accumulator 5 carries: 0 nothing, 1 the focal term, 2 the exp-log term.
// G2: 1 = gamma is 2 (compile time), 0 = it is not, -1 = decided per call from gamma_is_two.
template <int MODE, bool HAS_WMAP, bool SOFT = true, int G2 = -1>
__device__ __forceinline__ void accumulate(float (&acc)[kAcc], float x, float t, float w,
                                           float gamma, bool gamma_is_two) {
  constexpr bool NEED_FOCAL = MODE == 1;
  const PixelTerms q = pixel_terms<NEED_FOCAL, SOFT>(x, t);
  const float bce_w = HAS_WMAP ? q.bce * w : q.bce;
  acc[0] += q.p;
  acc[1] += t;
  acc[2] = fmaf(q.p, t, acc[2]);
  acc[3] += bce_w;
  acc[4] = fmaf(q.bce, t, acc[4]);
  if (NEED_FOCAL) {
    const bool sq = G2 < 0 ? gamma_is_two : G2 == 1;
    const float m = sq ? q.one_m_pt * q.one_m_pt : pow_pos(q.one_m_pt, gamma);
    acc[5] = fmaf(m, q.bce, acc[5]);
  }
  if (MODE == 2) acc[5] += explog_f(explog_om(q.p, t), gamma);
  const float pb = q.p > 0.5f ? 1.0f : 0.0f;
  acc[6] += pb;
  acc[7] = fmaf(pb, t, acc[7]);
}

// ---- reduction tail shared by loss-forward and metrics kernels --------------
struct ReduceWs {
  unsigned int ticket;
  unsigned int pad[3];
  double partial[kMaxBlocks][kAcc];
};

// Returns true (to every thread of exactly one block) once all blocks have
// published their partials; that block then holds the grid totals in tot[].
__device__ __forceinline__ bool grid_reduce(float (&acc)[kAcc], ReduceWs* ws, double (&tot)[kAcc],
                                            double* s_tot /* smem [kAcc] */) {
  __shared__ float s_part[kThreads / 32][kAcc];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < kAcc; ++i) acc[i] = warp_sum(acc[i]);
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < kAcc; ++i) s_part[warp][i] = acc[i];
  }
  __syncthreads();
  if (threadIdx.x < kAcc) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) s += (double)s_part[w][threadIdx.x];
    ws->partial[blockIdx.x][threadIdx.x] = s;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(&ws->ticket, 1u);
    s_last = (tk == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return false;
  __threadfence();
  // One warp per accumulator, fixed block order -> deterministic.
  if (warp < kAcc) {
    double s = 0.0;
    for (int b = lane; b < (int)gridDim.x; b += 32) s += ws->partial[b][warp];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) s_tot[warp] = s;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < kAcc; ++i) tot[i] = s_tot[i];
  if (threadIdx.x == 0) ws->ticket = 0;  // leave the workspace reusable
  return true;
}

__device__ void write_confusion(double* stats, double n) {
  const double pb = stats[B2SEG_S_PB], tp = stats[B2SEG_S_PB_T], st = stats[B2SEG_S_T];
  stats[B2SEG_S_N] = n;
  stats[B2SEG_S_TP] = tp;
  stats[B2SEG_S_FP] = pb - tp;
  stats[B2SEG_S_FN] = st - tp;
  stats[B2SEG_S_TN] = n - pb - st + tp;
}

// Loss value + gradient coefficients from the six sums.
//   dL/dx_i = (A t_i + B) p_i (1-p_i) + (W0 + W1 t_i [+ Wm w_i]) (p_i - t_i)
//             + F ((1-pt)^g + g (1-pt)^(g-1) pt bce_i) (p_i - t_i)
__device__ void finish_loss(const b2seg_loss_spec& s, const double* tot, double n, double* stats,
                            float* loss_out, float* coef) {
  const double Sp = tot[0], St = tot[1], Spt = tot[2], Sbce = tot[3], Sbce_t = tot[4],
               Sfoc = tot[5];
  double loss = 0.0, A = 0.0, B = 0.0, W0 = 0.0, W1 = 0.0, Wm = 0.0, F = 0.0;

  if (s.w_dice != 0.0f || s.w_logcosh_dice != 0.0f) {
    const double sm = s.dice_smooth, U = Sp + St;
    const double dice = (2.0 * Spt + sm) / (U + sm);
    const double Ld = 1.0 - dice;
    const double a = -2.0 / (U + sm), b = (2.0 * Spt + sm) / ((U + sm) * (U + sm));
    loss += s.w_dice * Ld;
    A += s.w_dice * a;
    B += s.w_dice * b;
    if (s.w_logcosh_dice != 0.0f) {
      const double th = tanh(Ld);
      loss += s.w_logcosh_dice * log(cosh(Ld));
      A += s.w_logcosh_dice * th * a;
      B += s.w_logcosh_dice * th * b;
    }
  }
  if (s.w_tversky != 0.0f) {
    const double al = s.tv_alpha, be = s.tv_beta, sm = s.tv_smooth, g = s.tv_gamma;
    const double TP = Spt, FP = Sp - Spt, FN = St - Spt;
    const double D = TP + al * FP + be * FN + sm;
    const double Lt = 1.0 - (TP + sm) / D;
    double val = Lt, fac = 1.0;
    if (g != 1.0) {
      val = pow(Lt, g);
      fac = g * pow(Lt, g - 1.0);
    }
    loss += s.w_tversky * val;
    A += s.w_tversky * fac * (-(D - (TP + sm) * (1.0 - al - be)) / (D * D));
    B += s.w_tversky * fac * ((TP + sm) * al / (D * D));
  }
  if (s.w_sens_spec != 0.0f) {
    const double w = s.ss_w, eps = s.ss_eps;
    const double TP = Spt, FN = St - Spt, FP = Sp - Spt, TN = n - Sp - St + Spt;
    const double sens = TP / (TP + FN + eps), spec = TN / (TN + FP + eps);
    loss += s.w_sens_spec * (1.0 - (w * sens + (1.0 - w) * spec));
    const double ds = 1.0 / (St + eps), dn = 1.0 / (n - St + eps);
    A += s.w_sens_spec * (-w * ds - (1.0 - w) * dn);
    B += s.w_sens_spec * ((1.0 - w) * dn);
  }
  if (s.w_bce != 0.0f) {
    double w0 = 1.0, w1 = 0.0;
    if (s.bce_mode == 1) {
      w1 = (double)s.bce_beta - 1.0;
    } else if (s.bce_mode == 2) {
      const double beta = 1.0 - St / n;
      w0 = 1.0 - beta;
      w1 = 2.0 * beta - 1.0;
    }
    if (s.bce_mode == 3) {  // Sbce already carries the per-pixel weight map
      loss += s.w_bce * Sbce / n;
      Wm += s.w_bce / n;
    } else {
      loss += s.w_bce * (w0 * Sbce + w1 * Sbce_t) / n;
      W0 += s.w_bce * w0 / n;
      W1 += s.w_bce * w1 / n;
    }
  }
  if (s.w_focal != 0.0f) {
    loss += s.w_focal * s.focal_alpha * Sfoc / n;
    F += s.w_focal * s.focal_alpha / n;
  }
  if (s.w_explog_dice != 0.0f) {  // w (-log(clamp(1 - dice_loss, 1e-7)))^g, losses.py:186-187
    const double sm = s.dice_smooth, U = Sp + St, g = s.el_gamma_dice;
    const double dice = (2.0 * Spt + sm) / (U + sm);
    const double a = -2.0 / (U + sm), b = (2.0 * Spt + sm) / ((U + sm) * (U + sm));
    const bool clamped = !(dice > 1e-7);
    const double u = -log(clamped ? 1e-7 : dice);
    loss += s.w_explog_dice * (u > 0.0 ? pow(u, g) : 0.0);
    const double fac = (clamped || !(u > 0.0)) ? 0.0 : g * pow(u, g - 1.0) / dice;  // d/d(dice_loss)
    A += s.w_explog_dice * fac * a;
    B += s.w_explog_dice * fac * b;
  }
  const bool explog = s.w_explog_cross != 0.0f;
  if (explog) {  // Sfoc carries sum f(bce_i)
    loss += s.w_explog_cross * Sfoc / n;
    F = s.w_explog_cross / n;
  }
  for (int i = 0; i < kAcc; ++i) stats[i] = tot[i];
  write_confusion(stats, n);
  stats[B2SEG_S_LOSS] = loss;
  *loss_out = (float)loss;
  coef[0] = (float)A;
  coef[1] = (float)B;
  coef[2] = (float)W0;
  coef[3] = (float)W1;
  coef[4] = (float)F;
  coef[5] = explog ? s.el_gamma_cross : s.focal_gamma;
  coef[6] = (float)Wm;
  coef[7] = explog ? 2.0f : 0.0f;  // what coef[4] multiplies: the focal or the exp-log pixel term
}

// ---- forward kernel -----------------------------------------------------------
template <int NEED_FOCAL, bool HAS_WMAP, bool SOFT, int G2 = -1>
__device__ __forceinline__ void accumulate4(float (&acc)[kAcc], const float4& x, const float4& t,
                                            const float4& w, float gamma, bool g2) {
  accumulate<NEED_FOCAL, HAS_WMAP, SOFT, G2>(acc, x.x, t.x, w.x, gamma, g2);
  accumulate<NEED_FOCAL, HAS_WMAP, SOFT, G2>(acc, x.y, t.y, w.y, gamma, g2);
  accumulate<NEED_FOCAL, HAS_WMAP, SOFT, G2>(acc, x.z, t.z, w.z, gamma, g2);
  accumulate<NEED_FOCAL, HAS_WMAP, SOFT, G2>(acc, x.w, t.w, w.w, gamma, g2);
}

template <int NEED_FOCAL, bool HAS_WMAP, bool VEC4>
__global__ void __launch_bounds__(kThreads)
    seg_loss_fwd_kernel(const float* __restrict__ x, const float* __restrict__ t,
                        const float* __restrict__ wm, long long n, LossSpecDev spec, ReduceWs* ws,
                        double* stats, float* loss_out, float* coef) {
  float acc[kAcc];
#pragma unroll
  for (int i = 0; i < kAcc; ++i) acc[i] = 0.0f;
  const float gamma = NEED_FOCAL == 2 ? spec.s.el_gamma_cross : spec.s.focal_gamma;
  const bool g2 = gamma == 2.0f;
  const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
  const long long nthreads = (long long)gridDim.x * kThreads;

  if (VEC4) {
    const long long n4 = n >> 2;
    const float4* x4 = reinterpret_cast<const float4*>(x);
    const float4* t4 = reinterpret_cast<const float4*>(t);
    const float4* w4 = reinterpret_cast<const float4*>(wm);
    long long i = tid;
    // two independent 128-bit loads per stream in flight per iteration
    for (; i + nthreads < n4; i += 2 * nthreads) {
      const float4 xa = ldg_stream_f4(x4 + i), xb = ldg_stream_f4(x4 + i + nthreads);
      const float4 ta = ldg_stream_f4(t4 + i), tb = ldg_stream_f4(t4 + i + nthreads);
      float4 wa = make_float4(1.f, 1.f, 1.f, 1.f), wb = wa;
      if (HAS_WMAP) {
        wa = ldg_stream_f4(w4 + i);
        wb = ldg_stream_f4(w4 + i + nthreads);
      }
      if (NEED_FOCAL == 1 && is_hard8(ta, tb)) {
        if (g2) {  // the factory default (losses.py:68): (1 - pt)^2 is one multiply
          accumulate4<NEED_FOCAL, HAS_WMAP, false, 1>(acc, xa, ta, wa, gamma, true);
          accumulate4<NEED_FOCAL, HAS_WMAP, false, 1>(acc, xb, tb, wb, gamma, true);
        } else {
          accumulate4<NEED_FOCAL, HAS_WMAP, false, 0>(acc, xa, ta, wa, gamma, false);
          accumulate4<NEED_FOCAL, HAS_WMAP, false, 0>(acc, xb, tb, wb, gamma, false);
        }
      } else {
        accumulate4<NEED_FOCAL, HAS_WMAP, true>(acc, xa, ta, wa, gamma, g2);
        accumulate4<NEED_FOCAL, HAS_WMAP, true>(acc, xb, tb, wb, gamma, g2);
      }
    }
    for (; i < n4; i += nthreads) {
      const float4 xa = ldg_stream_f4(x4 + i);
      const float4 ta = ldg_stream_f4(t4 + i);
      float4 wa = make_float4(1.f, 1.f, 1.f, 1.f);
      if (HAS_WMAP) wa = ldg_stream_f4(w4 + i);
      accumulate<NEED_FOCAL, HAS_WMAP>(acc, xa.x, ta.x, wa.x, gamma, g2);
      accumulate<NEED_FOCAL, HAS_WMAP>(acc, xa.y, ta.y, wa.y, gamma, g2);
      accumulate<NEED_FOCAL, HAS_WMAP>(acc, xa.z, ta.z, wa.z, gamma, g2);
      accumulate<NEED_FOCAL, HAS_WMAP>(acc, xa.w, ta.w, wa.w, gamma, g2);
    }
    for (long long j = (n4 << 2) + tid; j < n; j += nthreads)
      accumulate<NEED_FOCAL, HAS_WMAP>(acc, x[j], t[j], HAS_WMAP ? wm[j] : 1.0f, gamma, g2);
  } else {
    for (long long j = tid; j < n; j += nthreads)
      accumulate<NEED_FOCAL, HAS_WMAP>(acc, x[j], t[j], HAS_WMAP ? wm[j] : 1.0f, gamma, g2);
  }

  __shared__ double s_tot[kAcc];
  double tot[kAcc];
  if (grid_reduce(acc, ws, tot, s_tot) && threadIdx.x == 0)
    finish_loss(spec.s, tot, (double)n, stats, loss_out, coef);
}

// ---- backward kernel ----------------------------------------------------------
template <int MODE, bool HAS_WMAP, bool SOFT = true>
__device__ __forceinline__ float pixel_grad(float x, float t, float w, float A, float B, float W0,
                                            float W1, float F, float Wm, float gamma,
                                            bool gamma_is_two, float go) {
  constexpr bool NEED_FOCAL = MODE == 1;
  const PixelTerms q = pixel_terms<NEED_FOCAL, SOFT>(x, t);
  const float d = q.p - t;
  float g = (fmaf(A, t, B)) * q.p * (1.0f - q.p);
  float wl = fmaf(W1, t, W0);
  if (HAS_WMAP) wl = fmaf(Wm, w, wl);
  if (NEED_FOCAL) {
    const float m = q.one_m_pt, pt = 1.0f - q.one_m_pt;
    float mg, mg1;  // m^g, g*m^(g-1)
    if (gamma_is_two) {
      mg = m * m;
      mg1 = 2.0f * m;
    } else {
      mg1 = gamma * pow_pos(m, gamma - 1.0f);
      mg = pow_pos(m, gamma);
      if (gamma == 1.0f) mg1 = 1.0f;
    }
    wl = fmaf(F, fmaf(mg1 * pt, q.bce, mg), wl);
  }
  if (MODE == 2) wl = fmaf(F, explog_df(explog_om(q.p, t), gamma), wl);
  g = fmaf(wl, d, g);
  return g * go;
}

struct BwdCoef {
  float A, B, W0, W1, F, gamma, Wm, go;
  bool g2;
};
template <int NEED_FOCAL, bool HAS_WMAP, bool SOFT>
__device__ __forceinline__ float4 pixel_grad4(const float4& x, const float4& t, const float4& w,
                                              const BwdCoef& c) {
  float4 g;
  g.x = pixel_grad<NEED_FOCAL, HAS_WMAP, SOFT>(x.x, t.x, w.x, c.A, c.B, c.W0, c.W1, c.F, c.Wm, c.gamma, c.g2, c.go);
  g.y = pixel_grad<NEED_FOCAL, HAS_WMAP, SOFT>(x.y, t.y, w.y, c.A, c.B, c.W0, c.W1, c.F, c.Wm, c.gamma, c.g2, c.go);
  g.z = pixel_grad<NEED_FOCAL, HAS_WMAP, SOFT>(x.z, t.z, w.z, c.A, c.B, c.W0, c.W1, c.F, c.Wm, c.gamma, c.g2, c.go);
  g.w = pixel_grad<NEED_FOCAL, HAS_WMAP, SOFT>(x.w, t.w, w.w, c.A, c.B, c.W0, c.W1, c.F, c.Wm, c.gamma, c.g2, c.go);
  return g;
}

template <int NEED_FOCAL, bool HAS_WMAP, bool VEC4>
__device__ __forceinline__ void seg_loss_bwd_body(const float* __restrict__ x, const float* __restrict__ t,
                                                  const float* __restrict__ wm, long long n,
                                                  const BwdCoef& c, float* __restrict__ gx) {
  const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
  const long long nthreads = (long long)gridDim.x * kThreads;
#define B2_PG(xx, tt, ww) \
  pixel_grad<NEED_FOCAL, HAS_WMAP, true>(xx, tt, ww, c.A, c.B, c.W0, c.W1, c.F, c.Wm, c.gamma, c.g2, c.go)
  if (VEC4) {
    const long long n4 = n >> 2;
    const float4* x4 = reinterpret_cast<const float4*>(x);
    const float4* t4 = reinterpret_cast<const float4*>(t);
    const float4* w4 = reinterpret_cast<const float4*>(wm);
    float4* g4 = reinterpret_cast<float4*>(gx);
    long long i = tid;
    for (; i + nthreads < n4; i += 2 * nthreads) {
      const float4 xa = ldg_stream_f4(x4 + i), xb = ldg_stream_f4(x4 + i + nthreads);
      const float4 ta = ldg_stream_f4(t4 + i), tb = ldg_stream_f4(t4 + i + nthreads);
      float4 wa = make_float4(1.f, 1.f, 1.f, 1.f), wb = wa;
      if (HAS_WMAP) {
        wa = ldg_stream_f4(w4 + i);
        wb = ldg_stream_f4(w4 + i + nthreads);
      }
      float4 ga, gb;
      if (NEED_FOCAL == 1 && is_hard8(ta, tb)) {
        ga = pixel_grad4<NEED_FOCAL, HAS_WMAP, false>(xa, ta, wa, c);
        gb = pixel_grad4<NEED_FOCAL, HAS_WMAP, false>(xb, tb, wb, c);
      } else {
        ga = pixel_grad4<NEED_FOCAL, HAS_WMAP, true>(xa, ta, wa, c);
        gb = pixel_grad4<NEED_FOCAL, HAS_WMAP, true>(xb, tb, wb, c);
      }
      stg_stream_f4(g4 + i, ga);
      stg_stream_f4(g4 + i + nthreads, gb);
    }
    for (; i < n4; i += nthreads) {
      const float4 xa = ldg_stream_f4(x4 + i);
      const float4 ta = ldg_stream_f4(t4 + i);
      float4 wa = make_float4(1.f, 1.f, 1.f, 1.f);
      if (HAS_WMAP) wa = ldg_stream_f4(w4 + i);
      stg_stream_f4(g4 + i, pixel_grad4<NEED_FOCAL, HAS_WMAP, true>(xa, ta, wa, c));
    }
    for (long long j = (n4 << 2) + tid; j < n; j += nthreads)
      gx[j] = B2_PG(x[j], t[j], HAS_WMAP ? wm[j] : 1.0f);
  } else {
    for (long long j = tid; j < n; j += nthreads)
      gx[j] = B2_PG(x[j], t[j], HAS_WMAP ? wm[j] : 1.0f);
  }
#undef B2_PG
}

// The focal coefficient is only known on the device (the forward kernel wrote it), so the
// kernel branches once, uniformly, between the focal and the focal-free instantiation.
template <int NEED_FOCAL_UNUSED, bool HAS_WMAP, bool VEC4>
__global__ void __launch_bounds__(kThreads)
    seg_loss_bwd_kernel(const float* __restrict__ x, const float* __restrict__ t,
                        const float* __restrict__ wm, long long n,
                        const float* __restrict__ coef, const float* __restrict__ grad_out,
                        float* __restrict__ gx) {
  BwdCoef c;
  c.A = coef[0]; c.B = coef[1]; c.W0 = coef[2]; c.W1 = coef[3]; c.F = coef[4]; c.gamma = coef[5];
  c.Wm = coef[6];
  c.go = grad_out ? *grad_out : 1.0f;
  c.g2 = c.gamma == 2.0f;
  if (c.F == 0.0f)
    seg_loss_bwd_body<0, HAS_WMAP, VEC4>(x, t, wm, n, c, gx);
  else if (coef[7] == 2.0f)
    seg_loss_bwd_body<2, HAS_WMAP, VEC4>(x, t, wm, n, c, gx);
  else
    seg_loss_bwd_body<1, HAS_WMAP, VEC4>(x, t, wm, n, c, gx);
}

// ---- metrics from probabilities (utils.py:120-151) --------------------------------
template <bool VEC4>
__global__ void __launch_bounds__(kThreads)
    seg_metrics_probs_kernel(const float* __restrict__ p, const float* __restrict__ t,
                             long long n, ReduceWs* ws, double* stats) {
  float acc[kAcc];
#pragma unroll
  for (int i = 0; i < kAcc; ++i) acc[i] = 0.0f;
  const long long tid = (long long)blockIdx.x * kThreads + threadIdx.x;
  const long long nthreads = (long long)gridDim.x * kThreads;
#define B2_ACC(pp, tt)                         \
  {                                            \
    const float pb = (pp) > 0.5f ? 1.f : 0.f;  \
    acc[0] += (pp);                            \
    acc[1] += (tt);                            \
    acc[2] = fmaf((pp), (tt), acc[2]);         \
    acc[6] += pb;                              \
    acc[7] = fmaf(pb, (tt), acc[7]);           \
  }
  if (VEC4) {
    const long long n4 = n >> 2;
    const float4* p4 = reinterpret_cast<const float4*>(p);
    const float4* t4 = reinterpret_cast<const float4*>(t);
    long long i = tid;
    for (; i + nthreads < n4; i += 2 * nthreads) {
      const float4 pa = ldg_stream_f4(p4 + i), pb2 = ldg_stream_f4(p4 + i + nthreads);
      const float4 ta = ldg_stream_f4(t4 + i), tb = ldg_stream_f4(t4 + i + nthreads);
      B2_ACC(pa.x, ta.x) B2_ACC(pa.y, ta.y) B2_ACC(pa.z, ta.z) B2_ACC(pa.w, ta.w)
      B2_ACC(pb2.x, tb.x) B2_ACC(pb2.y, tb.y) B2_ACC(pb2.z, tb.z) B2_ACC(pb2.w, tb.w)
    }
    for (; i < n4; i += nthreads) {
      const float4 pa = ldg_stream_f4(p4 + i);
      const float4 ta = ldg_stream_f4(t4 + i);
      B2_ACC(pa.x, ta.x) B2_ACC(pa.y, ta.y) B2_ACC(pa.z, ta.z) B2_ACC(pa.w, ta.w)
    }
    for (long long j = (n4 << 2) + tid; j < n; j += nthreads) B2_ACC(p[j], t[j])
  } else {
    for (long long j = tid; j < n; j += nthreads) B2_ACC(p[j], t[j])
  }
#undef B2_ACC
  __shared__ double s_tot[kAcc];
  double tot[kAcc];
  if (grid_reduce(acc, ws, tot, s_tot) && threadIdx.x == 0) {
    for (int i = 0; i < kAcc; ++i) stats[i] = tot[i];
    write_confusion(stats, (double)n);
    stats[B2SEG_S_LOSS] = 0.0;
  }
}

// Grid = what is needed, capped at ONE resident wave of this instantiation (grid-stride loops: a
// partial second wave would leave SMs idle for the tail).
template <auto Kern>
int resident_blocks() {
  static int per_sm = 0;
  if (per_sm == 0) {
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, Kern, kThreads, 0) != cudaSuccess || nb < 1) nb = 1;
    per_sm = nb;
  }
  return per_sm * b2_num_sms();
}

int pick_blocks(long long n, int per_thread, int resident) {
  long long b = (n + (long long)kThreads * per_thread - 1) / ((long long)kThreads * per_thread);
  const long long cap = resident;
  if (b > cap) b = cap;
  if (b > kMaxBlocks) b = kMaxBlocks;
  if (b < 1) b = 1;
  return (int)b;
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace

extern "C" size_t b2seg_seg_reduce_workspace_bytes(void) { return sizeof(ReduceWs); }

extern "C" int b2seg_seg_loss_fwd(const float* logits, const float* target,
                                  const float* weight_map, int64_t n,
                                  const b2seg_loss_spec* spec, void* workspace, double* stats,
                                  float* loss_out, float* coef, void* stream) {
  B2_REQUIRE(logits && target && spec && workspace && stats && loss_out && coef,
             "seg_loss_fwd: null argument");
  B2_REQUIRE(n > 0, "seg_loss_fwd: n must be positive (got %lld)", (long long)n);
  B2_REQUIRE(spec->bce_mode >= 0 && spec->bce_mode <= 3, "seg_loss_fwd: bad bce_mode %d",
             spec->bce_mode);
  const bool wmap = spec->bce_mode == 3 && spec->w_bce != 0.0f;
  B2_REQUIRE(!wmap || weight_map, "seg_loss_fwd: bce_mode 3 needs a weight map");
  B2_REQUIRE(spec->w_focal == 0.0f || spec->focal_gamma > 0.0f,
             "seg_loss_fwd: focal_gamma must be > 0");
  const bool focal = spec->w_focal != 0.0f;
  const bool explog = spec->w_explog_cross != 0.0f;
  B2_REQUIRE(!(focal && explog), "seg_loss_fwd: the focal and the exp-log pixel terms are exclusive");
  B2_REQUIRE((!explog || spec->el_gamma_cross > 0.0f) &&
                 (spec->w_explog_dice == 0.0f || spec->el_gamma_dice > 0.0f),
             "seg_loss_fwd: exp-log exponents must be > 0");
  const bool vec = aligned16(logits) && aligned16(target) && (!wmap || aligned16(weight_map));
  LossSpecDev sd;
  sd.s = *spec;
  cudaStream_t st = (cudaStream_t)stream;
  ReduceWs* ws = (ReduceWs*)workspace;
#define LAUNCH(F, W, V)                                                                            \
  seg_loss_fwd_kernel<F, W, V>                                                                     \
      <<<pick_blocks(n, 16, resident_blocks<seg_loss_fwd_kernel<F, W, V>>()), kThreads, 0, st>>>(  \
          logits, target, weight_map, n, sd, ws, stats, loss_out, coef)
  if (focal) {
    if (wmap) { if (vec) LAUNCH(1, true, true); else LAUNCH(1, true, false); }
    else      { if (vec) LAUNCH(1, false, true); else LAUNCH(1, false, false); }
  } else if (explog) {
    if (wmap) { if (vec) LAUNCH(2, true, true); else LAUNCH(2, true, false); }
    else      { if (vec) LAUNCH(2, false, true); else LAUNCH(2, false, false); }
  } else {
    if (wmap) { if (vec) LAUNCH(0, true, true); else LAUNCH(0, true, false); }
    else      { if (vec) LAUNCH(0, false, true); else LAUNCH(0, false, false); }
  }
#undef LAUNCH
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_seg_loss_bwd(const float* logits, const float* target,
                                  const float* weight_map, int64_t n, const float* coef,
                                  const float* grad_out, float* grad_logits, void* stream) {
  B2_REQUIRE(logits && target && coef && grad_logits, "seg_loss_bwd: null argument");
  B2_REQUIRE(n > 0, "seg_loss_bwd: n must be positive (got %lld)", (long long)n);
  const bool wmap = weight_map != nullptr;
  const bool vec = aligned16(logits) && aligned16(target) && aligned16(grad_logits) &&
                   (!wmap || aligned16(weight_map));
  cudaStream_t st = (cudaStream_t)stream;
  // The focal coefficient lives on the device: the kernel branches once, uniformly, between its
  // focal and focal-free bodies (the first template argument is unused).
#define LAUNCH(W, V)                                                                             \
  seg_loss_bwd_kernel<1, W, V>                                                                   \
      <<<pick_blocks(n, 8, resident_blocks<seg_loss_bwd_kernel<1, W, V>>()), kThreads, 0, st>>>( \
          logits, target, weight_map, n, coef, grad_out, grad_logits)
  if (wmap) { if (vec) LAUNCH(true, true); else LAUNCH(true, false); }
  else      { if (vec) LAUNCH(false, true); else LAUNCH(false, false); }
#undef LAUNCH
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

// Loss value and gradient coefficients from given sums (stats[0..7]) of `n` pixels: the second
// half of the forward kernel as its own launch, for sums that were combined across data-parallel
// replicas after the local reduction (Dice / Tversky normalise by global-batch totals).
namespace {
__global__ void seg_loss_finish_kernel(LossSpecDev spec, double n, double* stats, float* loss_out,
                                       float* coef) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double tot[kAcc];
    for (int i = 0; i < kAcc; ++i) tot[i] = stats[i];
    if (n <= 0.0) n = stats[B2SEG_S_N];  // pixel count summed on the device with the other sums
    finish_loss(spec.s, tot, n, stats, loss_out, coef);
  }
}
}  // namespace

extern "C" int b2seg_seg_loss_finish(const b2seg_loss_spec* spec, int64_t n_total, double* stats,
                                     float* loss_out, float* coef, void* stream) {
  B2_REQUIRE(spec && stats && loss_out && coef, "seg_loss_finish: null argument");
  B2_REQUIRE(n_total >= 0, "seg_loss_finish: n must be >= 0 (got %lld)", (long long)n_total);
  LossSpecDev sd;
  sd.s = *spec;
  seg_loss_finish_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(sd, (double)n_total, stats, loss_out, coef);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

extern "C" int b2seg_seg_metrics_from_probs(const float* probs, const float* target, int64_t n,
                                            void* workspace, double* stats, void* stream) {
  B2_REQUIRE(probs && target && workspace && stats, "seg_metrics: null argument");
  B2_REQUIRE(n > 0, "seg_metrics: n must be positive (got %lld)", (long long)n);
  const bool vec = aligned16(probs) && aligned16(target);
  cudaStream_t st = (cudaStream_t)stream;
  ReduceWs* ws = (ReduceWs*)workspace;
  if (vec)
    seg_metrics_probs_kernel<true>
        <<<pick_blocks(n, 16, resident_blocks<seg_metrics_probs_kernel<true>>()), kThreads, 0, st>>>(probs, target, n, ws, stats);
  else
    seg_metrics_probs_kernel<false>
        <<<pick_blocks(n, 16, resident_blocks<seg_metrics_probs_kernel<false>>()), kThreads, 0, st>>>(probs, target, n, ws, stats);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}

// ---- allocation-free variant -------------------------------------------------------------------
// b2seg_seg_metrics_from_probs with its reduction scratch in library-owned static device memory,
// for callers that may not call cudaMalloc (code under graph capture, serving loops, harnesses
// that forbid it).  No counterpart in the reference.
//
// g_noalloc_ws is one ReduceWs (75 792 B: a ticket plus kMaxBlocks x kAcc fp64 block partials) that
// every GPU loading the library holds until exit.  In HOST code the name is only a stand-in for
// the symbol: the device address comes from cudaGetSymbolAddress, looked up once per device with
// that device current, and that pointer is what the kernel gets.  All calls on a device share the
// array, so two calls that overlap on the GPU (different streams, no ordering) race on it.
namespace {
static __device__ __align__(256) unsigned char g_noalloc_ws[sizeof(ReduceWs)];

constexpr int kNoallocMaxDevices = 64;
ReduceWs* g_noalloc_ptr[kNoallocMaxDevices];  // device address of g_noalloc_ws, per device
std::mutex g_noalloc_mu;

// Device address of the static scratch on the CURRENT device (looked up on first use there).
int noalloc_ws(ReduceWs** out) {
  int dev = -1;
  B2_CHECK_CUDA(cudaGetDevice(&dev));
  B2_REQUIRE(dev >= 0 && dev < kNoallocMaxDevices, "seg_metrics_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* ptr = nullptr;
    B2_CHECK_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_ws));
    B2_REQUIRE(ptr != nullptr, "seg_metrics_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[dev] = (ReduceWs*)ptr;
  }
  *out = g_noalloc_ptr[dev];
  return B2SEG_OK;
}
}  // namespace

extern "C" int b2seg_seg_metrics_from_probs_noalloc(const float* probs, const float* target,
                                                    int64_t n, double* stats, void* stream) {
  B2_REQUIRE(probs && target && stats, "seg_metrics_noalloc: null argument");
  B2_REQUIRE(n > 0, "seg_metrics_noalloc: n must be positive (got %lld)", (long long)n);
  B2_REQUIRE(n <= (int64_t)B2SEG_NOALLOC_MAX_PIXELS,
             "seg_metrics_noalloc: n = %lld is beyond the static-scratch entry point's maximum of %lld "
             "pixels; use b2seg_seg_metrics_from_probs with a caller workspace",
             (long long)n, (long long)B2SEG_NOALLOC_MAX_PIXELS);
  ReduceWs* ws = nullptr;
  const int rc = noalloc_ws(&ws);
  if (rc != B2SEG_OK) return rc;
  const bool vec = aligned16(probs) && aligned16(target);
  cudaStream_t st = (cudaStream_t)stream;
  // The array keeps whatever the previous call (possibly an aborted one) left: the arrival ticket
  // must start at zero; the partial rows are overwritten before they are read.
  B2_CHECK_CUDA(cudaMemsetAsync(ws, 0, offsetof(ReduceWs, partial), st));
  int blocks;
  if (vec) {
    blocks = pick_blocks(n, 16, resident_blocks<seg_metrics_probs_kernel<true>>());
    seg_metrics_probs_kernel<true><<<blocks, kThreads, 0, st>>>(probs, target, n, ws, stats);
  } else {
    blocks = pick_blocks(n, 16, resident_blocks<seg_metrics_probs_kernel<false>>());
    seg_metrics_probs_kernel<false><<<blocks, kThreads, 0, st>>>(probs, target, n, ws, stats);
  }
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}
