// Multi-tensor Adam: one launch updates every parameter of a param group.
//
// Replaces `optimizer.step()` of the reference's training loop (train.py:49) for the optimizer its
// drivers build -- torch.optim.Adam(lr, weight_decay) at train.py:331-335 and
// cross_validation.py:93-97 -- whose stock implementation walks ~250 small tensors with several
// elementwise kernels each.  Here the parameters are cut into fixed-size chunks described by a
// device table; a CTA streams one chunk: read p, g, m, v (16 B/elem), write p, m, v (12 B/elem).
// HBM-bound: 28 B per parameter element.
//
// Arithmetic follows torch's single-tensor Adam (amsgrad=False, maximize=False) step by step:
//   g' = g + wd * p;  m += (g' - m) * (1 - b1);  v = v * b2 + (1 - b2) * g' * g';
//   p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
#include "common.cuh"

namespace {

constexpr int kAdamThreads = 256;

struct AdamHyper {
  float lr_over_bc1, bc2_sqrt, beta1, beta2, eps, wd;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamHyper& h) {
  if (h.wd != 0.0f) g = fmaf(h.wd, p, g);
  m = fmaf(g - m, 1.0f - h.beta1, m);
  v = fmaf(v, h.beta2, (1.0f - h.beta2) * g * g);
  const float denom = sqrtf(v) / h.bc2_sqrt + h.eps;
  p = p - h.lr_over_bc1 * (m / denom);
}

__global__ void __launch_bounds__(kAdamThreads)
    adam_chunks_kernel(const b2seg_adam_chunk* __restrict__ chunks, AdamHyper h) {
  pdl_prologue();
  const b2seg_adam_chunk c = chunks[blockIdx.x];
  float* __restrict__ p = c.p;
  const float* __restrict__ g = c.g;
  float* __restrict__ m = c.m;
  float* __restrict__ v = c.v;
  const int n = c.n;
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                     reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0;
  int done = 0;
  if (vec) {
    const int n4 = n >> 2;
    float4* p4 = reinterpret_cast<float4*>(p);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* v4 = reinterpret_cast<float4*>(v);
    for (int i = threadIdx.x; i < n4; i += kAdamThreads) {
      float4 pp = p4[i], mm = m4[i], vv = v4[i];
      const float4 gg = g4[i];
      adam_one(pp.x, gg.x, mm.x, vv.x, h);
      adam_one(pp.y, gg.y, mm.y, vv.y, h);
      adam_one(pp.z, gg.z, mm.z, vv.z, h);
      adam_one(pp.w, gg.w, mm.w, vv.w, h);
      p4[i] = pp; m4[i] = mm; v4[i] = vv;
    }
    done = n4 << 2;
  }
  for (int i = done + threadIdx.x; i < n; i += kAdamThreads) {
    float pp = p[i], mm = m[i], vv = v[i];
    adam_one(pp, g[i], mm, vv, h);
    p[i] = pp; m[i] = mm; v[i] = vv;
  }
}

}  // namespace

extern "C" int b2seg_adam_step(const b2seg_adam_chunk* chunks, int nchunks, float lr, float beta1,
                               float beta2, float eps, float weight_decay, double step,
                               void* stream) {
  B2_REQUIRE(chunks && nchunks > 0, "adam_step: empty chunk table");
  B2_REQUIRE(step >= 1.0, "adam_step: step %g must be >= 1", step);
  B2_REQUIRE(beta1 >= 0.0f && beta1 < 1.0f && beta2 >= 0.0f && beta2 < 1.0f && eps >= 0.0f,
             "adam_step: bad hyper-parameters");
  AdamHyper h;
  const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
  h.lr_over_bc1 = (float)((double)lr / bc1);
  h.bc2_sqrt = (float)sqrt(bc2);
  h.beta1 = beta1; h.beta2 = beta2; h.eps = eps; h.wd = weight_decay;
  B2_CHECK_CUDA(b2_launch(adam_chunks_kernel, dim3((unsigned)nchunks), dim3(kAdamThreads), 0,
                          (cudaStream_t)stream, chunks, h));
  B2_CHECK_CUDA(cudaGetLastError());
  return B2SEG_OK;
}
