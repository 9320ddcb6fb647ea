// b2seg_ew_create: planned-op wrapper around the HBM-bound layer kernels (bn.cu, stem_pool.cu,
// attn.cu, se.cu).  Grid/block/shared-memory sizes are fixed at creation; run() is launch-only.
#include <new>
#include <string.h>

#include "ew.cuh"

int ew_plan_se(const b2seg_ew_desc* d, EwPlan* plan);
int ew_dot_rows(long long P, int C);
int ew_se_rows(int HW, int C);

namespace {

struct EwOp : b2seg_op {
  EwArgs args;
  EwPlan plan;
  int kind;
  const char* name() const override { return "ew"; }
  int launches() const override { return plan.launches; }
  int run(cudaStream_t st) override { return plan.fn(args, plan.grid, plan.block, plan.smem, st); }
};

}  // namespace

extern "C" int b2seg_ew_create(const b2seg_ew_desc* d, b2seg_op** out) {
  B2_REQUIRE(d && out, "ew_create: null argument");
  EwPlan plan;
  memset(&plan, 0, sizeof(plan));
  int rc;
  switch (d->kind) {
    case B2SEG_EW_BN_FINALIZE: case B2SEG_EW_BN_APPLY: case B2SEG_EW_BN_BWD_REDUCE:
    case B2SEG_EW_BN_BWD_FINALIZE: case B2SEG_EW_BN_BWD_APPLY: case B2SEG_EW_COLSUM:
    case B2SEG_EW_COLSUM_FINALIZE: case B2SEG_EW_PACK: case B2SEG_EW_FILL:
    case B2SEG_EW_BN_APPLY_FUSED: case B2SEG_EW_BN_BWD_FUSED:
      rc = ew_plan_bn(d, &plan);
      break;
    case B2SEG_EW_STEM_FWD: case B2SEG_EW_STEM_WGRAD: case B2SEG_EW_MAXPOOL_FWD:
    case B2SEG_EW_MAXPOOL_BWD:
      rc = ew_plan_stem(d, &plan);
      break;
    case B2SEG_EW_DOT1_FWD: case B2SEG_EW_DOT1_BWD: case B2SEG_EW_GATE_APPLY:
    case B2SEG_EW_GATE_BWD:
      rc = ew_plan_attn(d, &plan);
      break;
    case B2SEG_EW_SE_POOL: case B2SEG_EW_SE_MLP: case B2SEG_EW_SE_SCALE:
    case B2SEG_EW_SE_BWD_REDUCE: case B2SEG_EW_SE_MLP_BWD: case B2SEG_EW_SE_BWD_APPLY:
      rc = ew_plan_se(d, &plan);
      break;
    default:
      b2seg_set_error("ew_create: unknown kind %d", d->kind);
      return B2SEG_EINVAL;
  }
  if (rc != B2SEG_OK) return rc;
  if (plan.smem > (plan.smem_limit ? plan.smem_limit : (size_t)48 * 1024)) {
    b2seg_set_error("ew_create: kind %d needs %zu bytes of shared memory", d->kind, plan.smem);
    return B2SEG_EINVAL;
  }
  auto* op = new (std::nothrow) EwOp();
  B2_REQUIRE(op, "ew_create: out of host memory");
  for (int k = 0; k < B2SEG_EW_NPTR; ++k) op->args.p[k] = d->p[k];
  for (int k = 0; k < B2SEG_EW_NINT; ++k) op->args.i[k] = d->i[k];
  for (int k = 0; k < 4; ++k) op->args.f[k] = d->f[k];
  op->plan = plan;
  op->kind = d->kind;
  *out = op;
  return B2SEG_OK;
}

extern "C" int b2seg_ew_partial_rows(int kind, int64_t P, int C) {
  switch (kind) {
    case B2SEG_EW_BN_BWD_REDUCE:
      return (C >= 8 && C % 8 == 0 && P >= 1) ? ew_rows(P, C, kBnBwdReduceBlocksPerSm) : 0;
    case B2SEG_EW_COLSUM:
      return (C >= 8 && C % 8 == 0 && P >= 1) ? ew_rows(P, C) : 0;
    case B2SEG_EW_DOT1_FWD: case B2SEG_EW_DOT1_BWD: case B2SEG_EW_GATE_BWD:
      return (C >= 8 && C % 8 == 0 && P >= 1) ? ew_dot_rows(P, C) : 0;
    case B2SEG_EW_SE_POOL: case B2SEG_EW_SE_BWD_REDUCE:  /* P = H*W of one sample */
      return (C >= 8 && C % 8 == 0 && P >= 1) ? ew_se_rows((int)P, C) : 0;
    default:
      return 0;
  }
}
