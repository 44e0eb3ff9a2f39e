// Optimizer arithmetic shared by the step kernels (dgs_step.cu) and the single-launch step (dgs_direct.cu).
//
// Formulas follow torch 2.11 torch/optim/adam.py (_multi_tensor_adam) and torch/optim/sgd.py (_multi_tensor_sgd) op by
// op, in fp32 -- what optimizer.step() computes at reference harness/determined/pytorch/_pytorch_context.py:914-921.
#pragma once

#include "dgs_common.cuh"

namespace dgs {

// ---- device-side hyper-parameter tables (fp32 constants derived on the host in double) ----------
struct AdamHp {
  long long begin, end;
  float decay;     // decoupled: 1 - lr*wd ; L2: wd
  float w1;        // 1 - beta1
  float beta2;
  float w2;        // 1 - beta2
  float eps;
  float neg_step;  // -(lr / bias_correction1)
  float inv_bc2_sqrt;  // 1 / sqrt(1 - beta2^t)
  int flags;
};
struct SgdHp {
  long long begin, end;
  float wd, momentum, w_g /* 1 - dampening */, neg_lr;
  int flags;
  int pad_;
};
__device__ __forceinline__ float fast_sqrt(float x) {
  float r;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

__device__ __forceinline__ void adam_update(float& p, float& m, float& v, float g, const AdamHp& h,
                                            float clip) {
  g *= clip;
  if (h.flags & DGS_OPT_MAXIMIZE) g = -g;
  if (h.flags & DGS_ADAM_DECOUPLED_WD) p *= h.decay;
  else g = g + h.decay * p;
  m = m + h.w1 * (g - m);                      // _foreach_lerp_(exp_avg, grad, 1 - beta1)
  v = v * h.beta2;                             // _foreach_mul_(exp_avg_sq, beta2)
  v = v + h.w2 * (g * g);                      // _foreach_addcmul_(exp_avg_sq, grad, grad, 1 - beta2)
  // sqrt.approx / rcp.approx (<= 2 ulp each) instead of the IEEE sequences: at 14 B/element (bf16) the
  // correctly-rounded sqrt + two divisions made the kernel issue-bound instead of HBM-bound.
  const float denom = fast_sqrt(v) * h.inv_bc2_sqrt + h.eps;
  p = p + h.neg_step * __fdividef(m, denom);   // _foreach_addcdiv_(param, exp_avg, denom, step_size)
}

__device__ __forceinline__ void sgd_update(float& p, float& b, float g, const SgdHp& h, float clip,
                                           bool has_buf) {
  g *= clip;
  if (h.flags & DGS_OPT_MAXIMIZE) g = -g;
  if (h.wd != 0.0f) g = g + h.wd * p;
  if (has_buf) {
    if (h.flags & DGS_SGD_FIRST_STEP) b = g;   // buf = clone(grad)
    else b = b * h.momentum + h.w_g * g;       // buf.mul_(momentum).add_(grad, alpha=1-dampening)
    g = (h.flags & DGS_SGD_NESTEROV) ? g + h.momentum * b : b;
  }
  p = p + h.neg_lr * g;
}

// host side: the kernels' fp32 constants from the Python doubles of param_groups, the way torch derives them (double
// arithmetic, then one rounding to fp32)
inline void fill_adam_hp(const dgs_adam_range& r, AdamHp& h) {
  h.begin = r.begin;
  h.end = r.end;
  h.flags = r.flags;
  if (r.flags & DGS_ADAM_DECOUPLED_WD) h.decay = (float)(1.0 - r.lr * r.weight_decay);
  else h.decay = (float)r.weight_decay;
  h.w1 = (float)(1.0 - r.beta1);
  h.beta2 = (float)r.beta2;
  h.w2 = (float)(1.0 - r.beta2);
  h.eps = (float)r.eps;
  h.neg_step = (float)(-r.step_size);
  h.inv_bc2_sqrt = (float)(1.0 / r.bias_correction2_sqrt);
}

inline void fill_sgd_hp(const dgs_sgd_range& r, SgdHp& h) {
  h.begin = r.begin;
  h.end = r.end;
  h.flags = r.flags;
  h.pad_ = 0;
  h.wd = (float)r.weight_decay;
  h.momentum = (float)r.momentum;
  h.w_g = (float)(1.0 - r.dampening);
  h.neg_lr = (float)(-r.lr);
}

}  // namespace dgs
