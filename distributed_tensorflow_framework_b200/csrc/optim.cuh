// Shared optimizer update rule (used by the flat optimizer kernel and by the fused collective kernels).
#pragma once
#include <cuda_runtime.h>

namespace dtf {

// kind: 0 = SGD, 1 = momentum (v = mu*v + g; p -= lr*v, TF MomentumOptimizer), 2 = Adam (TF AdamOptimizer form,
// lr_t = lr*sqrt(1-b2^t)/(1-b1^t) computed on the host). wd adds wd*p to the gradient.
__device__ __forceinline__ float optim_update(int kind, float p, float g, float* m, float* v, long long i, float lr,
                                              float b1, float b2, float eps, float wd, float lr_t) {
  g = fmaf(wd, p, g);
  if (kind == 0) return p - lr * g;
  if (kind == 1) {
    const float vel = fmaf(b1, m[i], g);
    m[i] = vel;
    return p - lr * vel;
  }
  const float mm = fmaf(b1, m[i], (1.f - b1) * g);
  const float vv = fmaf(b2, v[i], (1.f - b2) * g * g);
  m[i] = mm;
  v[i] = vv;
  return p - lr_t * mm / (sqrtf(vv) + eps);
}

struct OptimArgs {
  int kind;
  float lr, b1, b2, eps, wd, gscale, lr_t;
};

// Device-resident hyper-parameters (optional): hyper[0] = learning rate (written by the host whenever the schedule
// moves it), hyper[1] = Adam's bias-corrected rate lr_t (recomputed on the device every step by hyper_step_kernel),
// word 2 = the step counter as int32. A kernel launched from a replayed CUDA graph reads them at run time, so a
// learning-rate schedule (LearningRate handler, /root/reference/distribute_learningrate.py:12-50) and Adam's bias
// correction follow the real step instead of the values frozen at capture.
__device__ __forceinline__ void optim_load_hyper(OptimArgs& o, const float* __restrict__ hyper) {
  if (hyper != nullptr) {
    o.lr = hyper[0];
    o.lr_t = hyper[1];
  }
}

}  // namespace dtf
