// The six optimizers of the reference's factory (reference trainer/task.py:41-56) as one update rule
// per element, shared by the sparse (embedding / wide rows) and dense ps-side apply kernels and
// compiled for the host as well, so that the CPU tests can hold the device math against the NumPy
// restatement of TF-1 semantics in distributed_tensorflow_b200/optim.
//
//   id  name      slot 1 (init)          slot 2 (init)            TF-1 defaults
//   0   sgd       -                      -
//   1   adadelta  accum (0)              accum_update (0)         rho 0.95, eps 1e-8
//   2   adagrad   accumulator (0.1)      -
//   3   adam      m (0)                  v (0)                    beta1 0.9, beta2 0.999, eps 1e-8 (lr_t per step)
//   4   ftrl      accum (0.1)            linear (0)               lr_power -0.5, l1 = l2 = 0
//   5   rmsprop   ms (1.0)               mom (0)                  decay 0.9, momentum 0, eps 1e-10
#pragma once

#include <math.h>

#if defined(__CUDACC__)
#define DTF_HD __host__ __device__ __forceinline__
#else
#define DTF_HD inline
#endif

namespace dtf {

enum WdOptimizer : int { WD_SGD = 0, WD_ADADELTA = 1, WD_ADAGRAD = 2, WD_ADAM = 3, WD_FTRL = 4, WD_RMSPROP = 5 };  // ids shared with dtf_native.h

struct OptConst {
  float lr;       // SGD / Adagrad / ... step; for Adam: lr * sqrt(1 - b2^t) / (1 - b1^t)
  float beta1, beta2, eps;
};

constexpr bool opt_has_slot1(int opt) { return opt != WD_SGD; }
constexpr bool opt_has_slot2(int opt) { return opt == WD_ADAM || opt == WD_ADADELTA || opt == WD_FTRL || opt == WD_RMSPROP; }
constexpr float opt_slot1_init(int opt) { return (opt == WD_ADAGRAD || opt == WD_FTRL) ? 0.1f : (opt == WD_RMSPROP ? 1.0f : 0.0f); }

template <int OPT>
DTF_HD void opt_update(float& w, float g, float& s1, float& s2, const OptConst& c) {
  if (OPT == WD_SGD) {
    w -= c.lr * g;
  } else if (OPT == WD_ADAGRAD) {
    s1 += g * g;
#if defined(__CUDA_ARCH__)
    w -= c.lr * g * rsqrtf(s1);
#else
    w -= c.lr * g / sqrtf(s1);
#endif
  } else if (OPT == WD_ADAM) {
    s1 = c.beta1 * s1 + (1.f - c.beta1) * g;
    s2 = c.beta2 * s2 + (1.f - c.beta2) * g * g;
    w -= c.lr * s1 / (sqrtf(s2) + c.eps);
  } else if (OPT == WD_ADADELTA) {
    const float rho = 0.95f, eps = 1e-8f;
    s1 = rho * s1 + (1.f - rho) * g * g;
    const float u = sqrtf(s2 + eps) / sqrtf(s1 + eps) * g;
    s2 = rho * s2 + (1.f - rho) * u * u;
    w -= c.lr * u;
  } else if (OPT == WD_FTRL) {
    // lr_power = -0.5, l1 = l2 = 0: sigma = (sqrt(n') - sqrt(n)) / lr; z += g - sigma w; w = -z / (sqrt(n') / lr)
    const float n_new = s1 + g * g;
    const float sq_new = sqrtf(n_new);
    const float sigma = (sq_new - sqrtf(s1)) / c.lr;
    s2 += g - sigma * w;
    w = -s2 * c.lr / sq_new;
    s1 = n_new;
  } else {  // WD_RMSPROP, momentum 0
    const float decay = 0.9f, eps = 1e-10f;
    s1 = s1 + (g * g - s1) * (1.f - decay);
    s2 = c.lr * g / sqrtf(s1 + eps);   // mom = momentum * mom + lr g / sqrt(ms + eps), momentum = 0
    w -= s2;
  }
}

}  // namespace dtf
