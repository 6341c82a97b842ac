// CPU optimizer apply kernels with TensorFlow-1.x semantics and defaults (use_locking=False):
// the parameter-server side of optimizer.minimize() in reference trainer/task.py:41-53,138.
// Formulas are restated from SURVEY.md section 2B-B7; slot naming lives in ps.cc.
#include <math.h>

#include "../../include/dtf_native.h"

// The three optimizers that see large tensors get an AVX2 clone next to the baseline (SSE) build; glibc's
// ifunc resolver picks one at load time, so the same .so runs on any x86-64.  sqrt and divide stay the
// IEEE instructions (no rsqrt approximations, no FMA contraction), i.e. results are bit-identical to the
// scalar loops.  The names start with hogwild_apply_ so that csrc/native/tsan.supp can name them.
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__) && !defined(__SANITIZE_THREAD__)
#define DTFN_CLONES __attribute__((target_clones("avx2", "default"), optimize("O3", "no-math-errno", "no-trapping-math")))
#else
#define DTFN_CLONES
#endif

DTFN_CLONES static void hogwild_apply_sgd(float* __restrict__ var, const float* __restrict__ g, size_t n, float lr) {
  for (size_t i = 0; i < n; ++i) var[i] -= lr * g[i];
}

DTFN_CLONES static void hogwild_apply_adagrad(float* __restrict__ var, float* __restrict__ accum,
                                              const float* __restrict__ g, size_t n, float lr) {
  for (size_t i = 0; i < n; ++i) {
    const float a = accum[i] + g[i] * g[i];
    accum[i] = a;
    var[i] -= lr * g[i] / sqrtf(a);
  }
}

DTFN_CLONES static void hogwild_apply_adam(float* __restrict__ var, float* __restrict__ m, float* __restrict__ v,
                                           const float* __restrict__ g, size_t n, float lr_t, float beta1, float beta2,
                                           float eps) {
  for (size_t i = 0; i < n; ++i) {
    const float mi = beta1 * m[i] + (1.f - beta1) * g[i];
    const float vi = beta2 * v[i] + (1.f - beta2) * g[i] * g[i];
    m[i] = mi;
    v[i] = vi;
    var[i] -= lr_t * mi / (sqrtf(vi) + eps);
  }
}

extern "C" {

void dtfn_apply_sgd(float* var, const float* g, size_t n, float lr) { hogwild_apply_sgd(var, g, n, lr); }

void dtfn_apply_adagrad(float* var, float* accum, const float* g, size_t n, float lr) {
  hogwild_apply_adagrad(var, accum, g, n, lr);
}

void dtfn_apply_adam(float* var, float* m, float* v, const float* g, size_t n, float lr, float beta1, float beta2,
                     float eps, float beta1_power, float beta2_power) {
  const float lr_t = lr * sqrtf(1.f - beta2_power) / (1.f - beta1_power);
  hogwild_apply_adam(var, m, v, g, n, lr_t, beta1, beta2, eps);
}

void dtfn_apply_adadelta(float* var, float* accum, float* accum_update, const float* g, size_t n, float lr,
                         float rho, float eps) {
  for (size_t i = 0; i < n; ++i) {
    const float a = rho * accum[i] + (1.f - rho) * g[i] * g[i];
    const float u = sqrtf(accum_update[i] + eps) / sqrtf(a + eps) * g[i];
    accum[i] = a;
    accum_update[i] = rho * accum_update[i] + (1.f - rho) * u * u;
    var[i] -= lr * u;
  }
}

void dtfn_apply_ftrl(float* var, float* accum, float* linear, const float* g, size_t n, float lr, float l1,
                     float l2, float lr_power) {
  for (size_t i = 0; i < n; ++i) {
    const float na = accum[i] + g[i] * g[i];
    float sigma, quad;
    if (lr_power == -0.5f) {
      sigma = (sqrtf(na) - sqrtf(accum[i])) / lr;
      quad = sqrtf(na) / lr + 2.f * l2;
    } else {
      sigma = (powf(na, -lr_power) - powf(accum[i], -lr_power)) / lr;
      quad = powf(na, -lr_power) / lr + 2.f * l2;
    }
    const float z = linear[i] + g[i] - sigma * var[i];
    linear[i] = z;
    const float sgn = z > 0.f ? 1.f : (z < 0.f ? -1.f : 0.f);
    var[i] = fabsf(z) > l1 ? (sgn * l1 - z) / quad : 0.f;
    accum[i] = na;
  }
}

void dtfn_apply_rmsprop(float* var, float* ms, float* mom, const float* g, size_t n, float lr, float decay,
                        float momentum, float eps) {
  for (size_t i = 0; i < n; ++i) {
    const float s = ms[i] + (g[i] * g[i] - ms[i]) * (1.f - decay);
    ms[i] = s;
    const float mo = mom[i] * momentum + lr * g[i] / sqrtf(s + eps);
    mom[i] = mo;
    var[i] -= mo;
  }
}

}  // extern "C"
