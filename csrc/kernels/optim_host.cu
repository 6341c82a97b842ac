// Host build of the per-element optimizer rule the ps-side apply kernels use (wd_optim.cuh), exported so
// that the CPU test-suite can compare it with the NumPy optimizers without a GPU
// (tests/test_optimizers.py).  No device code here.
#include "../../include/dtf_kernels.h"
#include "wd_optim.cuh"

using namespace dtf;

template <int OPT>
static void run(float* w, const float* g, float* s1, float* s2, long long n, const OptConst& c) {
  float d1 = 0.f, d2 = 0.f;
  for (long long i = 0; i < n; ++i) opt_update<OPT>(w[i], g[i], s1 ? s1[i] : d1, s2 ? s2[i] : d2, c);
}

extern "C" int dtf_opt_update_host(int optimizer, float* w, const float* g, float* slot1, float* slot2, long long n,
                                   float lr, long long step) {
  OptConst c;
  c.beta1 = 0.9f; c.beta2 = 0.999f; c.eps = 1e-8f; c.lr = lr;
  if (optimizer == WD_ADAM) {
    const double b1t = pow(0.9, static_cast<double>(step)), b2t = pow(0.999, static_cast<double>(step));
    c.lr = static_cast<float>(lr * sqrt(1.0 - b2t) / (1.0 - b1t));
  }
  if ((opt_has_slot1(optimizer) && !slot1) || (opt_has_slot2(optimizer) && !slot2)) return -2;
  switch (optimizer) {
    case WD_SGD: run<WD_SGD>(w, g, slot1, slot2, n, c); break;
    case WD_ADADELTA: run<WD_ADADELTA>(w, g, slot1, slot2, n, c); break;
    case WD_ADAGRAD: run<WD_ADAGRAD>(w, g, slot1, slot2, n, c); break;
    case WD_ADAM: run<WD_ADAM>(w, g, slot1, slot2, n, c); break;
    case WD_FTRL: run<WD_FTRL>(w, g, slot1, slot2, n, c); break;
    case WD_RMSPROP: run<WD_RMSPROP>(w, g, slot1, slot2, n, c); break;
    default: return -2;
  }
  return 0;
}

extern "C" float dtf_opt_slot1_init(int optimizer) { return opt_slot1_init(optimizer); }
extern "C" int dtf_opt_num_slots(int optimizer) { return (opt_has_slot1(optimizer) ? 1 : 0) + (opt_has_slot2(optimizer) ? 1 : 0); }
