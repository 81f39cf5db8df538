// Per-element optimizer math shared by the single-GPU pass (optim.cu) and the fused
// reduce-scatter + optimizer + all-gather kernel (comm/rs_opt_ag.cu).
#pragma once
#include <cuda_runtime.h>

namespace zoo {

enum DenseOpt : int {
  OPT_SGD = 0,          // momentum / dampening / nesterov / weight decay (BigDL SGD)
  OPT_ADAM = 1,         // torch.optim.Adam: eps added after the bias-corrected sqrt
  OPT_ADAGRAD = 2,      // BigDL Adagrad (accumulated squared gradients)
  OPT_ADAM_ZOO = 3,     // zoo keras Adam (Adam.scala:59-106): lr*sqrt(bc2)/bc1 * m/(sqrt(v)+eps)
  OPT_ADAMW_BERT = 4,   // zoo AdamWeightDecay (AdamWeightDecay.scala:75-135): no bias correction
  OPT_FTRL = 5,         // BigDL Ftrl (TF FtrlOptimizer math); m = accumulator, v = linear
  OPT_RMSPROP = 6,      // BigDL RMSprop
  OPT_ADAMW = 7         // torch.optim.AdamW (decoupled weight decay, bias-corrected)
};

// hyper[] (device, fp32) slots
enum HyperSlot : int {
  H_LR = 0, H_B1 = 1, H_B2 = 2, H_EPS = 3, H_WD = 4, H_STEP = 5, H_GSCALE = 6,
  H_X1 = 7,   // sgd: dampening      ftrl: l1
  H_X2 = 8,   // sgd: nesterov flag  ftrl: l2
  H_X3 = 9,   // ftrl: lr_power (negative, e.g. -0.5)
  ZOO_HYPER_LEN = 16
};

struct OptHyper {
  int opt;
  float lr, b1, b2, eps, wd, step, gscale, x1, x2, x3;
  float bc1, bc2;  // bias corrections (Adam family)
};

__host__ __device__ inline bool opt_needs_m(int opt) { return opt != OPT_SGD; }  // SGD: m only with momentum
__host__ __device__ inline bool opt_needs_v(int opt) {
  return opt == OPT_ADAM || opt == OPT_ADAM_ZOO || opt == OPT_ADAMW_BERT || opt == OPT_FTRL || opt == OPT_ADAMW;
}

__device__ __forceinline__ OptHyper load_hyper(const float* __restrict__ hyper, int opt) {
  OptHyper h;
  h.opt = opt;
  h.lr = hyper[H_LR]; h.b1 = hyper[H_B1]; h.b2 = hyper[H_B2]; h.eps = hyper[H_EPS]; h.wd = hyper[H_WD];
  h.step = hyper[H_STEP]; h.gscale = hyper[H_GSCALE]; h.x1 = hyper[H_X1]; h.x2 = hyper[H_X2]; h.x3 = hyper[H_X3];
  h.bc1 = 1.f; h.bc2 = 1.f;
  if (opt == OPT_ADAM || opt == OPT_ADAM_ZOO || opt == OPT_ADAMW) {
    const float t = fmaxf(h.step, 1.f);
    h.bc1 = 1.f - powf(h.b1, t);
    h.bc2 = 1.f - powf(h.b2, t);
  }
  return h;
}

// p, m, v updated in place; g is the raw (summed) gradient, scaled by h.gscale here.
__device__ __forceinline__ void zoo_opt_update(float& p, float g, float& m, float& v, const OptHyper& h) {
  g *= h.gscale;
  switch (h.opt) {
    case OPT_ADAM: {
      g += h.wd * p;
      m = h.b1 * m + (1.f - h.b1) * g;
      v = h.b2 * v + (1.f - h.b2) * g * g;
      p -= h.lr * (m / h.bc1) / (sqrtf(v / h.bc2) + h.eps);
      break;
    }
    case OPT_ADAMW: {
      p -= h.lr * h.wd * p;
      m = h.b1 * m + (1.f - h.b1) * g;
      v = h.b2 * v + (1.f - h.b2) * g * g;
      p -= h.lr * (m / h.bc1) / (sqrtf(v / h.bc2) + h.eps);
      break;
    }
    case OPT_ADAM_ZOO: {
      m = h.b1 * m + (1.f - h.b1) * g;
      v = h.b2 * v + (1.f - h.b2) * g * g;
      p -= h.lr * sqrtf(h.bc2) / h.bc1 * m / (sqrtf(v) + h.eps);
      break;
    }
    case OPT_ADAMW_BERT: {
      m = h.b1 * m + (1.f - h.b1) * g;
      v = h.b2 * v + (1.f - h.b2) * g * g;
      p -= h.lr * (m / (sqrtf(v) + h.eps) + h.wd * p);
      break;
    }
    case OPT_ADAGRAD: {
      g += h.wd * p;
      m += g * g;
      p -= h.lr * g / (sqrtf(m) + h.eps);
      break;
    }
    case OPT_FTRL: {
      const float l1 = h.x1, l2 = h.x2, lrp = h.x3;
      const float acc_new = m + g * g;
      float sigma;
      if (lrp == -0.5f) sigma = (sqrtf(acc_new) - sqrtf(m)) / h.lr;
      else sigma = (powf(acc_new, -lrp) - powf(m, -lrp)) / h.lr;
      v += g - sigma * p;
      const float quad = (lrp == -0.5f ? sqrtf(acc_new) : powf(acc_new, -lrp)) / h.lr + 2.f * l2;
      const float av = fabsf(v);
      p = av > l1 ? (copysignf(l1, v) - v) / quad : 0.f;
      m = acc_new;
      break;
    }
    case OPT_RMSPROP: {
      g += h.wd * p;
      m = h.b2 * m + (1.f - h.b2) * g * g;
      p -= h.lr * g / (sqrtf(m) + h.eps);
      break;
    }
    default: {  // OPT_SGD
      g += h.wd * p;
      if (h.b1 != 0.f) {
        m = (h.step <= 1.f) ? g : h.b1 * m + (1.f - h.x1) * g;
        g = (h.x2 != 0.f) ? g + h.b1 * m : m;
      }
      p -= h.lr * g;
      break;
    }
  }
}

}  // namespace zoo
