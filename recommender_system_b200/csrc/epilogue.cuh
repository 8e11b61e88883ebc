// Dense-layer epilogue shared by the CUDA-core and tcgen05 towers:
// none / relu / PReLU(alpha[row % alpha_rows, n]) / Dice / sigmoid / tanh.
// PReLU follows keras.layers.PReLU.call (relu(x) - alpha*relu(-x)); alpha_rows = T reproduces the
// [T,h]-shaped alpha Keras builds for Dense(h, activation=PReLU()) on a [B,T,.] tensor
// (DL/layer/interaction.py:366).  Dice follows DL/layer/interaction.py:422-425.
#pragma once
#include "common.cuh"

namespace ctrhot {

struct Epilogue {
  const float* p;     // PRELU: alpha[alpha_rows][N]; DICE: [3][N] = alphas, mean, var
  int act;
  int alpha_rows;
  float eps;

  static Epilogue from(const ctrhot_dense_desc& d) {
    Epilogue e;
    e.p = d.act_param;
    e.act = d.act;
    e.alpha_rows = d.alpha_rows > 0 ? d.alpha_rows : 1;
    e.eps = d.dice_eps;
    return e;
  }

  __device__ __forceinline__ float apply(float t, int64_t row, int n, int N) const {
    switch (act) {
      case CTRHOT_ACT_RELU: return fmaxf(t, 0.f);
      case CTRHOT_ACT_PRELU: {
        const float a = __ldg(p + (size_t)(row % alpha_rows) * N + n);
        return fmaxf(t, 0.f) - a * fmaxf(-t, 0.f);
      }
      case CTRHOT_ACT_DICE: {
        const float a = __ldg(p + n), mu = __ldg(p + N + n), var = __ldg(p + 2 * N + n);
        const float pr = 1.0f / (1.0f + expf(-(t - mu) / sqrtf(var + eps)));
        return a * (1.0f - pr) * t + pr * t;
      }
      case CTRHOT_ACT_SIGMOID: return 1.0f / (1.0f + expf(-t));
      case CTRHOT_ACT_TANH: return tanhf(t);
      default: return t;
    }
  }
};

}  // namespace ctrhot
