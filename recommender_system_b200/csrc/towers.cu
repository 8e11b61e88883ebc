// what the remaining towers need on top of the Dense-stack kernels (mlp_tc.cu / fused.cu / mlp_simt.cu).
//   residual_relu : out = relu(x + y)                               ResLayer.call, DL/layer/interaction.py:277-278
//   mmoe_combine  : out_t[b, h] = sum_e softmax_e(g_t[b, :])[e] * expert[b, h, e]     mmoe_layer.call, interaction.py:492-509
//   lau_concat    : a[b, t, :] = [q_b, k_bt, q_b - k_bt, q_b * k_bt]    LocalActivationUnit.call, DL/layer/core.py:96-101
// All are one pass over their data (HBM-bound); the contractions around them run on the tensor pipe through DenseStack.
#include "common.cuh"

namespace ctrhot {

__global__ void residual_relu_kernel(const float* __restrict__ x, int64_t x_ld, const float* __restrict__ y, int64_t y_ld,
                                     int64_t M, int N, float* __restrict__ out, int64_t out_ld) {
  const int64_t total = M * N;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = i / N;
    const int n = (int)(i - m * N);
    out[m * out_ld + n] = fmaxf(x[m * x_ld + n] + y[m * y_ld + n], 0.f);
  }
}

// expert [B, H * E] (column h * E + e: the memory layout of the reference's [in, hidden, E] weight, so the experts are ONE
// Dense layer), gate [B, T * E] logits (bias added).  One thread per (b, h), all T tasks: the expert values are read once.
__global__ void mmoe_combine_kernel(const float* __restrict__ expert, int64_t e_ld, const float* __restrict__ gate, int64_t g_ld,
                                    int64_t B, int H, int E, int T, float* __restrict__ out, int64_t out_task_stride,
                                    int64_t out_ld) {
  const int64_t total = B * H;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / H;
    const int h = (int)(i - b * H);
    const float* ex = expert + b * e_ld + (int64_t)h * E;
    for (int t = 0; t < T; ++t) {
      const float* g = gate + b * g_ld + (int64_t)t * E;
      float mx = -INFINITY;
      for (int e = 0; e < E; ++e) mx = fmaxf(mx, g[e]);
      float den = 0.f, num = 0.f;
      for (int e = 0; e < E; ++e) {
        const float w = expf(g[e] - mx);
        den += w;
        num = fmaf(w, ex[e], num);
      }
      out[(int64_t)t * out_task_stride + b * out_ld + h] = num / den;
    }
  }
}

__global__ void lau_concat_kernel(const float* __restrict__ q, int64_t q_ld, const float* __restrict__ keys, int64_t k_ld,
                                  int64_t B, int T, int K, float* __restrict__ out, int64_t out_ld) {
  const int64_t total = B * T * K;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t bt = i / K;
    const int c = (int)(i - bt * K);
    const int64_t b = bt / T;
    const float qv = q[b * q_ld + c], kv = keys[bt * k_ld + c];
    float* o = out + bt * out_ld;
    o[c] = qv; o[K + c] = kv; o[2 * K + c] = qv - kv; o[3 * K + c] = qv * kv;
  }
}

static inline int grid1d(int64_t n, int block) {
  int64_t g = (n + block - 1) / block;
  const int64_t cap = (int64_t)kNumSMs * 16;
  return (int)(g > cap ? cap : (g < 1 ? 1 : g));
}

}  // namespace ctrhot

using namespace ctrhot;

extern "C" int ctrhot_residual_relu_f32(const float* x, int64_t x_ld, const float* y, int64_t y_ld, int64_t M, int N,
                                        float* out, int64_t out_ld, void* stream) {
  if (!x || !y || !out || M < 0 || N <= 0 || x_ld < N || y_ld < N || out_ld < N) return CTRHOT_EARG;
  if (M == 0) return CTRHOT_OK;
  residual_relu_kernel<<<grid1d(M * N, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, x_ld, y, y_ld, M, N, out, out_ld);
  CTRHOT_LAUNCH_OK("ctrhot_residual_relu_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_mmoe_combine_f32(const float* expert, int64_t e_ld, const float* gate, int64_t g_ld, int64_t B, int H,
                                       int E, int T, float* out, int64_t out_task_stride, int64_t out_ld, void* stream) {
  if (!expert || !gate || !out || B < 0 || H <= 0 || E <= 0 || T <= 0 || e_ld < (int64_t)H * E || g_ld < (int64_t)T * E ||
      out_ld < H) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  mmoe_combine_kernel<<<grid1d(B * H, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(expert, e_ld, gate, g_ld, B, H, E, T, out,
                                                                                       out_task_stride, out_ld);
  CTRHOT_LAUNCH_OK("ctrhot_mmoe_combine_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_lau_concat_f32(const float* query, int64_t q_ld, const float* keys, int64_t k_ld, int64_t B, int T, int K,
                                     float* out, int64_t out_ld, void* stream) {
  if (!query || !keys || !out || B < 0 || T <= 0 || K <= 0 || q_ld < K || k_ld < K || out_ld < 4 * (int64_t)K) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  lau_concat_kernel<<<grid1d(B * T * K, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(query, q_ld, keys, k_ld, B, T, K, out,
                                                                                          out_ld);
  CTRHOT_LAUNCH_OK("ctrhot_lau_concat_f32");
  return CTRHOT_OK;
}
