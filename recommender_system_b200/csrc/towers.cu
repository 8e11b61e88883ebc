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
  CTRHOT_RANGE("ctrhot_residual_relu_f32");
  if (!x || !y || !out || M < 0 || N <= 0 || x_ld < N || y_ld < N || out_ld < N) return CTRHOT_EARG;
  if (M == 0) return CTRHOT_OK;
  residual_relu_kernel<<<grid1d(M * N, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, x_ld, y, y_ld, M, N, out, out_ld);
  CTRHOT_LAUNCH_OK("ctrhot_residual_relu_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_mmoe_combine_f32(const float* expert, int64_t e_ld, const float* gate, int64_t g_ld, int64_t B, int H,
                                       int E, int T, float* out, int64_t out_task_stride, int64_t out_ld, void* stream) {
  CTRHOT_RANGE("ctrhot_mmoe_combine_f32");
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
  CTRHOT_RANGE("ctrhot_lau_concat_f32");
  if (!query || !keys || !out || B < 0 || T <= 0 || K <= 0 || q_ld < K || k_ld < K || out_ld < 4 * (int64_t)K) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  lau_concat_kernel<<<grid1d(B * T * K, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(query, q_ld, keys, k_ld, B, T, K, out,
                                                                                          out_ld);
  CTRHOT_LAUNCH_OK("ctrhot_lau_concat_f32");
  return CTRHOT_OK;
}

// ---- AttentionSequencePoolingLayer.call, DL/layer/sequence.py:242-276 (the part after the LocalActivationUnit) --------------
// s_t = valid(t) ? score[b, t] : (weight_normalization ? -2^32 + 1 : 0);  p = weight_normalization ? softmax_T(s) : s;
// return_score ? out[b, t] = p_t : out[b, e] = sum_t p_t keys[b, t, e].  valid(t) = mask[b, t] != 0 (keras mask) or t < lengths[b]
// (tf.sequence_mask of keys_length).  One CTA per sample.
namespace ctrhot {
__global__ void __launch_bounds__(128)
seq_attn_pool_kernel(const float* __restrict__ scores, int64_t score_ld, const float* __restrict__ keys, int64_t key_ld_b,
                     const uint8_t* __restrict__ mask, int64_t mask_ld, const int32_t* __restrict__ lengths, int T, int E,
                     int wn, int ret, float* __restrict__ out, int64_t out_ld) {
  extern __shared__ float sp[];                              // [T] weights
  __shared__ float red[4];
  const int64_t b = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const float pad = wn ? -4294967295.0f : 0.f;
  float m = -INFINITY;
  for (int t = tid; t < T; t += 128) {
    const bool ok = mask ? mask[b * mask_ld + t] != 0 : t < lengths[b];
    const float s = ok ? scores[b * score_ld + t] : pad;
    sp[t] = s;
    m = fmaxf(m, s);
  }
  if (wn) {
    m = warp_max(m);
    if (lane == 0) red[w] = m;
    __syncthreads();
    m = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
    __syncthreads();
    float sum = 0.f;
    for (int t = tid; t < T; t += 128) { const float e = __expf(sp[t] - m); sp[t] = e; sum += e; }
    sum = warp_sum(sum);
    if (lane == 0) red[w] = sum;
    __syncthreads();
    const float inv = 1.f / ((red[0] + red[1]) + (red[2] + red[3]));
    for (int t = tid; t < T; t += 128) sp[t] *= inv;
  }
  __syncthreads();
  if (ret) {
    for (int t = tid; t < T; t += 128) out[b * out_ld + t] = sp[t];
    return;
  }
  const float* kb = keys + b * key_ld_b;
  for (int e = tid; e < E; e += 128) {
    float a0 = 0.f, a1 = 0.f;
    int t = 0;
    for (; t + 1 < T; t += 2) { a0 = fmaf(sp[t], kb[(int64_t)t * E + e], a0); a1 = fmaf(sp[t + 1], kb[(int64_t)(t + 1) * E + e], a1); }
    if (t < T) a0 = fmaf(sp[t], kb[(int64_t)t * E + e], a0);
    out[b * out_ld + e] = a0 + a1;
  }
}
}  // namespace ctrhot

extern "C" int ctrhot_seq_attn_pool_f32(const float* scores, int64_t score_ld, const float* keys, int64_t key_ld_b,
                                        const uint8_t* mask, int64_t mask_ld, const int32_t* lengths, int64_t B, int T, int E,
                                        int weight_normalization, int return_score, float* out, int64_t out_ld, void* stream) {
  CTRHOT_RANGE("ctrhot_seq_attn_pool_f32");
  if (!scores || !keys || !out || (!mask && !lengths) || B < 0 || T < 1 || E < 1 || score_ld < T || key_ld_b < (int64_t)T * E ||
      (mask && mask_ld < T) || out_ld < (return_score ? T : E)) return CTRHOT_EARG;
  if ((size_t)T * sizeof(float) > 200u * 1024u) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  const size_t smem = (size_t)T * sizeof(float);
  if (smem > 48u * 1024u)
    CTRHOT_CUDA_OK(cudaFuncSetAttribute(ctrhot::seq_attn_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "seq_attn_pool attr");
  ctrhot::seq_attn_pool_kernel<<<(unsigned)B, 128, smem, static_cast<cudaStream_t>(stream)>>>(
      scores, score_ld, keys, key_ld_b, mask, mask_ld, lengths, T, E, weight_normalization, return_score, out, out_ld);
  CTRHOT_LAUNCH_OK("ctrhot_seq_attn_pool_f32");
  return CTRHOT_OK;
}

// ---- AUGRU: AUGRUCell.call (DL/layer/activation.py:113-142) unrolled over time as rnn_augru does without a mask (:237-244) ----
//   z = ra(x Wz + h Uz) * att_t;  r = ra(x Wr + h Ur);  hh = act(x Wh + (r * h) Uh);  h' = z h + (1 - z) hh      (no biases)
// Sequential in T, parallel over samples: a CTA of 3H threads (thread = gate column) carries AS samples through all T steps,
// weights [E + H, 3H] resident in shared memory, 8 FMAs per weight read.
namespace ctrhot {
constexpr int AS = 8;
__device__ __forceinline__ float augru_ra(float v, int hard) {
  return hard ? fminf(fmaxf(0.2f * v + 0.5f, 0.f), 1.f) : 1.0f / (1.0f + __expf(-v));
}
__global__ void augru_kernel(const float* __restrict__ x, int64_t x_ld_b, const float* __restrict__ att, int64_t att_ld,
                             const float* __restrict__ Wk, const float* __restrict__ Uk, const float* __restrict__ h0,
                             int64_t B, int T, int E, int H, int ra_hard, int act_linear, float* __restrict__ seq_out,
                             float* __restrict__ last_out) {
  extern __shared__ float sm[];
  const int G = 3 * H, j = threadIdx.x;
  float* sW = sm;                       // [E][3H]
  float* sU = sW + (size_t)E * G;       // [H][3H]
  float* xs = sm + (((size_t)(E + H) * G + 3) & ~(size_t)3);   // [E][AS], 16-byte aligned for the float4 reads
  float* hs = xs + E * AS;              // [H][AS]
  float* rh = hs + H * AS;              // [H][AS]
  float* zs = rh + H * AS;              // [H][AS]
  for (int i = j; i < E * G; i += G) sW[i] = __ldg(Wk + i);
  for (int i = j; i < H * G; i += G) sU[i] = __ldg(Uk + i);
  for (int64_t b0 = (int64_t)blockIdx.x * AS; b0 < B; b0 += (int64_t)gridDim.x * AS) {
    const int ns = (int)((B - b0 < AS) ? (B - b0) : AS);
    __syncthreads();
    for (int i = j; i < H * AS; i += G) {
      const int h = i / AS, s = i - h * AS;
      hs[i] = (h0 && s < ns) ? h0[(b0 + s) * H + h] : 0.f;
    }
    for (int t = 0; t < T; ++t) {
      for (int i = j; i < E * AS; i += G) {
        const int s = i / E, e = i - s * E;                        // coalesced over e
        xs[e * AS + s] = (s < ns) ? x[(b0 + s) * x_ld_b + (int64_t)t * E + e] : 0.f;
      }
      __syncthreads();
      float acc[AS];
#pragma unroll
      for (int s = 0; s < AS; ++s) acc[s] = 0.f;
      for (int e = 0; e < E; ++e) {
        const float w = sW[e * G + j];
        const float4 a = *reinterpret_cast<const float4*>(xs + e * AS), b = *reinterpret_cast<const float4*>(xs + e * AS + 4);
        acc[0] = fmaf(a.x, w, acc[0]); acc[1] = fmaf(a.y, w, acc[1]); acc[2] = fmaf(a.z, w, acc[2]); acc[3] = fmaf(a.w, w, acc[3]);
        acc[4] = fmaf(b.x, w, acc[4]); acc[5] = fmaf(b.y, w, acc[5]); acc[6] = fmaf(b.z, w, acc[6]); acc[7] = fmaf(b.w, w, acc[7]);
      }
      if (j < 2 * H) {
        for (int h = 0; h < H; ++h) {
          const float w = sU[h * G + j];
          const float4 a = *reinterpret_cast<const float4*>(hs + h * AS), b = *reinterpret_cast<const float4*>(hs + h * AS + 4);
          acc[0] = fmaf(a.x, w, acc[0]); acc[1] = fmaf(a.y, w, acc[1]); acc[2] = fmaf(a.z, w, acc[2]); acc[3] = fmaf(a.w, w, acc[3]);
          acc[4] = fmaf(b.x, w, acc[4]); acc[5] = fmaf(b.y, w, acc[5]); acc[6] = fmaf(b.z, w, acc[6]); acc[7] = fmaf(b.w, w, acc[7]);
        }
        if (j < H) {
#pragma unroll
          for (int s = 0; s < AS; ++s)
            zs[j * AS + s] = augru_ra(acc[s], ra_hard) * (s < ns ? att[(b0 + s) * att_ld + t] : 0.f);
        } else {
#pragma unroll
          for (int s = 0; s < AS; ++s) rh[(j - H) * AS + s] = augru_ra(acc[s], ra_hard) * hs[(j - H) * AS + s];
        }
      }
      __syncthreads();
      if (j >= 2 * H) {
        const int u = j - 2 * H;
        for (int h = 0; h < H; ++h) {
          const float w = sU[h * G + j];
          const float4 a = *reinterpret_cast<const float4*>(rh + h * AS), b = *reinterpret_cast<const float4*>(rh + h * AS + 4);
          acc[0] = fmaf(a.x, w, acc[0]); acc[1] = fmaf(a.y, w, acc[1]); acc[2] = fmaf(a.z, w, acc[2]); acc[3] = fmaf(a.w, w, acc[3]);
          acc[4] = fmaf(b.x, w, acc[4]); acc[5] = fmaf(b.y, w, acc[5]); acc[6] = fmaf(b.z, w, acc[6]); acc[7] = fmaf(b.w, w, acc[7]);
        }
#pragma unroll
        for (int s = 0; s < AS; ++s) {
          const float hh = act_linear ? acc[s] : tanhf(acc[s]);
          const float z = zs[u * AS + s];
          const float hn = z * hs[u * AS + s] + (1.f - z) * hh;
          hs[u * AS + s] = hn;
          if (s < ns) {
            if (seq_out) seq_out[((b0 + s) * T + t) * H + u] = hn;
            if (last_out && t == T - 1) last_out[(b0 + s) * H + u] = hn;
          }
        }
      }
      __syncthreads();
    }
  }
}
}  // namespace ctrhot

extern "C" int ctrhot_augru_f32(const float* x, int64_t x_ld_b, const float* att, int64_t att_ld, const float* kernel,
                                const float* recurrent_kernel, const float* h0, int64_t B, int T, int E, int H,
                                int recurrent_hard_sigmoid, int linear_activation, float* seq_out, float* last_out,
                                void* stream) {
  CTRHOT_RANGE("ctrhot_augru_f32");
  if (!x || !att || !kernel || !recurrent_kernel || (!seq_out && !last_out) || B < 0 || T < 1 || E < 1 || H < 1 ||
      x_ld_b < (int64_t)T * E || att_ld < T) return CTRHOT_EARG;
  if (3 * H > 1024) return CTRHOT_ESHAPE;
  const size_t smem = ((((size_t)(E + H) * 3 * H + 3) & ~(size_t)3) + (size_t)(E + 3 * H) * ctrhot::AS) * sizeof(float);
  if (smem > 227u * 1024u) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  CTRHOT_CUDA_OK(cudaFuncSetAttribute(ctrhot::augru_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "augru attr");
  int64_t groups = (B + ctrhot::AS - 1) / ctrhot::AS;
  int64_t grid = groups < (int64_t)ctrhot::kNumSMs * 2 ? groups : (int64_t)ctrhot::kNumSMs * 2;
  ctrhot::augru_kernel<<<(int)grid, 3 * H, smem, static_cast<cudaStream_t>(stream)>>>(
      x, x_ld_b, att, att_ld, kernel, recurrent_kernel, h0, B, T, E, H, recurrent_hard_sigmoid, linear_activation, seq_out,
      last_out);
  CTRHOT_LAUNCH_OK("ctrhot_augru_f32");
  return CTRHOT_OK;
}
