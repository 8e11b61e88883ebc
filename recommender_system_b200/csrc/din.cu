// R9 DIN local-activation attention (DL/layer/interaction.py:369-406), exact-fp32 CUDA-core path.
// One CTA per sample: the T x K history (gathered straight from the table, or read from `key`)
// lives in shared memory for the whole unit, so neither the tiled query, the 4K-wide concat
// [q, k, q-k, q*k] nor the hidden activations ever touch HBM.  Masked softmax uses the reference's
// -2^32+1 fill (an all-padding history gives the uniform 1/T average, not zeros).
// The tcgen05 version of the same unit is in din_tc.cu.
#include "common.cuh"
#include "epilogue.cuh"

namespace ctrhot {

constexpr int kMaxAttLayers = 4;
struct AttLayers {
  const float* w[kMaxAttLayers];      // [in,out] or NULL = element-wise activation only (the 'dice' branch, N3)
  const float* b[kMaxAttLayers];
  Epilogue ep[kMaxAttLayers];
  int in_dim[kMaxAttLayers], out_dim[kMaxAttLayers];
  int n;
};

template <int IDKIND, bool GATHER>
__global__ void __launch_bounds__(256)
din_attention_kernel(const float* __restrict__ table, int64_t V, const void* __restrict__ hist_ids,
                     int64_t hist_ld, const void* __restrict__ cand_ids, int64_t cand_ld,
                     const float* __restrict__ query, int64_t q_ld, const float* __restrict__ key,
                     int64_t key_ld_b, const float* __restrict__ mask, int64_t B, int T, int K,
                     AttLayers L, const float* __restrict__ out_w, const float* __restrict__ out_b,
                     int hmax, float* __restrict__ out, int64_t out_ld, float* __restrict__ item_out,
                     int64_t item_ld, int* __restrict__ err_flag) {
  extern __shared__ float sm[];
  float* sK = sm;                       // [T][K]
  float* sQ = sK + (size_t)T * K;       // [K]
  float* sS = sQ + K;                   // [T] scores / probabilities
  float* sM = sS + T;                   // [T] mask
  float* bufA = sM + T;                 // [T][hmax]
  float* bufB = bufA + (size_t)T * hmax;
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    __syncthreads();
    // ---- stage q, keys, mask
    if (GATHER) {
      const int64_t cid = IdLoad<IDKIND>::at(cand_ids, b * cand_ld);
      const bool cok = cid >= 0 && cid < V;
      if (!cok && err_flag && tid == 0) *err_flag = 1;
      for (int c = tid; c < K; c += nt) {
        const float t = cok ? ldg_nc_f1(table + cid * K + c) : 0.f;
        sQ[c] = t;
        if (item_out) item_out[b * item_ld + c] = t;
      }
      for (int e = tid; e < T * K; e += nt) {
        const int t = e / K, c = e - t * K;
        const int64_t id = IdLoad<IDKIND>::at(hist_ids, b * hist_ld + t);
        const bool ok = id >= 0 && id < V;
        if (!ok && err_flag) *err_flag = 1;
        sK[e] = ok ? ldg_nc_f1(table + id * K + c) : 0.f;
        if (c == 0) sM[t] = (id != 0) ? 1.f : 0.f;      // DL/model/din.py:80
      }
    } else {
      for (int c = tid; c < K; c += nt) sQ[c] = __ldg(query + b * q_ld + c);
      for (int e = tid; e < T * K; e += nt) sK[e] = ldg_nc_f1(key + b * key_ld_b + e);
      for (int t = tid; t < T; t += nt) sM[t] = __ldg(mask + b * (int64_t)T + t);
    }
    __syncthreads();
    // ---- hidden layers; layer 0 reads the virtual concat [q, k, q-k, q*k]
    const float* in = nullptr;
    int in_w = 4 * K;
    float* bufs[2] = {bufA, bufB};
    int cur = 0;
    for (int l = 0; l < L.n; ++l) {
      const int N = L.out_dim[l];
      float* o = bufs[cur];
      if (L.w[l] == nullptr) {
        // element-wise activation over the current input (in-place when it is already a buffer)
        for (int e = tid; e < T * in_w; e += nt) {
          const int t = e / in_w, i = e - t * in_w;
          float a;
          if (in == nullptr) {
            const int seg = i / K, c = i - seg * K;
            const float q = sQ[c], kk = sK[t * K + c];
            a = seg == 0 ? q : (seg == 1 ? kk : (seg == 2 ? q - kk : q * kk));
          } else {
            a = in[(size_t)t * in_w + i];
          }
          o[(size_t)t * in_w + i] = L.ep[l].apply(a, t, i, in_w);
        }
        __syncthreads();
        if (in != nullptr && in != o) { /* wrote to the other buffer */ }
        in = o;
        cur ^= 1;
        continue;
      }
      const float* W = L.w[l];
      constexpr int TT = 4;
      const int groups = (T + TT - 1) / TT;
      for (int idx = tid; idx < groups * N; idx += nt) {
        const int g = idx / N, n = idx - g * N;
        const int t0 = g * TT;
        float acc[TT] = {0.f, 0.f, 0.f, 0.f};
        if (in == nullptr) {
          for (int c = 0; c < K; ++c) {
            const float q = sQ[c];
            const float w0 = __ldg(W + (size_t)c * N + n), w1 = __ldg(W + (size_t)(K + c) * N + n);
            const float w2 = __ldg(W + (size_t)(2 * K + c) * N + n), w3 = __ldg(W + (size_t)(3 * K + c) * N + n);
#pragma unroll
            for (int u = 0; u < TT; ++u) {
              const int t = min(t0 + u, T - 1);
              const float kk = sK[t * K + c];
              // same term order as the concat: q*W0 + k*W1 + (q-k)*W2 + (q*k)*W3
              acc[u] = fmaf(q * kk, w3, fmaf(q - kk, w2, fmaf(kk, w1, fmaf(q, w0, acc[u]))));
            }
          }
        } else {
          for (int i = 0; i < in_w; ++i) {
            const float w = __ldg(W + (size_t)i * N + n);
#pragma unroll
            for (int u = 0; u < TT; ++u) acc[u] = fmaf(in[(size_t)min(t0 + u, T - 1) * in_w + i], w, acc[u]);
          }
        }
        const float bb = L.b[l] ? __ldg(L.b[l] + n) : 0.f;
#pragma unroll
        for (int u = 0; u < TT; ++u)
          if (t0 + u < T) o[(size_t)(t0 + u) * N + n] = L.ep[l].apply(acc[u] + bb, t0 + u, n, N);
      }
      __syncthreads();
      in = o;
      in_w = N;
      cur ^= 1;
    }
    // ---- score = Dense(1), mask fill, softmax over T
    const int warp = tid >> 5, lane = tid & 31, nw = nt >> 5;
    for (int t = warp; t < T; t += nw) {
      float s = 0.f;
      for (int i = lane; i < in_w; i += 32) {
        float a;
        if (in == nullptr) {
          const int seg = i / K, c = i - seg * K;
          const float q = sQ[c], kk = sK[t * K + c];
          a = seg == 0 ? q : (seg == 1 ? kk : (seg == 2 ? q - kk : q * kk));
        } else {
          a = in[(size_t)t * in_w + i];
        }
        s = fmaf(a, __ldg(out_w + i), s);
      }
      s = warp_sum(s);
      if (lane == 0) sS[t] = (sM[t] == 0.f) ? -4294967295.0f : s + (out_b ? __ldg(out_b) : 0.f);
    }
    __syncthreads();
    if (warp == 0) {
      float mx = -INFINITY;
      for (int t = lane; t < T; t += 32) mx = fmaxf(mx, sS[t]);
      mx = warp_max(mx);
      float sum = 0.f;
      for (int t = lane; t < T; t += 32) { const float e = expf(sS[t] - mx); sS[t] = e; sum += e; }
      sum = warp_sum(sum);
      const float inv = 1.0f / sum;
      for (int t = lane; t < T; t += 32) sS[t] *= inv;
    }
    __syncthreads();
    // ---- out = p @ value  (value == key)
    for (int c = tid; c < K; c += nt) {
      float acc = 0.f;
      for (int t = 0; t < T; ++t) acc = fmaf(sS[t], sK[t * K + c], acc);
      out[b * out_ld + c] = acc;
    }
  }
}

static int build_layers(const ctrhot_dense_desc* h, int n, int K, AttLayers& L, int& hmax, int& last_w) {
  if (n < 0 || n > kMaxAttLayers) return CTRHOT_ESHAPE;
  L.n = n;
  int w = 4 * K;
  hmax = 1;
  for (int l = 0; l < n; ++l) {
    L.w[l] = h[l].weight; L.b[l] = h[l].bias; L.ep[l] = Epilogue::from(h[l]);
    if (h[l].weight == nullptr) {
      if (h[l].act == CTRHOT_ACT_NONE) return CTRHOT_EARG;
      L.in_dim[l] = w; L.out_dim[l] = w;
    } else {
      if (h[l].in_dim != w || h[l].out_dim <= 0) return CTRHOT_EARG;
      L.in_dim[l] = w; L.out_dim[l] = h[l].out_dim; w = h[l].out_dim;
    }
    if ((h[l].act == CTRHOT_ACT_PRELU || h[l].act == CTRHOT_ACT_DICE) && !h[l].act_param) return CTRHOT_EARG;
    if (w > hmax) hmax = w;
  }
  last_w = w;
  return CTRHOT_OK;
}

template <int IDKIND, bool GATHER>
static int launch_att(const float* table, int64_t V, const void* hist_ids, int64_t hist_ld,
                      const void* cand_ids, int64_t cand_ld, const float* query, int64_t q_ld,
                      const float* key, int64_t key_ld_b, const float* mask, int64_t B, int T, int K,
                      const ctrhot_dense_desc* h_layers, int n_layers, const float* out_w,
                      const float* out_b, float* out, int64_t out_ld, float* item_out, int64_t item_ld,
                      int* err_flag, void* stream) {
  AttLayers L;
  int hmax = 0, last_w = 0;
  int rc = build_layers(h_layers, n_layers, K, L, hmax, last_w);
  if (rc != CTRHOT_OK) return rc;
  const size_t smem = ((size_t)T * K + K + 2 * T + 2 * (size_t)T * hmax) * sizeof(float);
  if (smem > 220 * 1024) return CTRHOT_ESHAPE;
  auto kern = din_attention_kernel<IDKIND, GATHER>;
  CTRHOT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "din attr");
  int64_t grid = B < (int64_t)kNumSMs * 4 ? B : (int64_t)kNumSMs * 4;
  kern<<<(int)grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(
      table, V, hist_ids, hist_ld, cand_ids, cand_ld, query, q_ld, key, key_ld_b, mask, B, T, K, L,
      out_w, out_b, hmax, out, out_ld, item_out, item_ld, err_flag);
  CTRHOT_LAUNCH_OK("ctrhot_din_attention");
  return CTRHOT_OK;
}

}  // namespace ctrhot

using namespace ctrhot;

extern "C" int ctrhot_din_attention_f32(const float* query, int64_t q_ld, const float* key,
                                        int64_t key_ld_b, const float* mask, int64_t B, int T, int K,
                                        const ctrhot_dense_desc* h_layers, int n_layers,
                                        const float* out_w, const float* out_b, float* out,
                                        int64_t out_ld, void* stream) {
  CTRHOT_RANGE("ctrhot_din_attention_f32");
  if (!query || !key || !mask || !out || !out_w || B < 0 || T <= 0 || K <= 0) return CTRHOT_EARG;
  if (q_ld < K || key_ld_b < (int64_t)T * K || out_ld < K || (n_layers > 0 && !h_layers)) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  return launch_att<CTRHOT_IDS_I32, false>(nullptr, 0, nullptr, 0, nullptr, 0, query, q_ld, key, key_ld_b,
                                           mask, B, T, K, h_layers, n_layers, out_w, out_b, out, out_ld,
                                           nullptr, 0, nullptr, stream);
}

extern "C" int ctrhot_din_attention_gather_f32(const float* table, int64_t V, int K, const void* hist_ids,
                                               int ids_kind, int64_t hist_ld, const void* cand_ids,
                                               int64_t cand_ld, int64_t B, int T,
                                               const ctrhot_dense_desc* h_layers, int n_layers,
                                               const float* out_w, const float* out_b, float* att_out,
                                               int64_t att_ld, float* item_out, int64_t item_ld,
                                               int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_din_attention_gather_f32");
  if (!table || !hist_ids || !cand_ids || !att_out || !out_w || B < 0 || T <= 0 || K <= 0 || V <= 0)
    return CTRHOT_EARG;
  if (hist_ld < T || cand_ld < 1 || att_ld < K || (item_out && item_ld < K) || (n_layers > 0 && !h_layers))
    return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  if (ids_kind == CTRHOT_IDS_I32)
    return launch_att<CTRHOT_IDS_I32, true>(table, V, hist_ids, hist_ld, cand_ids, cand_ld, nullptr, 0, nullptr,
                                            0, nullptr, B, T, K, h_layers, n_layers, out_w, out_b, att_out,
                                            att_ld, item_out, item_ld, err_flag, stream);
  if (ids_kind == CTRHOT_IDS_F32)
    return launch_att<CTRHOT_IDS_F32, true>(table, V, hist_ids, hist_ld, cand_ids, cand_ld, nullptr, 0, nullptr,
                                            0, nullptr, B, T, K, h_layers, n_layers, out_w, out_b, att_out,
                                            att_ld, item_out, item_ld, err_flag, stream);
  return CTRHOT_EARG;
}
