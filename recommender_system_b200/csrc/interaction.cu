// K2-K5 stand-alone interaction layers (CUDA-core fp32): Wide, FM (dense / one-hot),
// the reference's FFM, PNN inner / outer product, DCN cross.  One warp per sample unless noted.
// All are HBM/L2-bound integer+fp32 work; no tensor cores (interaction outputs must stay
// within 1e-5 relative of the fp32 reference, which TF32 cannot give).
#include "common.cuh"

namespace ctrhot {

#define WARP_IDS()                                                                  \
  const int lane = threadIdx.x & 31;                                                \
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); \
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);

// ---------------------------------------------------------------- R2 WideLayer
__global__ void __launch_bounds__(256)
wide_dense_kernel(const float* __restrict__ x, int64_t x_ld, int64_t n, int64_t B,
                  const float* __restrict__ w0, const float* __restrict__ w, float* __restrict__ out) {
  WARP_IDS();
  for (int64_t b = warp; b < B; b += nwarps) {
    const float* row = x + b * x_ld;
    float acc = 0.f;
    for (int64_t j = lane; j < n; j += 32) {
      const float t = ldg_nc_f1(row + j);
      if (t != 0.f) acc = fmaf(t, __ldg(w + j), acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) out[b] = acc + __ldg(w0);
  }
}

template <int IDKIND>
__global__ void __launch_bounds__(256)
wide_onehot_kernel(const float* __restrict__ dense, int64_t dense_ld, int D,
                   const int32_t* __restrict__ dense_col, const void* __restrict__ ids,
                   int64_t ids_ld, int F, const int64_t* __restrict__ col_off, int64_t B,
                   const float* __restrict__ w0, const float* __restrict__ w, float* __restrict__ out) {
  WARP_IDS();
  for (int64_t b = warp; b < B; b += nwarps) {
    float acc = 0.f;
    for (int j = lane; j < D; j += 32)
      acc = fmaf(__ldg(dense + b * dense_ld + j), __ldg(w + __ldg(dense_col + j)), acc);
    for (int f = lane; f < F; f += 32) {
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      const int64_t c0 = col_off[f], c1 = col_off[f + 1];
      if (id >= 0 && id < c1 - c0) acc += ldg_nc_f1(w + c0 + id);
    }
    acc = warp_sum(acc);
    if (lane == 0) out[b] = acc + __ldg(w0);
  }
}

// ---------------------------------------------------------------- R3 FMLayer, dense x
// lane c < kfm owns factor c: s_c = sum_j x_j v[j,c], t_c = sum_j x_j^2 v[j,c]^2; lane kfm owns the
// linear term.  Non-zero columns are found with a ballot so a one-hot row costs ~nnz v-row reads.
__global__ void __launch_bounds__(256)
fm_dense_kernel(const float* __restrict__ x, int64_t x_ld, int64_t n, int64_t B,
                const float* __restrict__ w0, const float* __restrict__ w1,
                const float* __restrict__ v, int kfm, float* __restrict__ out, int apply_sigmoid) {
  WARP_IDS();
  for (int64_t b = warp; b < B; b += nwarps) {
    const float* row = x + b * x_ld;
    float s = 0.f, t = 0.f;
    for (int64_t j0 = 0; j0 < n; j0 += 32) {
      const int64_t j = j0 + lane;
      const float xj = (j < n) ? ldg_nc_f1(row + j) : 0.f;
      unsigned m = __ballot_sync(0xffffffffu, xj != 0.f);
      while (m) {
        const int src = __ffs(m) - 1;
        m &= m - 1;
        const float xs = __shfl_sync(0xffffffffu, xj, src);
        const int64_t js = j0 + src;
        if (lane < kfm) {
          const float vv = __ldg(v + js * kfm + lane);
          s = fmaf(xs, vv, s);
          t = fmaf(xs * xs, vv * vv, t);
        } else if (lane == kfm) {
          s = fmaf(xs, __ldg(w1 + js), s);
        }
      }
    }
    float r = (lane < kfm) ? 0.5f * (s * s - t) : (lane == kfm ? s : 0.f);
    r = warp_sum(r);
    if (lane == 0) {
      r += __ldg(w0);
      out[b] = apply_sigmoid ? 1.0f / (1.0f + expf(-r)) : r;
    }
  }
}

// ---------------------------------------------------------------- R3 FMLayer, one-hot as gather
template <int IDKIND>
__global__ void __launch_bounds__(256)
fm_onehot_kernel(const float* __restrict__ dense, int64_t dense_ld, int D,
                 const int32_t* __restrict__ dense_col, const void* __restrict__ ids,
                 int64_t ids_ld, int F, const int64_t* __restrict__ col_off, int64_t B,
                 const float* __restrict__ w0, const float* __restrict__ w1,
                 const float* __restrict__ v, int kfm, float* __restrict__ out, int apply_sigmoid) {
  WARP_IDS();
  for (int64_t b = warp; b < B; b += nwarps) {
    float s = 0.f, t = 0.f;
    for (int j = 0; j < D; ++j) {
      const float xs = __ldg(dense + b * dense_ld + j);
      if (xs == 0.f) continue;
      const int64_t js = __ldg(dense_col + j);
      if (lane < kfm) {
        const float vv = __ldg(v + js * kfm + lane);
        s = fmaf(xs, vv, s);
        t = fmaf(xs * xs, vv * vv, t);
      } else if (lane == kfm) {
        s = fmaf(xs, __ldg(w1 + js), s);
      }
    }
    for (int f = 0; f < F; ++f) {
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      const int64_t c0 = col_off[f], c1 = col_off[f + 1];
      if (id < 0 || id >= c1 - c0) continue;
      const int64_t js = c0 + id;
      if (lane < kfm) {
        const float vv = ldg_nc_f1(v + js * kfm + lane);
        s += vv;
        t = fmaf(vv, vv, t);
      } else if (lane == kfm) {
        s += ldg_nc_f1(w1 + js);
      }
    }
    float r = (lane < kfm) ? 0.5f * (s * s - t) : (lane == kfm ? s : 0.f);
    r = warp_sum(r);
    if (lane == 0) {
      r += __ldg(w0);
      out[b] = apply_sigmoid ? 1.0f / (1.0f + expf(-r)) : r;
    }
  }
}

// ---------------------------------------------------------------- R5 FFMLayer (reference formula, N4)
// field_f[i,:] = sum_j x_j v[j,i,:]  (a [field_num*k] vector: one gathered v-row per non-zero feature),
// inter = sum_{i<j} <field_f_i, field_f_j> = 0.5 (|sum_i field_f_i|^2 - sum_i |field_f_i|^2).
// Element e = i*k + c lives in lane e%32, slot e/32; k must divide 32 or be a multiple of 32 so
// that a lane's elements all share c = e % k (c group = lane % k for k<=32).
template <int IDKIND, int NS>
__global__ void __launch_bounds__(128)
ffm_kernel(const float* __restrict__ dense, int64_t dense_ld, int D, const void* __restrict__ ids,
           int64_t ids_ld, int F, const int64_t* __restrict__ col_off, int64_t B,
           const float* __restrict__ w0, const float* __restrict__ w, const float* __restrict__ v,
           int field_num, int k, float* __restrict__ out, int apply_sigmoid) {
  WARP_IDS();
  const int len = field_num * k;
  for (int64_t b = warp; b < B; b += nwarps) {
    float acc[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) acc[s] = 0.f;
    float lin = 0.f;
    for (int j = 0; j < D; ++j) {
      const float xs = __ldg(dense + b * dense_ld + j);
      if (xs == 0.f) continue;
      const float* vr = v + (size_t)j * len;
#pragma unroll
      for (int s = 0; s < NS; ++s) {
        const int e = lane + 32 * s;
        if (e < len) acc[s] = fmaf(xs, __ldg(vr + e), acc[s]);
      }
      if (lane == 0) lin = fmaf(xs, __ldg(w + j), lin);
    }
    for (int f = 0; f < F; ++f) {
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      const int64_t c0 = col_off[f], c1 = col_off[f + 1];
      if (id < 0 || id >= c1 - c0) continue;          // tf.one_hot: out-of-range -> zero row
      const float* vr = v + (size_t)(c0 + id) * len;
#pragma unroll
      for (int s = 0; s < NS; ++s) {
        const int e = lane + 32 * s;
        if (e < len) acc[s] += ldg_nc_f1(vr + e);
      }
      if (lane == 0) lin += ldg_nc_f1(w + c0 + id);
    }
    float sq = 0.f, tot = 0.f;
#pragma unroll
    for (int s = 0; s < NS; ++s) { sq = fmaf(acc[s], acc[s], sq); tot += acc[s]; }
    sq = warp_sum(sq);
    // tot currently sums this lane's elements (all share c = lane % k when k <= 32)
    float tsq;
    if (k <= 32) {
      for (int o = 16; o >= k; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
      tsq = (lane < k) ? tot * tot : 0.f;
      tsq = warp_sum(tsq);
    } else {
      // k multiple of 32: element e -> c = e % k = lane + 32*(s % (k/32)); regroup slots
      const int g = k >> 5;
      tsq = 0.f;
      for (int c = 0; c < g; ++c) {
        float tc = 0.f;
#pragma unroll
        for (int s = 0; s < NS; ++s) if ((s % g) == c) tc += acc[s];
        tsq = fmaf(tc, tc, tsq);
      }
      tsq = warp_sum(tsq);
    }
    if (lane == 0) {
      const float r = lin + __ldg(w0) + 0.5f * (tsq - sq);
      out[b] = apply_sigmoid ? 1.0f / (1.0f + expf(-r)) : r;
    }
  }
}

// ---------------------------------------------------------------- R6 InnerProductLayer
// One CTA per TS samples: E tile [TS][F*k] staged in smem, thread p computes <e_i, e_j>.
__global__ void __launch_bounds__(256)
pnn_inner_kernel(const float* __restrict__ e, int64_t e_ld, int F, int k, int64_t B,
                 float* __restrict__ out, int64_t out_ld) {
  extern __shared__ float sm[];
  const int P = F * (F - 1) / 2;
  const int fk = F * k;
  const int fkp = fk + 1;
  constexpr int TS = 8;
  int* prow = reinterpret_cast<int*>(sm + TS * fkp);
  int* pcol = prow + P;
  for (int i = threadIdx.x; i < F - 1; i += blockDim.x) {
    int base = i * (2 * F - i - 1) / 2;
    for (int j = i + 1; j < F; ++j) { prow[base + j - i - 1] = i; pcol[base + j - i - 1] = j; }
  }
  for (int64_t b0 = (int64_t)blockIdx.x * TS; b0 < B; b0 += (int64_t)gridDim.x * TS) {
    __syncthreads();
    for (int t = threadIdx.x; t < TS * fk; t += blockDim.x) {
      const int s = t / fk, c = t - s * fk;
      sm[s * fkp + c] = (b0 + s < B) ? ldg_nc_f1(e + (b0 + s) * e_ld + c) : 0.f;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < TS * P; t += blockDim.x) {
      const int s = t / P, p = t - s * P;
      if (b0 + s >= B) continue;
      const float* a = sm + s * fkp + prow[p] * k;
      const float* c = sm + s * fkp + pcol[p] * k;
      float acc = 0.f;
      for (int d = 0; d < k; ++d) acc = fmaf(a[d], c[d], acc);
      out[(b0 + s) * out_ld + p] = acc;
    }
  }
}

// ---------------------------------------------------------------- R7 OuterProductLayer
// out[b,p] = sum_a q[a] * (sum_c W[a,p,c] p_[c]),  p_=e_i, q=e_j.  CTA = (pair chunk, sample tile):
// W[:,p,:] for PC pairs staged in smem once and reused by TS samples.
__global__ void __launch_bounds__(256)
pnn_outer_kernel(const float* __restrict__ e, int64_t e_ld, int F, int k, int64_t B,
                 const float* __restrict__ W, float* __restrict__ out, int64_t out_ld, int PC, int TS) {
  extern __shared__ float sm[];
  const int P = F * (F - 1) / 2;
  const int fk = F * k;
  const int kk = k * k;
  float* sW = sm;                                  // [PC][k*k + 1]  (a-major: W[a,p,c] -> sW[pp][a*k+c])
  float* sE = sW + (size_t)PC * (kk + 1);          // [TS][fk + 1]
  const int p0 = blockIdx.x * PC;
  const int np = min(PC, P - p0);
  for (int t = threadIdx.x; t < np * kk; t += blockDim.x) {
    const int pp = t / kk, r = t - pp * kk;
    const int a = r / k, c = r - a * k;
    sW[(size_t)pp * (kk + 1) + r] = __ldg(W + ((size_t)a * P + (p0 + pp)) * k + c);
  }
  for (int64_t b0 = (int64_t)blockIdx.y * TS; b0 < B; b0 += (int64_t)gridDim.y * TS) {
    __syncthreads();
    for (int t = threadIdx.x; t < TS * fk; t += blockDim.x) {
      const int s = t / fk, c = t - s * fk;
      sE[s * (fk + 1) + c] = (b0 + s < B) ? ldg_nc_f1(e + (b0 + s) * e_ld + c) : 0.f;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < TS * np; t += blockDim.x) {
      const int s = t / np, pp = t - s * np;
      if (b0 + s >= B) continue;
      const int p = p0 + pp;
      // invert row-major pair index p -> (i,j)
      int i = 0, rem = p;
      while (rem >= F - 1 - i) { rem -= F - 1 - i; ++i; }
      const int j = i + 1 + rem;
      const float* pi = sE + s * (fk + 1) + i * k;
      const float* qj = sE + s * (fk + 1) + j * k;
      const float* w = sW + (size_t)pp * (kk + 1);
      float acc = 0.f;
      for (int a = 0; a < k; ++a) {
        float t2 = 0.f;
        for (int c = 0; c < k; ++c) t2 = fmaf(w[a * k + c], pi[c], t2);
        acc = fmaf(t2, qj[a], acc);
      }
      out[(b0 + s) * out_ld + p] = acc;
    }
  }
}

// ---------------------------------------------------------------- R8 CrossLayer (x already in HBM)
// One warp per sample, x0 / x_l held in a per-warp smem slab (any d); all L layers in one launch.
__global__ void __launch_bounds__(128)
cross_kernel(const float* __restrict__ x, int64_t x_ld, int d, int64_t B, int L,
             const float* __restrict__ w, const float* __restrict__ bia, float* __restrict__ out,
             int64_t out_ld, const float* __restrict__ final_w, float* __restrict__ dot_out) {
  extern __shared__ float sm[];
  WARP_IDS();
  float* x0 = sm + (size_t)(threadIdx.x >> 5) * 2 * d;
  float* xl = x0 + d;
  for (int64_t b = warp; b < B; b += nwarps) {
    for (int j = lane; j < d; j += 32) { const float t = ldg_nc_f1(x + b * x_ld + j); x0[j] = t; xl[j] = t; }
    __syncwarp();
    for (int l = 0; l < L; ++l) {
      const float* wl = w + (size_t)l * d;
      const float* bl = bia + (size_t)l * d;
      float s = 0.f;
      for (int j = lane; j < d; j += 32) s = fmaf(xl[j], __ldg(wl + j), s);
      s = warp_sum(s);
      for (int j = lane; j < d; j += 32) xl[j] = fmaf(x0[j], s, __ldg(bl + j)) + xl[j];
      __syncwarp();
    }
    if (out) for (int j = lane; j < d; j += 32) out[b * out_ld + j] = xl[j];
    if (final_w) {
      float s = 0.f;
      for (int j = lane; j < d; j += 32) s = fmaf(xl[j], __ldg(final_w + j), s);
      s = warp_sum(s);
      if (lane == 0) dot_out[b] = s;
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------- R10 Dice (elementwise)
__global__ void dice_kernel(const float* __restrict__ x, int64_t total, int h,
                            const float* __restrict__ alphas, const float* __restrict__ mean,
                            const float* __restrict__ var, float eps, float* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % h);
    const float t = x[i];
    const float p = 1.0f / (1.0f + expf(-(t - __ldg(mean + c)) / sqrtf(__ldg(var + c) + eps)));
    out[i] = __ldg(alphas + c) * (1.0f - p) * t + p * t;
  }
}

static inline int warp_grid(int64_t B, int warps_per_cta) {
  int64_t g = (B + warps_per_cta - 1) / warps_per_cta;
  const int64_t cap = (int64_t)kNumSMs * 16;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace ctrhot

using namespace ctrhot;
#define ST(s) static_cast<cudaStream_t>(s)
#define IDS_OK(kind) ((kind) == CTRHOT_IDS_I32 || (kind) == CTRHOT_IDS_F32)

extern "C" int ctrhot_wide_dense_f32(const float* x, int64_t x_ld, int64_t n, int64_t B,
                                     const float* w0, const float* w, float* out, void* stream) {
  if (!x || !w0 || !w || !out || n <= 0 || B < 0 || x_ld < n) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  wide_dense_kernel<<<warp_grid(B, 8), 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w, out);
  CTRHOT_LAUNCH_OK("ctrhot_wide_dense_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_wide_onehot_f32(const float* dense, int64_t dense_ld, int D,
                                      const int32_t* dense_col, const void* ids, int ids_kind,
                                      int64_t ids_ld, int F, const int64_t* col_off, int64_t B,
                                      const float* w0, const float* w, float* out, void* stream) {
  if (!w0 || !w || !out || B < 0 || D < 0 || F < 0 || !IDS_OK(ids_kind)) return CTRHOT_EARG;
  if ((D > 0 && (!dense || !dense_col)) || (F > 0 && (!ids || !col_off))) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  if (ids_kind == CTRHOT_IDS_I32)
    wide_onehot_kernel<CTRHOT_IDS_I32><<<warp_grid(B, 8), 256, 0, ST(stream)>>>(
        dense, dense_ld, D, dense_col, ids, ids_ld, F, col_off, B, w0, w, out);
  else
    wide_onehot_kernel<CTRHOT_IDS_F32><<<warp_grid(B, 8), 256, 0, ST(stream)>>>(
        dense, dense_ld, D, dense_col, ids, ids_ld, F, col_off, B, w0, w, out);
  CTRHOT_LAUNCH_OK("ctrhot_wide_onehot_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_fm_dense_f32(const float* x, int64_t x_ld, int64_t n, int64_t B,
                                   const float* w0, const float* w1, const float* v, int kfm,
                                   float* out, int apply_sigmoid, void* stream) {
  if (!x || !w0 || !w1 || !out || n <= 0 || B < 0 || x_ld < n || kfm < 0) return CTRHOT_EARG;
  if (kfm > 0 && !v) return CTRHOT_EARG;
  if (kfm > 31) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  fm_dense_kernel<<<warp_grid(B, 8), 256, 0, ST(stream)>>>(x, x_ld, n, B, w0, w1, v, kfm, out,
                                                          apply_sigmoid);
  CTRHOT_LAUNCH_OK("ctrhot_fm_dense_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_fm_onehot_f32(const float* dense, int64_t dense_ld, int D,
                                    const int32_t* dense_col, const void* ids, int ids_kind,
                                    int64_t ids_ld, int F, const int64_t* col_off, int64_t B,
                                    const float* w0, const float* w1, const float* v, int kfm,
                                    float* out, int apply_sigmoid, void* stream) {
  if (!w0 || !w1 || !out || B < 0 || D < 0 || F < 0 || kfm < 0 || !IDS_OK(ids_kind)) return CTRHOT_EARG;
  if ((D > 0 && (!dense || !dense_col)) || (F > 0 && (!ids || !col_off)) || (kfm > 0 && !v)) return CTRHOT_EARG;
  if (kfm > 31) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  if (ids_kind == CTRHOT_IDS_I32)
    fm_onehot_kernel<CTRHOT_IDS_I32><<<warp_grid(B, 8), 256, 0, ST(stream)>>>(
        dense, dense_ld, D, dense_col, ids, ids_ld, F, col_off, B, w0, w1, v, kfm, out, apply_sigmoid);
  else
    fm_onehot_kernel<CTRHOT_IDS_F32><<<warp_grid(B, 8), 256, 0, ST(stream)>>>(
        dense, dense_ld, D, dense_col, ids, ids_ld, F, col_off, B, w0, w1, v, kfm, out, apply_sigmoid);
  CTRHOT_LAUNCH_OK("ctrhot_fm_onehot_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_ffm_onehot_f32(const float* dense, int64_t dense_ld, int D, const void* ids,
                                     int ids_kind, int64_t ids_ld, int F, const int64_t* col_off,
                                     int64_t B, const float* w0, const float* w, const float* v,
                                     int field_num, int k, float* out, int apply_sigmoid, void* stream) {
  if (!w0 || !w || !v || !out || B < 0 || D < 0 || F < 0 || field_num <= 0 || k <= 0 || !IDS_OK(ids_kind))
    return CTRHOT_EARG;
  if ((D > 0 && !dense) || (F > 0 && (!ids || !col_off))) return CTRHOT_EARG;
  const bool kok = (k <= 32) ? (32 % k == 0) : (k % 32 == 0);
  const int len = field_num * k;
  if (!kok || len > 64 * 32) return CTRHOT_ESHAPE;
  if (B == 0) return CTRHOT_OK;
  const int ns = (len + 31) / 32;
  const int grid = warp_grid(B, 4);
#define L2(KIND, NS) ffm_kernel<KIND, NS><<<grid, 128, 0, ST(stream)>>>(dense, dense_ld, D, ids, ids_ld, F, \
      col_off, B, w0, w, v, field_num, k, out, apply_sigmoid)
#define L1(KIND) if (ns <= 4) L2(KIND, 4); else if (ns <= 8) L2(KIND, 8); else if (ns <= 16) L2(KIND, 16); \
      else if (ns <= 32) L2(KIND, 32); else L2(KIND, 64);
  if (ids_kind == CTRHOT_IDS_I32) { L1(CTRHOT_IDS_I32) } else { L1(CTRHOT_IDS_F32) }
#undef L1
#undef L2
  CTRHOT_LAUNCH_OK("ctrhot_ffm_onehot_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_pnn_inner_f32(const float* e, int64_t e_ld, int F, int k, int64_t B,
                                    float* out, int64_t out_ld, void* stream) {
  const int P = F * (F - 1) / 2;
  if (!e || !out || F < 2 || k <= 0 || B < 0 || e_ld < (int64_t)F * k || out_ld < P) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  const size_t smem = (size_t)8 * (F * k + 1) * sizeof(float) + (size_t)2 * P * sizeof(int);
  if (smem > 200 * 1024) return CTRHOT_ESHAPE;
  CTRHOT_CUDA_OK(cudaFuncSetAttribute(pnn_inner_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem), "pnn_inner attr");
  int64_t g = (B + 7) / 8;
  if (g > kNumSMs * 8) g = kNumSMs * 8;
  pnn_inner_kernel<<<(int)g, 256, smem, ST(stream)>>>(e, e_ld, F, k, B, out, out_ld);
  CTRHOT_LAUNCH_OK("ctrhot_pnn_inner_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_pnn_outer_f32(const float* e, int64_t e_ld, int F, int k, int64_t B,
                                    const float* W, float* out, int64_t out_ld, void* stream) {
  const int P = F * (F - 1) / 2;
  if (!e || !W || !out || F < 2 || k <= 0 || B < 0 || e_ld < (int64_t)F * k || out_ld < P) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  const int TS = 32;
  int PC = (int)((160 * 1024 / sizeof(float) - (size_t)TS * (F * k + 1)) / (k * k + 1));
  if (PC < 1) return CTRHOT_ESHAPE;
  if (PC > 32) PC = 32;
  if (PC > P) PC = P;
  const size_t smem = ((size_t)PC * (k * k + 1) + (size_t)TS * (F * k + 1)) * sizeof(float);
  CTRHOT_CUDA_OK(cudaFuncSetAttribute(pnn_outer_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem), "pnn_outer attr");
  const int gx = (P + PC - 1) / PC;
  int64_t gy = (B + TS - 1) / TS;
  const int64_t cap = (kNumSMs * 4 + gx - 1) / gx;
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  pnn_outer_kernel<<<dim3(gx, (unsigned)gy), 256, smem, ST(stream)>>>(e, e_ld, F, k, B, W, out, out_ld, PC, TS);
  CTRHOT_LAUNCH_OK("ctrhot_pnn_outer_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_cross_f32(const float* x, int64_t x_ld, int d, int64_t B, int L,
                                const float* w, const float* b, float* out, int64_t out_ld,
                                const float* final_w, float* dot, void* stream) {
  if (!x || d <= 0 || B < 0 || L < 0 || x_ld < d || (L > 0 && (!w || !b))) return CTRHOT_EARG;
  if (!out && !final_w) return CTRHOT_EARG;
  if ((out && out_ld < d) || (final_w && !dot)) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  const size_t smem = (size_t)4 * 2 * d * sizeof(float);
  if (smem > 220 * 1024) return CTRHOT_ESHAPE;
  CTRHOT_CUDA_OK(cudaFuncSetAttribute(cross_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem), "cross attr");
  cross_kernel<<<warp_grid(B, 4), 128, smem, ST(stream)>>>(x, x_ld, d, B, L, w, b, out, out_ld, final_w, dot);
  CTRHOT_LAUNCH_OK("ctrhot_cross_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_dice_f32(const float* x, int64_t rows, int h, const float* alphas,
                               const float* mean, const float* var, float eps, float* out, void* stream) {
  if (!x || !alphas || !mean || !var || !out || rows < 0 || h <= 0) return CTRHOT_EARG;
  const int64_t total = rows * h;
  if (total == 0) return CTRHOT_OK;
  int64_t g = (total + 255) / 256;
  if (g > kNumSMs * 16) g = kNumSMs * 16;
  dice_kernel<<<(int)g, 256, 0, ST(stream)>>>(x, total, h, alphas, mean, var, eps, out);
  CTRHOT_LAUNCH_OK("ctrhot_dice_f32");
  return CTRHOT_OK;
}
