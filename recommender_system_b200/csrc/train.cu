// the step on the RIGHT of the hot path -- one SGD step of DeepFM, so that compile_fit
// (DL/utils/compile_fit.py:9-15: SGD(lr), loss 'binary_crossentropy', model.fit over batches) works on the drop-in model.
//
//   forward   x = [dense | embedding rows]                         (DeepFM.call, DL/model/deepFM.py:23-31)
//             fm = w0 + x w1 + 1/2 sum_f [(x v)_f^2 - (x^2)(v^2)_f]  (FMLayer.call, DL/layer/interaction.py:106-114)
//             h_l = dropout(relu(h_{l-1} W_l + b_l)),  dnn = h_L w_out + b_out   (DNNLayer.call, interaction.py:40-46)
//             p = sigmoid(1/2 (fm + dnn))
//   loss      mean_b BCE(y_b, clip(p_b, 1e-7, 1 - 1e-7))  +  reg_w sum w1^2  +  reg_v sum v^2     (keras binary_crossentropy;
//             l2 regularisers of FMLayer.build, interaction.py:97-104)
//   backward  exact fp32 gradients of that loss; embedding rows get the gradient of the samples that looked them up
//   update    w <- w - lr * grad (keras SGD, no momentum); embedding rows by atomic adds (keras sums duplicate indices)
//
// Everything is fp32 on the CUDA cores: the reference trains at batch 32 (compile_fit's default), where the step is a chain of
// small launches and tensor-core tiles would be mostly padding.  One generic tiled GEMM serves the tower both ways and every
// weight update (C = beta C + alpha op(A) op(B), strided operands).  Dropout uses a stateless hash of (seed, layer, element):
// the mask of the forward pass is recomputed, not stored, in the backward pass; it cannot match TensorFlow's generator.
#include "common.cuh"

namespace ctrhot {

constexpr int TR_T = 32;                       // GEMM tile
constexpr int TR_MAXL = 8;

// Optional device-resident control block (ctrhot_deepfm_train_step_ctl_f32): lets ONE captured CUDA graph of the step be
// replayed over consecutive batches -- the kernels read the batch's first row and the step number from here, and the last
// kernel of the step advances them.  32 bytes, caller-owned.
struct TrCtl {
  unsigned long long step;       // added to the dropout seed
  long long row0;                // first row of this batch inside ids / dense / y
  float acc_loss;                // += (sum over the batch of the per-sample BCE) -- an epoch's running total
  float acc_correct;             // += number of correct predictions
  long long reserved;
};
__global__ void tr_advance_kernel(TrCtl* ctl, long long B) { ctl->step += 1; ctl->row0 += B; }

// C[m, n] = beta * C[m, n] + alpha * sum_k A(m, k) B(k, n);  A(m, k) = A[m * am + k * ak], B(k, n) = B[k * bk + n * bn]
__global__ void __launch_bounds__(256)
tr_gemm_kernel(const float* __restrict__ A, int64_t am, int64_t ak, const float* __restrict__ Bm, int64_t bk, int64_t bn,
               float* __restrict__ C, int64_t ldc, int M, int N, int K, float alpha, float beta,
               float* __restrict__ csum = nullptr, float csum_alpha = 0.f) {
  __shared__ float As[TR_T][TR_T + 1], Bs[TR_T][TR_T + 1];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;          // 16 x 16 threads, 2 x 2 outputs each
  const int m0 = blockIdx.y * TR_T, n0 = blockIdx.x * TR_T;
  float acc[2][2] = {{0.f, 0.f}, {0.f, 0.f}};
  float cs0 = 0.f, cs1 = 0.f;
  const bool a_kfast = ak == 1 && am != 1, b_kfast = bk == 1 && bn != 1;
  // split-K (gridDim.z > 1, only launched with beta == 1): slice z of the k range, added to C with atomics
  const int kchunk = gridDim.z > 1 ? ((K + (int)gridDim.z - 1) / (int)gridDim.z + TR_T - 1) / TR_T * TR_T : K;
  const int kbeg = blockIdx.z * kchunk;
  K = min(K, kbeg + kchunk);
  for (int k0 = kbeg; k0 < K; k0 += TR_T) {
    for (int e = threadIdx.x; e < TR_T * TR_T; e += 256) {
      // As[k][m], Bs[k][n]; consecutive threads walk whichever index is contiguous in memory (whole 128-byte lines either way;
      // the 33-float rows keep the transposed store free of bank conflicts)
      const int ra = a_kfast ? e % TR_T : e / TR_T, ca = a_kfast ? e / TR_T : e % TR_T;
      const int m = m0 + ca, ka = k0 + ra;
      As[ra][ca] = (m < M && ka < K) ? A[(int64_t)m * am + (int64_t)ka * ak] : 0.f;
      const int rb = b_kfast ? e % TR_T : e / TR_T, cb = b_kfast ? e / TR_T : e % TR_T;
      const int n = n0 + cb, kb = k0 + rb;
      Bs[rb][cb] = (n < N && kb < K) ? Bm[(int64_t)kb * bk + (int64_t)n * bn] : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < TR_T; ++k) {
      const float a0 = As[k][ty], a1 = As[k][ty + 16], b0 = Bs[k][tx], b1 = Bs[k][tx + 16];
      acc[0][0] = fmaf(a0, b0, acc[0][0]); acc[0][1] = fmaf(a0, b1, acc[0][1]);
      acc[1][0] = fmaf(a1, b0, acc[1][0]); acc[1][1] = fmaf(a1, b1, acc[1][1]);
      cs0 += b0; cs1 += b1;
    }
    __syncthreads();
  }
  // csum[n] += csum_alpha * sum_k B(k, n): the bias gradient of a Dense layer rides in its weight-gradient product (B = g, k = the
  // batch) -- the first row of CTAs has the tiles of g in shared memory anyway
  if (csum && blockIdx.y == 0 && ty == 0) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int n = n0 + tx + 16 * j;
      const float v = csum_alpha * (j ? cs1 : cs0);
      if (n < N) { if (gridDim.z > 1) atomicAdd(csum + n, v); else csum[n] += v; }
    }
  }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int m = m0 + ty + 16 * i, n = n0 + tx + 16 * j;
      if (m < M && n < N) {
        float* c = C + (int64_t)m * ldc + n;
        if (gridDim.z > 1) atomicAdd(c, alpha * acc[i][j]);
        else *c = (beta == 0.f ? 0.f : beta * *c) + alpha * acc[i][j];
      }
    }
}

__device__ __forceinline__ float tr_uniform(uint64_t seed, uint32_t layer, uint64_t idx) {
  uint64_t z = seed + 0x9E3779B97F4A7C15ull * (idx + 1) + ((uint64_t)layer << 56);     // splitmix64
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(z >> 40) * (1.0f / 16777216.0f);
}

// optional epilogue of the small-M product: h = dropout(relu(acc + bias)), the Dense layer of a tower in one launch
struct TrEpi {
  const float* bias;
  const TrCtl* ctl;
  uint64_t seed;
  float rate;
  uint32_t layer;
  int relu, on;
};
// The reference's batch of 32 (and batches up to 128 as gridDim.y blocks of 32 rows).  The 32 x 32 kernel above walks K in 32-wide steps, each a global load, two barriers
// and 32 FMA deep -- ~1.2 us per step, 7 steps for the 221-wide first layer.  Here the 8 warps of a CTA take one eighth of the k
// range each (A slice staged in shared memory by the warp itself, B read straight from global: every load independent), then the
// eight partial tiles are added through shared memory in a fixed order.  One pass instead of K / 32 dependent ones.
__global__ void __launch_bounds__(256)
tr_gemm_m32_kernel(const float* __restrict__ A, int64_t am, int64_t ak, const float* __restrict__ Bm, int64_t bk, int64_t bn,
                   float* __restrict__ C, int64_t ldc, int M, int N, int K, float alpha, float beta, TrEpi epi) {
  __shared__ float sm[8][32][33];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int n = blockIdx.x * 32 + lane;
  const int mb = blockIdx.y * 32;                                 // batches up to 128 rows: one CTA row per 32 of them
  A += (int64_t)mb * am;
  C += (int64_t)mb * ldc;
  M = min(32, M - mb);
  const int kper = (K + 7) / 8, kw0 = w * kper, kw1 = min(K, kw0 + kper);
  float acc[32];
#pragma unroll
  for (int m = 0; m < 32; ++m) acc[m] = 0.f;
  for (int kc = kw0; kc < kw1; kc += 32) {
    const int kn = min(32, kw1 - kc);
    // this warp's A[0:32, kc:kc+kn) -> sm[w][m][k]; lanes along whichever index is contiguous
    if (ak == 1) {
#pragma unroll 8
      for (int m = 0; m < 32; ++m) sm[w][m][lane] = (m < M && lane < kn) ? A[(int64_t)m * am + (kc + lane)] : 0.f;
    } else {
#pragma unroll 8
      for (int k = 0; k < 32; ++k) sm[w][lane][k] = (lane < M && k < kn) ? A[(int64_t)lane * am + (int64_t)(kc + k) * ak] : 0.f;
    }
    float b[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) b[k] = (n < N && k < kn) ? Bm[(int64_t)(kc + k) * bk + (int64_t)n * bn] : 0.f;
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 32; ++k) {
#pragma unroll
      for (int m = 0; m < 32; ++m) acc[m] = fmaf(sm[w][m][k], b[k], acc[m]);
    }
    __syncwarp();
  }
  __syncthreads();
#pragma unroll
  for (int m = 0; m < 32; ++m) sm[w][m][lane] = acc[m];
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int idx = threadIdx.x + 256 * i, m = idx >> 5, nn = idx & 31, nc = blockIdx.x * 32 + nn;
    float t = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) t += sm[q][m][nn];
    if (m < M && nc < N) {
      float* c = C + (int64_t)m * ldc + nc;
      t = (beta == 0.f ? 0.f : beta * *c) + alpha * t;
      if (epi.on == 2) {                                          // backward: t *= d dropout(relu(.)) at h = epi.bias[m, nc]
        float mk = 1.f;
        if (epi.rate > 0.f) {
          const uint64_t seed = epi.seed + (epi.ctl ? epi.ctl->step : 0ull);
          mk = (tr_uniform(seed, epi.layer, (uint64_t)(mb + m) * N + nc) >= epi.rate) ? 1.f / (1.f - epi.rate) : 0.f;
        }
        if (epi.relu && !(epi.bias[(int64_t)(mb + m) * N + nc] > 0.f)) mk = 0.f;
        t *= mk;
      } else if (epi.on) {                                        // same arithmetic and mask index as tr_bias_relu_drop_kernel
        t += epi.bias ? epi.bias[nc] : 0.f;
        if (epi.relu) t = fmaxf(t, 0.f);
        if (epi.rate > 0.f) {
          const uint64_t seed = epi.seed + (epi.ctl ? epi.ctl->step : 0ull);
          t = (tr_uniform(seed, epi.layer, (uint64_t)(mb + m) * N + nc) >= epi.rate) ? t * (1.f / (1.f - epi.rate)) : 0.f;
        }
      }
      *c = t;
    }
  }
}

// The same product on 64 x 64 output tiles for the batch-sized operands (M >= 128, N >= 64): 256 threads x (4 x 4) outputs, k in
// steps of 16, the next step's operands fetched into registers while this one is multiplied.  8 FFMA per shared-memory load
// against 1 in the 32 x 32 kernel above.
constexpr int TR_T64 = 64, TR_K64 = 16;
__global__ void __launch_bounds__(256)
tr_gemm64_kernel(const float* __restrict__ A, int64_t am, int64_t ak, const float* __restrict__ Bm, int64_t bk, int64_t bn,
                 float* __restrict__ C, int64_t ldc, int M, int N, int K, float alpha, float beta,
                 float* __restrict__ csum = nullptr, float csum_alpha = 0.f) {
  __shared__ __align__(16) float As[TR_K64][TR_T64 + 4], Bs[TR_K64][TR_T64 + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m0 = blockIdx.y * TR_T64, n0 = blockIdx.x * TR_T64;
  const bool a_kfast = ak == 1 && am != 1, b_kfast = bk == 1 && bn != 1;
  const int kchunk = gridDim.z > 1 ? ((K + (int)gridDim.z - 1) / (int)gridDim.z + TR_K64 - 1) / TR_K64 * TR_K64 : K;
  const int kbeg = blockIdx.z * kchunk;
  K = min(K, kbeg + kchunk);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  float ra[4], rb[4], cs[4] = {0.f, 0.f, 0.f, 0.f};
  auto fetch = [&](int k0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int e = threadIdx.x + 256 * q;
      const int ka = a_kfast ? e % TR_K64 : e / TR_T64, ma = a_kfast ? e / TR_K64 : e % TR_T64;
      ra[q] = (m0 + ma < M && k0 + ka < K) ? A[(int64_t)(m0 + ma) * am + (int64_t)(k0 + ka) * ak] : 0.f;
      const int kb = b_kfast ? e % TR_K64 : e / TR_T64, nb = b_kfast ? e / TR_K64 : e % TR_T64;
      rb[q] = (n0 + nb < N && k0 + kb < K) ? Bm[(int64_t)(k0 + kb) * bk + (int64_t)(n0 + nb) * bn] : 0.f;
    }
  };
  auto stash = [&]() {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int e = threadIdx.x + 256 * q;
      const int ka = a_kfast ? e % TR_K64 : e / TR_T64, ma = a_kfast ? e / TR_K64 : e % TR_T64;
      As[ka][ma] = ra[q];
      const int kb = b_kfast ? e % TR_K64 : e / TR_T64, nb = b_kfast ? e / TR_K64 : e % TR_T64;
      Bs[kb][nb] = rb[q];
    }
  };
  if (kbeg < K) fetch(kbeg);
  for (int k0 = kbeg; k0 < K; k0 += TR_K64) {
    stash();
    __syncthreads();
    if (k0 + TR_K64 < K) fetch(k0 + TR_K64);
#pragma unroll
    for (int k = 0; k < TR_K64; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
#pragma unroll
      for (int j = 0; j < 4; ++j) cs[j] += bv[j];
    }
    __syncthreads();
  }
  if (csum && blockIdx.y == 0 && ty == 0) {                       // the bias gradient, as in tr_gemm_kernel
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n < N) { if (gridDim.z > 1) atomicAdd(csum + n, csum_alpha * cs[j]); else csum[n] += csum_alpha * cs[j]; }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = m0 + ty * 4 + i, n = n0 + tx * 4 + j;
      if (m < M && n < N) {
        float* c = C + (int64_t)m * ldc + n;
        if (gridDim.z > 1) atomicAdd(c, alpha * acc[i][j]);
        else *c = (beta == 0.f ? 0.f : beta * *c) + alpha * acc[i][j];
      }
    }
}

// x[b] = [dense_b (D) | rows of the F fields (F * k)]   (reference column order)
template <int IDKIND>
__global__ void tr_gather_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k,
                                 const void* __restrict__ ids, int64_t ids_ld, const float* __restrict__ dense, int64_t dense_ld,
                                 int D, int64_t B, float* __restrict__ x, int n, int* __restrict__ err_flag,
                                 const TrCtl* __restrict__ ctl, int64_t x_ld = 0) {
  if (x_ld == 0) x_ld = n;
  const int64_t r0 = ctl ? ctl->row0 : 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / n;
    const int c = (int)(i - b * n);
    float v;
    if (c < D) {
      v = dense[(b + r0) * dense_ld + c];
    } else {
      const int f = (c - D) / k, e = (c - D) - f * k;
      const int64_t id = IdLoad<IDKIND>::at(ids, (b + r0) * ids_ld + f);
      const int64_t rows = row_off[f + 1] - row_off[f];
      const bool ok = id >= 0 && id < rows;
      if (!ok && err_flag) *err_flag = 1;
      v = ok ? arena[(row_off[f] + id) * k + e] : 0.f;
    }
    x[b * x_ld + c] = v;
  }
}

// one warp per sample: xv[b, f] = sum_j x_bj v_jf,  fm[b] = w0 + x.w1 + 1/2 sum_f (xv_f^2 - sum_j x_bj^2 v_jf^2)
__global__ void tr_fm_fwd_kernel(const float* __restrict__ x, int n, int64_t B, const float* __restrict__ w0,
                                 const float* __restrict__ w1, const float* __restrict__ v, int kf,
                                 float* __restrict__ xv, float* __restrict__ fm) {
  const int lane = threadIdx.x & 31;
  const int64_t b = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= B) return;
  float lin = 0.f, sq = 0.f, acc[32];
#pragma unroll
  for (int f = 0; f < 32; ++f) acc[f] = 0.f;
  for (int j = lane; j < n; j += 32) {
    const float xj = x[b * n + j];
    lin = fmaf(xj, w1[j], lin);
#pragma unroll
    for (int f = 0; f < 32; ++f)
      if (f < kf) { const float vv = v[(int64_t)j * kf + f]; acc[f] = fmaf(xj, vv, acc[f]); sq = fmaf(xj * xj, vv * vv, sq); }
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) { lin += __shfl_xor_sync(0xffffffffu, lin, o); sq += __shfl_xor_sync(0xffffffffu, sq, o); }
  float s1 = 0.f;
#pragma unroll
  for (int f = 0; f < 32; ++f) {
    if (f < kf) {
      float a = acc[f];
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
      if (lane == 0) xv[b * kf + f] = a;
      s1 = fmaf(a, a, s1);
    }
  }
  if (lane == 0) fm[b] = w0[0] + lin + 0.5f * (s1 - sq);
}

// h = dropout(relu(h + bias)) in place (h holds the GEMM result); mask recomputable from (seed, layer, element)
__global__ void tr_bias_relu_drop_kernel(float* __restrict__ h, const float* __restrict__ bias, int N, int64_t B, int relu,
                                         float rate, uint64_t seed, uint32_t layer, const TrCtl* __restrict__ ctl) {
  const float scale = rate > 0.f ? 1.f / (1.f - rate) : 1.f;
  if (ctl) seed += ctl->step;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * N; i += (int64_t)gridDim.x * blockDim.x) {
    float t = h[i] + (bias ? bias[i % N] : 0.f);
    if (relu) t = fmaxf(t, 0.f);
    if (rate > 0.f) t = (tr_uniform(seed, layer, (uint64_t)i) >= rate) ? t * scale : 0.f;
    h[i] = t;
  }
}

// p, per-sample loss / correctness and dz = dLoss/dlogit (loss = mean BCE over the batch, probabilities clipped like keras)
__global__ void tr_loss_kernel(const float* __restrict__ fm, const float* __restrict__ dnn, const float* __restrict__ b_out,
                               const float* __restrict__ y, int64_t B, float* __restrict__ dz, float* __restrict__ loss_out,
                               TrCtl* __restrict__ ctl, float zscale) {
  __shared__ float s_loss[256], s_ok[256];
  float l = 0.f, ok = 0.f;
  if (ctl) y += ctl->row0;
  for (int64_t b = threadIdx.x; b < B; b += blockDim.x) {
    const float z = zscale * (fm[b] + (dnn ? dnn[b] : 0.f) + (b_out ? b_out[0] : 0.f));   // 1/2 (fm + dnn): DeepFM; fm: FM
    const float p = 1.0f / (1.0f + __expf(-z));
    const float eps = 1e-7f;
    const float pc = fminf(fmaxf(p, eps), 1.f - eps);
    const float yb = y[b];
    l -= yb * __logf(pc) + (1.f - yb) * __logf(1.f - pc);
    ok += ((p > 0.5f) == (yb > 0.5f)) ? 1.f : 0.f;
    // d/dp of the clipped BCE is zero where the clip is active; times dp/dz = p (1 - p); times 1/B for the mean
    const float g = (p > eps && p < 1.f - eps) ? (pc - yb) : 0.f;
    dz[b] = g / (float)B;
  }
  s_loss[threadIdx.x] = l; s_ok[threadIdx.x] = ok;
  __syncthreads();
  for (int o = blockDim.x / 2; o >= 1; o >>= 1) {               // fixed order: the reported loss is deterministic
    if (threadIdx.x < o) { s_loss[threadIdx.x] += s_loss[threadIdx.x + o]; s_ok[threadIdx.x] += s_ok[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    loss_out[0] = s_loss[0] / (float)B; loss_out[1] = s_ok[0];
    if (ctl) { ctl->acc_loss += s_loss[0]; ctl->acc_correct += s_ok[0]; }
  }
}

// g[b, n] *= (h[b, n] > 0 ? dropout scale : 0): the derivative of dropout(relu(.)) at the stored activation
__global__ void tr_relu_drop_bwd_kernel(float* __restrict__ g, const float* __restrict__ h, int64_t total, int relu, float rate,
                                        uint64_t seed, uint32_t layer, const TrCtl* __restrict__ ctl) {
  const float scale = rate > 0.f ? 1.f / (1.f - rate) : 1.f;
  if (ctl) seed += ctl->step;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    float m = 1.f;
    if (rate > 0.f) m = (tr_uniform(seed, layer, (uint64_t)i) >= rate) ? scale : 0.f;
    if (relu && !(h[i] > 0.f)) m = 0.f;
    g[i] *= m;
  }
}

// out[n] = beta * out[n] + alpha * sum_b w_b A[b, n] (w = NULL: ones).  Block = 32 columns x 32 row groups; every thread sums
// its rows in order, the 32 partial sums of a column are added in a fixed order: deterministic for a given B below 2048 rows
__global__ void __launch_bounds__(1024)
tr_colsum_kernel(const float* __restrict__ A, int64_t lda, const float* __restrict__ w, int64_t B, int N,
                 int square, float* __restrict__ out, float alpha, float beta) {
  __shared__ float part[32][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int n = blockIdx.x * 32 + cx;
  float s = 0.f;
  // gridDim.y > 1 (launched only with beta == 1 and thousands of rows): row slices, added with atomics
  const int64_t rows = (B + gridDim.y - 1) / gridDim.y, b0 = blockIdx.y * rows, b1 = b0 + rows < B ? b0 + rows : B;
  if (n < N)
    for (int64_t b = b0 + ry; b < b1; b += 32) {
      const float a = A[b * lda + n];
      s = fmaf(w ? w[b] : 1.f, square ? a * a : a, s);
    }
  part[ry][cx] = s;
  __syncthreads();
  if (ry == 0 && n < N) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < 32; ++r) t += part[r][cx];
    if (gridDim.y > 1) atomicAdd(out + n, alpha * t);
    else out[n] = (beta == 0.f ? 0.f : beta * out[n]) + alpha * t;
  }
}

// dx[b, j] += dfm_b (w1_j + sum_f v_jf xv_bf - x_bj sum_f v_jf^2);  also t[b, f] = dfm_b xv_bf (the right factor of dv)
__global__ void tr_fm_bwd_kernel(const float* __restrict__ x, int n, int64_t B, const float* __restrict__ dz,
                                 const float* __restrict__ w1, const float* __restrict__ v, int kf,
                                 const float* __restrict__ xv, float* __restrict__ dx, float* __restrict__ t, float gscale) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / n;
    const int j = (int)(i - b * n);
    const float dfm = gscale * dz[b];
    if (dx) {                                              // (a model whose x is its input needs no gradient for it)
      float s = w1[j], vsq = 0.f;
      for (int f = 0; f < kf; ++f) { const float vv = v[(int64_t)j * kf + f]; s = fmaf(vv, xv[b * kf + f], s); vsq = fmaf(vv, vv, vsq); }
      dx[i] += dfm * (s - x[i] * vsq);
    }
    if (j < kf) t[b * kf + j] = dfm * xv[b * kf + j];
  }
}

// v[j, f] *= decay + lr * s2[j]   (the -x_j^2 v_jf part of dv and the l2 regulariser, before the GEMM adds -lr x^T t)
__global__ void tr_scale_rows_kernel(float* __restrict__ v, int n, int kf, const float* __restrict__ s2, float decay, float lr) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n * kf; i += gridDim.x * blockDim.x)
    v[i] *= decay + lr * s2[i / kf];
}

// arena rows -= lr * dx[b, D + f k + e]: atomic adds (a row looked up by several samples gets the sum, as keras' sparse SGD)
template <int IDKIND>
__global__ void tr_embed_sgd_kernel(float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k,
                                    const void* __restrict__ ids, int64_t ids_ld, int D, int64_t B, const float* __restrict__ dx,
                                    int n, float lr, const TrCtl* __restrict__ ctl) {
  const int fk = F * k;
  const int64_t r0 = ctl ? ctl->row0 : 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * fk; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / fk;
    const int c = (int)(i - b * fk), f = c / k, e = c - f * k;
    const int64_t id = IdLoad<IDKIND>::at(ids, (b + r0) * ids_ld + f);
    if (id >= 0 && id < row_off[f + 1] - row_off[f])
      atomicAdd(arena + (row_off[f] + id) * k + e, -lr * dx[b * n + D + c]);
  }
}

// ---------------------------------------------------------------- DCN (DL/model/dcn.py:15-34): cross layers, forward and backward
// One warp per sample.  x_{l+1} = x_0 s_l + b_l + x_l with s_l = x_l . w_l (CrossLayer.call, interaction.py:75-83); the rows
// x_1 .. x_L and the scalars s_l are kept for the backward pass; zc[b] = x_L . fk[0:n] is the cross half of the logit.
constexpr int TR_NJ = 32;                                       // columns per lane: n <= 1024
struct TrCross { const float* w[TR_MAXL]; const float* b[TR_MAXL]; };
__global__ void __launch_bounds__(256)
tr_cross_fwd_kernel(const float* __restrict__ x, int n, int64_t B, int L, TrCross cw, const float* __restrict__ fk,
                    float* __restrict__ XL, float* __restrict__ S, float* __restrict__ zc) {
  const int lane = threadIdx.x & 31;
  const int64_t b = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= B) return;
  float x0[TR_NJ], xl[TR_NJ];
#pragma unroll
  for (int i = 0; i < TR_NJ; ++i) {
    const int j = lane + 32 * i;
    x0[i] = j < n ? x[b * n + j] : 0.f;
    xl[i] = x0[i];
  }
  for (int l = 0; l < L; ++l) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < TR_NJ; ++i) { const int j = lane + 32 * i; if (j < n) s = fmaf(xl[i], cw.w[l][j], s); }
    s = warp_sum(s);
    if (lane == 0) S[(int64_t)l * B + b] = s;
#pragma unroll
    for (int i = 0; i < TR_NJ; ++i) {
      const int j = lane + 32 * i;
      if (j < n) { xl[i] = fmaf(x0[i], s, cw.b[l][j]) + xl[i]; XL[((int64_t)l * B + b) * n + j] = xl[i]; }
    }
  }
  float z = 0.f;
#pragma unroll
  for (int i = 0; i < TR_NJ; ++i) { const int j = lane + 32 * i; if (j < n) z = fmaf(xl[i], fk[j], z); }
  z = warp_sum(z);
  if (lane == 0) zc[b] = z;
}
// out[b] = a[b] (or 0) + v[b] * s[0]
__global__ void tr_axpy_scalar_kernel(const float* __restrict__ a, const float* __restrict__ v, const float* __restrict__ s,
                                      int64_t B, float* __restrict__ out) {
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < B; b += (int64_t)gridDim.x * blockDim.x)
    out[b] = (a ? a[b] : 0.f) + v[b] * s[0];
}
// g = dL/dx_{l+1}, from g_L = dz fk[0:n] down: G[l] = g (for db_l), C[l] = c = g . x_0 (for dw_l = sum_b c x_l),
// dx_0 += g s_l, g += c w_l (the OLD w_l: the updates run after this kernel); finally dx[b] += g + dx_0
__global__ void __launch_bounds__(256)
tr_cross_bwd_kernel(const float* __restrict__ x, int n, int64_t B, int L, TrCross cw, const float* __restrict__ fk,
                    const float* __restrict__ dz, const float* __restrict__ S, float* __restrict__ G, float* __restrict__ C,
                    float* __restrict__ dx) {
  const int lane = threadIdx.x & 31;
  const int64_t b = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= B) return;
  float x0[TR_NJ], g[TR_NJ], d0[TR_NJ];
  const float dzb = dz[b];
#pragma unroll
  for (int i = 0; i < TR_NJ; ++i) {
    const int j = lane + 32 * i;
    x0[i] = j < n ? x[b * n + j] : 0.f;
    g[i] = j < n ? dzb * fk[j] : 0.f;
    d0[i] = 0.f;
  }
  for (int l = L - 1; l >= 0; --l) {
    float c = 0.f;
#pragma unroll
    for (int i = 0; i < TR_NJ; ++i) {
      const int j = lane + 32 * i;
      if (j < n) { G[((int64_t)l * B + b) * n + j] = g[i]; c = fmaf(g[i], x0[i], c); }
    }
    c = warp_sum(c);
    if (lane == 0) C[(int64_t)l * B + b] = c;
    const float s = S[(int64_t)l * B + b];
#pragma unroll
    for (int i = 0; i < TR_NJ; ++i) {
      const int j = lane + 32 * i;
      if (j < n) { d0[i] = fmaf(g[i], s, d0[i]); g[i] = fmaf(c, cw.w[l][j], g[i]); }
    }
  }
#pragma unroll
  for (int i = 0; i < TR_NJ; ++i) { const int j = lane + 32 * i; if (j < n) dx[b * n + j] += g[i] + d0[i]; }
}

// ---------------------------------------------------------------- PNN (DL/model/pnn.py:74-85, the reference's own training loop)
// keras' binary_crossentropy on the model's RAW output (PNN.call has no sigmoid): p = clip(z, 1e-7, 1 - 1e-7),
// loss = mean -(y log p + (1 - y) log(1 - p)), d loss / dz = (p - y) / (p (1 - p)) / B inside the clip and 0 outside.
__global__ void tr_raw_bce_kernel(const float* __restrict__ z, const float* __restrict__ b_out, const float* __restrict__ y,
                                  int64_t B, float* __restrict__ dz, float* __restrict__ loss_out) {
  __shared__ float s_loss[256], s_ok[256];
  float l = 0.f, ok = 0.f;
  for (int64_t b = threadIdx.x; b < B; b += blockDim.x) {
    const float zz = z[b] + (b_out ? b_out[0] : 0.f);
    const float eps = 1e-7f;
    const float pc = fminf(fmaxf(zz, eps), 1.f - eps);
    const float yb = y[b];
    l -= yb * __logf(pc) + (1.f - yb) * __logf(1.f - pc);
    ok += ((zz > 0.5f) == (yb > 0.5f)) ? 1.f : 0.f;
    dz[b] = (zz > eps && zz < 1.f - eps) ? (pc - yb) / (pc * (1.f - pc)) / (float)B : 0.f;
  }
  s_loss[threadIdx.x] = l; s_ok[threadIdx.x] = ok;
  __syncthreads();
  for (int o = blockDim.x / 2; o >= 1; o >>= 1) {
    if (threadIdx.x < o) { s_loss[threadIdx.x] += s_loss[threadIdx.x + o]; s_ok[threadIdx.x] += s_ok[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { loss_out[0] = s_loss[0] / (float)B; loss_out[1] = s_ok[0]; }
}
__device__ __forceinline__ int tr_pair(int i, int j, int F) { return i * (2 * F - i - 1) / 2 + (j - i - 1); }   // i < j
// Product layers backward, one thread per (sample, field f, component t): with pair p = (i, j), i < j, upstream gi_p (inner) and
// go_p (outer):  de_i[c] += gi e_j[c] + go sum_a W[a, p, c] e_j[a];  de_j[a] += gi e_i[a] + go sum_c W[a, p, c] e_i[c].
// dx's flat columns (already holding the tower's gradient for them) are updated in place; its product columns are read.
__global__ void tr_pnn_prod_bwd_kernel(const float* __restrict__ x, int n, int64_t B, int F, int K, int P,
                                       const float* __restrict__ W, int col_in, int col_out, float* __restrict__ dx) {
  const int fk = F * K;
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < B * fk; i0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i0 / fk;
    const int c0 = (int)(i0 - b * fk), f = c0 / K, t = c0 - f * K;
    const float* e = x + b * n;
    const float* g = dx + b * n;
    float acc = 0.f;
    for (int o = 0; o < F; ++o) {
      if (o == f) continue;
      const bool first = f < o;                              // f plays e_i (index c) when it is the smaller field of the pair
      const int p = first ? tr_pair(f, o, F) : tr_pair(o, f, F);
      const float* eo = e + o * K;
      if (col_in >= 0) acc = fmaf(g[col_in + p], eo[t], acc);
      if (col_out >= 0) {
        float s = 0.f;
        if (first) { for (int a = 0; a < K; ++a) s = fmaf(W[((int64_t)a * P + p) * K + t], eo[a], s); }
        else { for (int c = 0; c < K; ++c) s = fmaf(W[((int64_t)t * P + p) * K + c], eo[c], s); }
        acc = fmaf(g[col_out + p], s, acc);
      }
    }
    dx[b * n + c0] += acc;
  }
}
// W[a, p, c] -= lr sum_b go[b, p] e_j[b, a] e_i[b, c]   (one thread per weight, batch summed in order: deterministic)
__global__ void tr_pnn_dw_kernel(const float* __restrict__ x, int n, int64_t B, int F, int K, int P, const float* __restrict__ dx,
                                 int col_out, float* __restrict__ W, float lr) {
  const int total = K * P * K;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int c = idx % K, p = (idx / K) % P, a = idx / (K * P);
    int i = 0, rem = p;
    while (rem >= F - 1 - i) { rem -= F - 1 - i; ++i; }
    const int j = i + 1 + rem;
    float s = 0.f;
    for (int64_t b = 0; b < B; ++b) s = fmaf(dx[b * n + col_out + p] * x[b * n + j * K + a], x[b * n + i * K + c], s);
    W[idx] -= lr * s;
  }
}

// ---------------------------------------------------------------- Wide&Deep (DL/model/wideDeep.py:14-34): the wide side on ids
// wide input = [dense (D) | dense again (D) | one 0/1 block per field] (DL/utils/dataset.py:53-58), never materialised:
// wide[b] = w0 + sum_j dense_bj (w_j + w_{D+j}) + sum_f [id_bf < width_f] w[col_off[f] + id_bf]
template <int IDKIND>
__global__ void tr_wide_fwd_kernel(const float* __restrict__ dense, int64_t dense_ld, int D, const void* __restrict__ ids,
                                   int64_t ids_ld, int F, const int64_t* __restrict__ col_off, const float* __restrict__ w0,
                                   const float* __restrict__ w, int64_t B, float* __restrict__ wide) {
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < B; b += (int64_t)gridDim.x * blockDim.x) {
    float s = w0[0];
    for (int j = 0; j < D; ++j) s = fmaf(dense[b * dense_ld + j], w[j] + w[D + j], s);
    for (int f = 0; f < F; ++f) {
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      if (id >= 0 && id < col_off[f + 1] - col_off[f]) s += w[col_off[f] + id];
    }
    wide[b] = s;
  }
}
__global__ void tr_scale_kernel(float* __restrict__ w, int64_t n, float a) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) w[i] *= a;
}
// w[col_off[f] + id] -= lr * gscale * dz[b]   (one thread per (sample, field); columns hit by several samples add up)
template <int IDKIND>
__global__ void tr_wide_onehot_sgd_kernel(const void* __restrict__ ids, int64_t ids_ld, int F, const int64_t* __restrict__ col_off,
                                          const float* __restrict__ dz, float gl, int64_t B, float* __restrict__ w) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * F; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / F;
    const int f = (int)(i - b * F);
    const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
    if (id >= 0 && id < col_off[f + 1] - col_off[f]) atomicAdd(w + col_off[f] + id, -gl * dz[b]);
  }
}

// ---------------------------------------------------------------- DeepCrossing (DL/model/deepCrossing.py:14-32): residual units
// out = relu(x + h)   (ResLayer.call, DL/layer/interaction.py:277-278)
__global__ void tr_residual_relu_kernel(const float* __restrict__ x, const float* __restrict__ h, int64_t total, float* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = fmaxf(x[i] + h[i], 0.f);
}
// gm = g where the unit's output is positive, else 0 (the derivative of that relu)
__global__ void tr_mask_copy_kernel(const float* __restrict__ g, const float* __restrict__ xo, int64_t total, float* __restrict__ gm) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x)
    gm[i] = xo[i] > 0.f ? g[i] : 0.f;
}
__global__ void tr_add_kernel(float* __restrict__ a, const float* __restrict__ b, int64_t total) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) a[i] += b[i];
}

// ---------------------------------------------------------------- NFM (DL/model/nfm.py:15-34) in training mode
// x[b] = [dense (D) | 0.5 ((sum_f e_f)^2 - sum_f e_f^2) (k)];  S[b, c] = sum_f e_fc is kept for the backward pass
template <int IDKIND>
__global__ void tr_nfm_bi_fwd_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k,
                                     const void* __restrict__ ids, int64_t ids_ld, const float* __restrict__ dense, int64_t dense_ld,
                                     int D, int64_t B, float* __restrict__ x, float* __restrict__ S, int* __restrict__ err_flag) {
  const int n = D + k;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / n;
    const int c = (int)(i - b * n);
    if (c < D) { x[i] = dense[b * dense_ld + c]; continue; }
    float s = 0.f, q = 0.f;
    for (int f = 0; f < F; ++f) {
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      const bool ok = id >= 0 && id < row_off[f + 1] - row_off[f];
      if (!ok && err_flag) *err_flag = 1;
      const float e = ok ? arena[(row_off[f] + id) * k + (c - D)] : 0.f;
      s += e; q = fmaf(e, e, q);
    }
    S[b * k + (c - D)] = s;
    x[i] = 0.5f * (s * s - q);
  }
}
// keras BatchNormalization, training: per column the batch mean and BIASED variance (two passes, fixed order), xhat, y = gamma
// xhat + beta, and the moving statistics  m <- momentum m + (1 - momentum) stat.  One block of 256 threads per column.
__global__ void __launch_bounds__(256)
tr_bn_fwd_kernel(const float* __restrict__ x, int n, int64_t B, const float* __restrict__ gamma, const float* __restrict__ beta,
                 float* __restrict__ mov_mean, float* __restrict__ mov_var, float momentum, float eps, float* __restrict__ xhat,
                 float* __restrict__ y, float* __restrict__ rstd_out) {
  __shared__ float red[256];
  const int j = blockIdx.x;
  float s = 0.f;
  for (int64_t b = threadIdx.x; b < B; b += 256) s += x[b * n + j];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o >= 1; o >>= 1) { if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
  const float mean = red[0] / (float)B;
  __syncthreads();
  float v = 0.f;
  for (int64_t b = threadIdx.x; b < B; b += 256) { const float d = x[b * n + j] - mean; v = fmaf(d, d, v); }
  red[threadIdx.x] = v;
  __syncthreads();
  for (int o = 128; o >= 1; o >>= 1) { if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
  const float var = red[0] / (float)B;
  const float rstd = 1.0f / sqrtf(var + eps);
  const float g = gamma[j], bt = beta[j];
  for (int64_t b = threadIdx.x; b < B; b += 256) {
    const float h = (x[b * n + j] - mean) * rstd;
    xhat[b * n + j] = h;
    y[b * n + j] = fmaf(g, h, bt);
  }
  if (threadIdx.x == 0) {
    rstd_out[j] = rstd;
    mov_mean[j] = momentum * mov_mean[j] + (1.f - momentum) * mean;
    mov_var[j] = momentum * mov_var[j] + (1.f - momentum) * var;
  }
}
// backward: dgamma = sum_b g xhat, dbeta = sum_b g, gx = gamma rstd (g - dbeta / B - xhat dgamma / B); then gamma, beta move
__global__ void __launch_bounds__(256)
tr_bn_bwd_kernel(const float* __restrict__ g, const float* __restrict__ xhat, int n, int64_t B, float* __restrict__ gamma,
                 float* __restrict__ beta, const float* __restrict__ rstd, float lr, float* __restrict__ gx) {
  __shared__ float r1[256], r2[256];
  const int j = blockIdx.x;
  float a = 0.f, c = 0.f;
  for (int64_t b = threadIdx.x; b < B; b += 256) { const float gg = g[b * n + j]; a = fmaf(gg, xhat[b * n + j], a); c += gg; }
  r1[threadIdx.x] = a; r2[threadIdx.x] = c;
  __syncthreads();
  for (int o = 128; o >= 1; o >>= 1) {
    if ((int)threadIdx.x < o) { r1[threadIdx.x] += r1[threadIdx.x + o]; r2[threadIdx.x] += r2[threadIdx.x + o]; }
    __syncthreads();
  }
  const float dgamma = r1[0], dbeta = r2[0];
  const float sc = gamma[j] * rstd[j], ib = 1.f / (float)B;
  for (int64_t b = threadIdx.x; b < B; b += 256)
    gx[b * n + j] = sc * (g[b * n + j] - dbeta * ib - xhat[b * n + j] * dgamma * ib);
  __syncthreads();
  if (threadIdx.x == 0) { gamma[j] -= lr * dgamma; beta[j] -= lr * dbeta; }
}
// Bi-Interaction backward: dE[b, f, c] = gx[b, D + c] (S[b, c] - e_fc), rows read from the arena BEFORE it moves
template <int IDKIND>
__global__ void tr_nfm_bi_bwd_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k,
                                     const void* __restrict__ ids, int64_t ids_ld, int D, int64_t B, const float* __restrict__ gx,
                                     const float* __restrict__ S, float* __restrict__ dE) {
  const int fk = F * k, n = D + k;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * fk; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / fk;
    const int r = (int)(i - b * fk), f = r / k, c = r - f * k;
    const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
    const bool ok = id >= 0 && id < row_off[f + 1] - row_off[f];
    const float e = ok ? arena[(row_off[f] + id) * k + c] : 0.f;
    dE[i] = gx[b * n + D + c] * (S[b * k + c] - e);
  }
}

// ---------------------------------------------------------------- AFM (DL/model/afm.py, AFMLayer DL/layer/interaction.py:322-351)
// x[b, c] = pool over pairs i < j of e_ic e_jc: sum ('att': the softmax over a size-1 axis makes every weight 1), mean, or max
// (with the arg-max pair kept).  sum = 0.5 (S^2 - Q); S[b, c] = sum_f e_fc is kept for the backward pass.
template <int IDKIND>
__global__ void tr_afm_fwd_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k,
                                  const void* __restrict__ ids, int64_t ids_ld, int64_t B, int mode, float* __restrict__ x,
                                  float* __restrict__ S, int32_t* __restrict__ amax, int* __restrict__ err_flag) {
  const int P = F * (F - 1) / 2;
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < B * k; i0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i0 / k;
    const int c = (int)(i0 - b * k);
    auto row = [&](int f) {
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      const bool ok = id >= 0 && id < row_off[f + 1] - row_off[f];
      if (!ok && err_flag) *err_flag = 1;
      return ok ? arena[(row_off[f] + id) * k + c] : 0.f;
    };
    if (mode == 2) {                                               // max over pairs, first maximum wins (numpy / tf argmax)
      float best = -INFINITY;
      int bi = 0, bj = 1;
      for (int i = 0; i < F - 1; ++i) {
        const float ei = row(i);
        for (int j = i + 1; j < F; ++j) {
          const float v = ei * row(j);
          if (v > best) { best = v; bi = i; bj = j; }
        }
      }
      x[i0] = best;
      amax[i0] = (bi << 16) | bj;
    } else {
      float s = 0.f, q = 0.f;
      for (int f = 0; f < F; ++f) { const float e = row(f); s += e; q = fmaf(e, e, q); }
      S[i0] = s;
      const float sum = 0.5f * (s * s - q);
      x[i0] = mode == 0 ? sum / (float)P : sum;
    }
  }
}
// q = sigmoid(z + b); p = sigmoid(q) (n_sigmoid 2: AFM.call) or q (AFMLayer alone); mean clipped BCE; dz = d loss / d z
__global__ void tr_afm_loss_kernel(const float* __restrict__ z, const float* __restrict__ b_out, const float* __restrict__ y,
                                   int64_t B, int n_sigmoid, float* __restrict__ dz, float* __restrict__ loss_out) {
  __shared__ float s_loss[256], s_ok[256];
  float l = 0.f, ok = 0.f;
  for (int64_t b = threadIdx.x; b < B; b += blockDim.x) {
    const float q = 1.0f / (1.0f + __expf(-(z[b] + b_out[0])));
    const float p = n_sigmoid == 2 ? 1.0f / (1.0f + __expf(-q)) : q;
    const float eps = 1e-7f;
    const float pc = fminf(fmaxf(p, eps), 1.f - eps);
    const float yb = y[b];
    l -= yb * __logf(pc) + (1.f - yb) * __logf(1.f - pc);
    ok += ((p > 0.5f) == (yb > 0.5f)) ? 1.f : 0.f;
    float g = (p > eps && p < 1.f - eps) ? (pc - yb) / (float)B : 0.f;           // d / d(logit of the last sigmoid)
    if (n_sigmoid == 2) g *= q * (1.f - q);
    dz[b] = g;
  }
  s_loss[threadIdx.x] = l; s_ok[threadIdx.x] = ok;
  __syncthreads();
  for (int o = blockDim.x / 2; o >= 1; o >>= 1) {
    if (threadIdx.x < o) { s_loss[threadIdx.x] += s_loss[threadIdx.x + o]; s_ok[threadIdx.x] += s_ok[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { loss_out[0] = s_loss[0] / (float)B; loss_out[1] = s_ok[0]; }
}
// max pooling backward: only the arg-max pair (i, j) of (b, c) receives gx:  dE[b, i, c] = gx e_jc,  dE[b, j, c] = gx e_ic
template <int IDKIND>
__global__ void tr_afm_max_bwd_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k,
                                      const void* __restrict__ ids, int64_t ids_ld, int64_t B, const float* __restrict__ gx,
                                      const int32_t* __restrict__ amax, float* __restrict__ dE) {
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < B * k; i0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i0 / k;
    const int c = (int)(i0 - b * k);
    for (int f = 0; f < F; ++f) dE[(b * F + f) * k + c] = 0.f;
    const int i = amax[i0] >> 16, j = amax[i0] & 0xFFFF;
    auto row = [&](int f) {
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      return (id >= 0 && id < row_off[f + 1] - row_off[f]) ? arena[(row_off[f] + id) * k + c] : 0.f;
    };
    const float g = gx[i0];
    dE[(b * F + i) * k + c] = g * row(j);
    dE[(b * F + j) * k + c] = g * row(i);
  }
}

// ---------------------------------------------------------------- FFM (DL/model/ffm.py:14-23, FFMLayer interaction.py:141-163)
// One warp per sample.  ff[b, e] (e = field * k + c) = sum over the sample's non-zero features i of x_i v[i, e]; S[c] = sum over
// fields of ff; z = w0 + sum_i x_i w_i + 0.5 (|S|^2 - |ff|^2).  The one-hot matrix is never built: the D counts and one row of v
// per sparse field (ids outside their column block: tf.one_hot's zero row).
#define TR_WARP_IDS()                                                               \
  const int lane = threadIdx.x & 31;                                                \
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); \
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
template <int IDKIND>
__global__ void __launch_bounds__(128)
tr_ffm_fwd_kernel(const float* __restrict__ dense, int64_t dense_ld, int D, const void* __restrict__ ids, int64_t ids_ld, int F,
                  const int64_t* __restrict__ col_off, int64_t B, const float* __restrict__ w0, const float* __restrict__ w,
                  const float* __restrict__ v, int len, int k, float* __restrict__ ff, float* __restrict__ S,
                  float* __restrict__ z) {
  TR_WARP_IDS();
  for (int64_t b = warp; b < B; b += nwarps) {
    float* fb = ff + b * len;
    float lin = 0.f, sq = 0.f;
    for (int e = lane; e < len; e += 32) {
      float a = 0.f;
      for (int j = 0; j < D; ++j) a = fmaf(dense[b * dense_ld + j], v[(size_t)j * len + e], a);
      for (int f = 0; f < F; ++f) {
        const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
        if (id >= 0 && id < col_off[f + 1] - col_off[f]) a += v[(size_t)(col_off[f] + id) * len + e];
      }
      fb[e] = a;
      sq = fmaf(a, a, sq);
    }
    for (int j = lane; j < D; j += 32) lin = fmaf(dense[b * dense_ld + j], w[j], lin);
    for (int f = lane; f < F; f += 32) {
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      if (id >= 0 && id < col_off[f + 1] - col_off[f]) lin += w[col_off[f] + id];
    }
    __syncwarp();
    float tsq = 0.f;
    for (int c = lane; c < k; c += 32) {
      float t = 0.f;
      for (int e = c; e < len; e += k) t += fb[e];
      S[b * k + c] = t;
      tsq = fmaf(t, t, tsq);
    }
    const float r = warp_sum(lin + 0.5f * (tsq - sq));
    if (lane == 0) z[b] = r + w0[0];
  }
}
// SGD on the rows the batch touched, after the l2 decay of all of w and v:  d z / d v[i, e] = x_i (S[e % k] - ff[e]),
// d z / d w_i = x_i.  One warp per (sample, non-zero feature); rows shared by several samples add up through atomics.
template <int IDKIND>
__global__ void __launch_bounds__(128)
tr_ffm_sgd_kernel(const float* __restrict__ dense, int64_t dense_ld, int D, const void* __restrict__ ids, int64_t ids_ld, int F,
                  const int64_t* __restrict__ col_off, int64_t B, const float* __restrict__ dz, float lr, int len, int k,
                  const float* __restrict__ ff, const float* __restrict__ S, float* __restrict__ w, float* __restrict__ v) {
  TR_WARP_IDS();
  const int nf = D + F;
  for (int64_t i = warp; i < B * nf; i += nwarps) {
    const int64_t b = i / nf;
    const int j = (int)(i - b * nf);
    int64_t col;
    float xs;
    if (j < D) { col = j; xs = dense[b * dense_ld + j]; }
    else {
      const int f = j - D;
      const int64_t id = IdLoad<IDKIND>::at(ids, b * ids_ld + f);
      if (id < 0 || id >= col_off[f + 1] - col_off[f]) continue;
      col = col_off[f] + id; xs = 1.f;
    }
    const float g = -lr * dz[b] * xs;
    if (g == 0.f) continue;
    if (lane == 0) atomicAdd(w + col, g);
    float* vr = v + (size_t)col * len;
    for (int e = lane; e < len; e += 32) atomicAdd(vr + e, g * (S[b * k + e % k] - ff[b * len + e]));
  }
}

// ---------------------------------------------------------------- DIN (DL/model/din.py:56-95 under compile_fit; Attention interaction.py:369-406)
// history rows k [B, T, K], the candidate's row q [B, K] (same table), the attention unit's input A = [q, k, q - k, q * k] and
// the mask (history id != 0).  q also goes to its place in the concatenated emb = [att | q | other].  Ids outside the table
// raise the flag and read as a zero row.
__global__ void tr_din_gather_kernel(const float* __restrict__ tab, int64_t rows, int K, const int32_t* __restrict__ hist,
                                     int64_t hist_ld, int T, const int32_t* __restrict__ cand, int64_t B, float* __restrict__ kbuf,
                                     float* __restrict__ qbuf, float* __restrict__ A, float* __restrict__ mask,
                                     float* __restrict__ emb, int n, int* __restrict__ err_flag) {
  const int64_t total = B * T * K;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % K);
    const int64_t bt = i / K;
    const int t = (int)(bt % T);
    const int64_t b = bt / T;
    const int64_t id = hist[b * hist_ld + t], ic = cand[b];
    const bool ok = id >= 0 && id < rows, okc = ic >= 0 && ic < rows;
    if ((!ok || !okc) && err_flag) atomicOr(err_flag, 1);
    const float kv = ok ? tab[id * K + c] : 0.f, qv = okc ? tab[ic * K + c] : 0.f;
    kbuf[i] = kv;
    float* a = A + bt * 4 * K;
    a[c] = qv; a[K + c] = kv; a[2 * K + c] = qv - kv; a[3 * K + c] = qv * kv;
    if (c == 0) mask[bt] = id != 0 ? 1.f : 0.f;
    if (t == 0) { qbuf[b * K + c] = qv; emb[b * n + K + c] = qv; }
  }
}
// pre += bias (kept for the backward pass), out = PReLU(pre) with alpha [alpha_rows, N] (row = r % alpha_rows: the position t for
// the attention unit's [T, h] alphas, 0 for the tower's [h])
__global__ void tr_bias_prelu_kernel(float* __restrict__ pre, const float* __restrict__ bias, const float* __restrict__ alpha,
                                     int64_t R, int N, int arows, float* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < R * N; i += (int64_t)gridDim.x * blockDim.x) {
    const int n = (int)(i % N);
    const int64_t r = i / N;
    const float p = pre[i] + bias[n];
    pre[i] = p;
    out[i] = p > 0.f ? p : alpha[(r % arows) * N + n] * p;
  }
}
// PReLU backward, one thread per alpha entry (t, n): g <- g * (pre > 0 ? 1 : alpha_old) over its rows r = b * arows + t, and
// alpha -= lr * sum_b g * min(pre, 0).  Rows in order: deterministic.
__global__ void tr_prelu_bwd_kernel(float* __restrict__ g, const float* __restrict__ pre, float* __restrict__ alpha, int64_t R, int N,
                                    int arows, float lr) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)arows * N) return;
  const int n = (int)(i % N);
  const int t = (int)(i / N);
  const float a = alpha[i];
  float ga = 0.f;
  for (int64_t r = t; r < R; r += arows) {
    const float p = pre[r * N + n], gg = g[r * N + n];
    if (p < 0.f) ga = fmaf(gg, p, ga);
    g[r * N + n] = p > 0.f ? gg : a * gg;
  }
  alpha[i] = a - lr * ga;
}
// scores -> masked softmax over the T positions -> att = sum_t pw_t k_t, written to emb[:, 0:K].  One warp per sample.
__global__ void __launch_bounds__(128)
tr_din_softmax_att_kernel(const float* __restrict__ s, const float* __restrict__ b3, const float* __restrict__ mask,
                          const float* __restrict__ kbuf, int T, int K, int64_t B, float* __restrict__ pw,
                          float* __restrict__ emb, int n) {
  TR_WARP_IDS();
  for (int64_t b = warp; b < B; b += nwarps) {
    float mx = -INFINITY;
    for (int t = lane; t < T; t += 32) {
      const float v = mask[b * T + t] != 0.f ? s[b * T + t] + b3[0] : -4294967295.0f;      // -2^32 + 1 (interaction.py:399)
      pw[b * T + t] = v;
      mx = fmaxf(mx, v);
    }
    for (int o = 16; o >= 1; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
    for (int t = lane; t < T; t += 32) { const float e = expf(pw[b * T + t] - mx); pw[b * T + t] = e; sum += e; }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    for (int t = lane; t < T; t += 32) pw[b * T + t] *= inv;
    __syncwarp();
    for (int c = lane; c < K; c += 32) {
      float a = 0.f;
      for (int t = 0; t < T; ++t) a = fmaf(pw[b * T + t], kbuf[(b * T + t) * K + c], a);
      emb[b * n + c] = a;
    }
  }
}
// one other-feature table into its columns of emb
__global__ void tr_din_other_gather_kernel(const float* __restrict__ tab, int64_t rows, int dim, const int32_t* __restrict__ ids,
                                           int64_t ids_ld, int64_t B, float* __restrict__ emb, int n, int* __restrict__ err_flag) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * dim; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / dim;
    const int c = (int)(i - b * dim);
    const int64_t id = ids[b * ids_ld];
    const bool ok = id >= 0 && id < rows;
    if (!ok && err_flag) atomicOr(err_flag, 1);
    emb[b * n + c] = ok ? tab[id * dim + c] : 0.f;
  }
}
__global__ void tr_din_other_sgd_kernel(float* __restrict__ tab, int64_t rows, int dim, const int32_t* __restrict__ ids,
                                        int64_t ids_ld, int64_t B, const float* __restrict__ gx, int n, float lr) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * dim; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / dim;
    const int c = (int)(i - b * dim);
    const int64_t id = ids[b * ids_ld];
    if (id >= 0 && id < rows) atomicAdd(tab + id * dim + c, -lr * gx[b * n + c]);
  }
}
__global__ void tr_copy_cols_kernel(const float* __restrict__ src, int64_t src_ld, int w, int64_t B, float* __restrict__ dst, int n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * w; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / w;
    dst[b * n + (i - b * w)] = src[b * src_ld + (i - b * w)];
  }
}
// attention pooling backward, one warp per sample: g_att = gx[b, 0:K];  g_p[t] = <g_att, k_t>;  gs[t] = mask ? pw (g_p - sum pw g_p)
// : 0  (tf.where passes no gradient to a masked score);  gk[b, t, :] = pw_t g_att  (non-zero on masked positions only when the
// whole history is masked and the softmax is uniform)
__global__ void __launch_bounds__(128)
tr_din_att_bwd_kernel(const float* __restrict__ gx, int n, const float* __restrict__ kbuf, const float* __restrict__ pw,
                      const float* __restrict__ mask, int T, int K, int64_t B, float* __restrict__ gs, float* __restrict__ gk) {
  TR_WARP_IDS();
  for (int64_t b = warp; b < B; b += nwarps) {
    float dot = 0.f;
    for (int t = lane; t < T; t += 32) {
      float gp = 0.f;
      for (int c = 0; c < K; ++c) gp = fmaf(gx[b * n + c], kbuf[(b * T + t) * K + c], gp);
      gs[b * T + t] = gp;
      dot = fmaf(pw[b * T + t], gp, dot);
    }
    dot = warp_sum(dot);
    for (int t = lane; t < T; t += 32)
      gs[b * T + t] = mask[b * T + t] != 0.f ? pw[b * T + t] * (gs[b * T + t] - dot) : 0.f;
    for (int i = lane; i < T * K; i += 32) gk[b * T * K + i] = pw[b * T + i / K] * gx[b * n + i % K];
  }
}
// A = [q, k, q - k, q * k] backward and the behaviour table's update, one thread per (sample, column):
// g_q = gx[b, K + c] + sum_t (gA0 + gA2 + gA3 k);  g_k = gk + gA1 - gA2 + gA3 q;  rows shared by samples / positions add up
__global__ void tr_din_concat_bwd_sgd_kernel(const float* __restrict__ gA, const float* __restrict__ gk, const float* __restrict__ gx,
                                             int n, const float* __restrict__ kbuf, const float* __restrict__ qbuf,
                                             const int32_t* __restrict__ hist, int64_t hist_ld, const int32_t* __restrict__ cand,
                                             int T, int K, int64_t B, int64_t rows, float lr, float* __restrict__ tab) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B * K; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / K;
    const int c = (int)(i - b * K);
    const float q = qbuf[i];
    float gq = gx[b * n + K + c];
    // a run of equal ids (the padding tail above all) is summed here and added once: T tiny adds into the padding row would each
    // round away against the row's own magnitude
    int64_t run_id = -1;
    float run = 0.f;
    for (int t = 0; t < T; ++t) {
      const int64_t bt = b * T + t;
      const float* a = gA + bt * 4 * K;
      const float kv = kbuf[bt * K + c];
      gq += a[c] + a[2 * K + c] + a[3 * K + c] * kv;
      const float g = gk[bt * K + c] + a[K + c] - a[2 * K + c] + a[3 * K + c] * q;
      const int64_t id = hist[b * hist_ld + t];
      if (id != run_id) {
        if (run != 0.f && run_id >= 0 && run_id < rows) atomicAdd(tab + run_id * K + c, -lr * run);
        run_id = id; run = 0.f;
      }
      run += g;
    }
    if (run != 0.f && run_id >= 0 && run_id < rows) atomicAdd(tab + run_id * K + c, -lr * run);
    const int64_t ic = cand[b];
    if (ic >= 0 && ic < rows) atomicAdd(tab + ic * K + c, -lr * gq);
  }
}

}  // namespace ctrhot

using namespace ctrhot;

namespace {
struct TrLayout { int64_t x, xv, fm, dnn, dz, t, s2, dx, g0, g1, h[TR_MAXL], total; int maxw; };
bool tr_layout(int64_t B, int n, int kf, const ctrhot_dense_desc* L, int n_layers, TrLayout& w) {
  if (!L || n_layers < 2 || n_layers > TR_MAXL || B < 0 || n <= 0 || kf < 1 || kf > 32) return false;
  if (L[0].in_dim != n || L[n_layers - 1].out_dim != 1) return false;
  int64_t o = 0;
  auto take = [&](int64_t cnt) { const int64_t at = o; o += (cnt + 3) / 4 * 4; return at; };
  w.maxw = n;
  for (int l = 0; l < n_layers; ++l) {
    if (l > 0 && L[l].in_dim != L[l - 1].out_dim) return false;
    if (l < n_layers - 1 && L[l].act != CTRHOT_ACT_RELU) return false;
    if (!L[l].weight) return false;
    if (L[l].out_dim > w.maxw) w.maxw = L[l].out_dim;
  }
  w.x = take(B * n); w.xv = take(B * kf); w.fm = take(B); w.dnn = take(B); w.dz = take(B); w.t = take(B * kf);
  w.s2 = take(n); w.dx = take(B * n); w.g0 = take(B * (int64_t)w.maxw); w.g1 = take(B * (int64_t)w.maxw);
  for (int l = 0; l < n_layers - 1; ++l) w.h[l] = take(B * (int64_t)L[l].out_dim);
  w.total = o;
  return true;
}
inline int tr_grid(int64_t total, int block) {
  int64_t g = (total + block - 1) / block;
  if (g > (int64_t)kNumSMs * 32) g = (int64_t)kNumSMs * 32;
  return g < 1 ? 1 : (int)g;
}
// column sums: beta == 1 over thousands of rows -> row slices added with atomics (one or a few CTAs would walk the whole batch)
inline void tr_colsum(cudaStream_t st, int gx, const float* A, int64_t lda, const float* w, int64_t B, int N, int square, float* out,
                      float alpha, float beta) {
  dim3 grid(gx, 1);
  if (beta == 1.f && B >= 2048) { const int64_t y = B / 512; grid.y = (unsigned)(y < 32 ? y : 32); }
  tr_colsum_kernel<<<grid, 1024, 0, st>>>(A, lda, w, B, N, square, out, alpha, beta);
}
inline void tr_gemm(cudaStream_t st, const float* A, int64_t am, int64_t ak, const float* Bm, int64_t bk, int64_t bn, float* C,
                    int64_t ldc, int M, int N, int K, float alpha, float beta, float* csum = nullptr, float csum_alpha = 0.f) {
  const bool big = M >= 128 && N >= 64 && K >= 64;
  const int T = big ? TR_T64 : TR_T;
  dim3 grid((N + T - 1) / T, (M + T - 1) / T);
  // the weight-gradient products (W += -lr X^T G: a few tiles of output, the batch as the k range) would walk thousands of rows
  // in a handful of CTAs: split the k range over enough CTAs for two per SM
  const int tiles = (int)(grid.x * grid.y);
  if (beta == 1.f && K >= 16 * TR_T && tiles < 2 * kNumSMs) {
    int z = (2 * kNumSMs + tiles - 1) / tiles;
    int zmax = K / (4 * TR_T);                                     // at least 4 k-tiles per slice,
    if (zmax > 32) zmax = 32;                                      // at most 32 roundings at the weight's own magnitude
    grid.z = z < zmax ? z : zmax;
  }
  // csum (B must be [K, N] row-major: bn == 1): fused in the tiled kernels, a separate column sum beside the small-M one
  if (M <= 128 && K >= 64 && grid.z == 1) {
    tr_gemm_m32_kernel<<<dim3((N + 31) / 32, (M + 31) / 32), 256, 0, st>>>(A, am, ak, Bm, bk, bn, C, ldc, M, N, K, alpha, beta, TrEpi{});
    if (csum) tr_colsum(st, (N + 31) / 32, Bm, bk, nullptr, K, N, 0, csum, csum_alpha, 1.f);
    return;
  }
  if (big) {
    tr_gemm64_kernel<<<grid, 256, 0, st>>>(A, am, ak, Bm, bk, bn, C, ldc, M, N, K, alpha, beta, csum, csum_alpha);
  } else {
    tr_gemm_kernel<<<grid, 256, 0, st>>>(A, am, ak, Bm, bk, bn, C, ldc, M, N, K, alpha, beta, csum, csum_alpha);
  }
}
inline bool tr_m32_ok(int M, int K) { return M <= 128 && K >= 64; }
// One Dense layer of a tower, forward: h = dropout(relu(in W + bias)).  At the reference's batch (M <= 32) the bias / ReLU / dropout
// ride in the product's epilogue -- one launch instead of two on a chain where every link costs 3-5 us.
inline void tr_dense_fwd(cudaStream_t st, const float* in, int in_dim, const float* W, const float* bias, int N, float* h, int M,
                         int relu, float rate, uint64_t seed, uint32_t layer, const TrCtl* ctl) {
  if (tr_m32_ok(M, in_dim)) {
    TrEpi e{bias, ctl, seed, rate, layer, relu, 1};
    tr_gemm_m32_kernel<<<dim3((N + 31) / 32, (M + 31) / 32), 256, 0, st>>>(in, in_dim, 1, W, N, 1, h, N, M, N, in_dim, 1.f, 0.f, e);
    return;
  }
  tr_gemm(st, in, in_dim, 1, W, N, 1, h, N, M, N, in_dim, 1.f, 0.f);
  tr_bias_relu_drop_kernel<<<tr_grid((int64_t)M * N, 256), 256, 0, st>>>(h, bias, N, M, relu, rate, seed, layer, ctl);
}
// Gradient w.r.t. a Dense layer's input through the OLD weights, dst [M, Kd] = g [M, N] W^T; when hprev (the activations of the
// layer below, [M, Kd]) is given and the small-M product runs, that layer's ReLU / dropout derivative is applied in the epilogue
// (what tr_relu_drop_bwd_kernel would do as the next launch) -- the loops test the same condition before launching that kernel.
inline void tr_dense_bwd_input(cudaStream_t st, const float* g, int N, const float* W, float* dst, int Kd, int M, const float* hprev,
                               float rate, uint64_t seed, uint32_t layer, const TrCtl* ctl) {
  if (hprev && tr_m32_ok(M, N)) {
    TrEpi e{hprev, ctl, seed, rate, layer, 1, 2};
    tr_gemm_m32_kernel<<<dim3((Kd + 31) / 32, (M + 31) / 32), 256, 0, st>>>(g, N, 1, W, 1, N, dst, Kd, M, Kd, N, 1.f, 0.f, e);
    return;
  }
  tr_gemm(st, g, N, 1, W, 1, N, dst, Kd, M, Kd, N, 1.f, 0.f);
}
// FM backward and update with d loss / d fm = gscale * dz:  w0, w1 = (1 - 2 lr reg_w) w1 - lr x^T dfm, and v with its -x^2 v part
// folded into a per-row scale (old w1 / v are used for dx, which is formed first; dx may be NULL)
static void tr_fm_update(cudaStream_t st, const float* x, int n, int64_t B, const float* dz, float gscale, float* fm_w0,
                         float* fm_w1, float* fm_v, int k_fm, float reg_w, float reg_v, float lr, const float* xv, float* t,
                         float* s2, float* dx) {
  const int M = (int)B;
  const int nb = dx ? n : k_fm;                            // without dx only t = dfm * xv is formed: B x k_fm elements
  tr_fm_bwd_kernel<<<tr_grid(B * nb, 256), 256, 0, st>>>(x, nb, B, dz, fm_w1, fm_v, k_fm, xv, dx, t, gscale);
  tr_colsum(st, 1, dz, 1, nullptr, B, 1, 0, fm_w0, -gscale * lr, 1.f);
  tr_colsum(st, (n + 31) / 32, x, n, dz, B, n, 0, fm_w1, -gscale * lr, 1.f - 2.f * lr * reg_w);
  tr_colsum(st, (n + 31) / 32, x, n, dz, B, n, 1, s2, gscale, 0.f);   // s2[j] = sum_b dfm_b x_bj^2
  tr_scale_rows_kernel<<<tr_grid((int64_t)n * k_fm, 256), 256, 0, st>>>(fm_v, n, k_fm, s2, 1.f - 2.f * lr * reg_v, lr);
  tr_gemm(st, x, 1, n, t, k_fm, 1, fm_v, k_fm, n, k_fm, M, -lr, 1.f);                      // v -= lr * x^T (dfm xv)
}
}  // namespace

extern "C" int64_t ctrhot_deepfm_train_workspace_bytes(int64_t B, int n, int k_fm, const ctrhot_dense_desc* h_layers,
                                                       int n_layers) {
  TrLayout w;
  if (!tr_layout(B, n, k_fm, h_layers, n_layers, w)) return CTRHOT_ESHAPE;
  return w.total * (int64_t)sizeof(float);
}

static int tr_step(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                   int64_t ids_ld, const float* dense, int64_t dense_ld, int D, const float* y,
                   int64_t B, float* fm_w0, float* fm_w1, float* fm_v, int k_fm, float reg_w,
                   float reg_v, const ctrhot_dense_desc* h_layers, int n_layers, float lr,
                   float dropout, uint64_t seed, float* loss_out, void* workspace,
                   int64_t workspace_bytes, int* err_flag, TrCtl* ctl, void* stream) {
  if (!arena || !row_off || !ids || !y || !fm_w0 || !fm_w1 || !fm_v || !loss_out || !workspace || F <= 0 || k <= 0 || D < 0 ||
      (D > 0 && !dense) || B <= 0 || ids_ld < F || (D > 0 && dense_ld < D)) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (!(dropout >= 0.f && dropout < 1.f) || !(lr >= 0.f)) return CTRHOT_EARG;
  const int n = D + F * k;
  TrLayout w;
  if (!tr_layout(B, n, k_fm, h_layers, n_layers, w)) return CTRHOT_ESHAPE;
  if (workspace_bytes < w.total * (int64_t)sizeof(float) || !aligned16(workspace)) return CTRHOT_EARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* ws = static_cast<float*>(workspace);
  float *x = ws + w.x, *xv = ws + w.xv, *fm = ws + w.fm, *dnn = ws + w.dnn, *dz = ws + w.dz, *t = ws + w.t, *s2 = ws + w.s2,
        *dx = ws + w.dx;
  const int NH = n_layers - 1;                                   // hidden layers; layer NH is Dense(h -> 1)
  const int M = (int)B;
  // ---- forward
  if (ids_kind == CTRHOT_IDS_I32)
    tr_gather_kernel<CTRHOT_IDS_I32><<<tr_grid(B * n, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, dense, dense_ld, D, B, x, n, err_flag, ctl);
  else
    tr_gather_kernel<CTRHOT_IDS_F32><<<tr_grid(B * n, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, dense, dense_ld, D, B, x, n, err_flag, ctl);
  tr_fm_fwd_kernel<<<(int)((B + 7) / 8), 256, 0, st>>>(x, n, B, fm_w0, fm_w1, fm_v, k_fm, xv, fm);
  const float* in = x;
  int in_dim = n;
  for (int l = 0; l < NH; ++l) {
    float* h = ws + w.h[l];
    const int N = h_layers[l].out_dim;
    tr_dense_fwd(st, in, in_dim, h_layers[l].weight, h_layers[l].bias, N, h, M, 1, dropout, seed, (uint32_t)l, ctl);
    in = h; in_dim = N;
  }
  tr_gemm(st, in, in_dim, 1, h_layers[NH].weight, 1, 1, dnn, 1, M, 1, in_dim, 1.f, 0.f);
  tr_loss_kernel<<<1, 256, 0, st>>>(fm, dnn, h_layers[NH].bias, y, B, dz, loss_out, ctl, 0.5f);
  // ---- backward through the tower: g = dLoss/d(pre-activation) of the current layer, [B, out_dim]
  float* g = ws + w.g0;
  float* gn = ws + w.g1;
  // d dnn = 1/2 dz; Dense(h -> 1): g_prev = 1/2 dz w_out^T (before w_out moves), then w_out, b_out
  {
    float* wo = const_cast<float*>(h_layers[NH].weight);
    tr_gemm(st, dz, 1, 1, wo, 0, 1, gn, in_dim, M, in_dim, 1, 0.5f, 0.f);                 // [B,1] x [1,h]: B(k=0, n) = wo[n]
    tr_gemm(st, in, 1, in_dim, dz, 1, 1, wo, 1, in_dim, 1, M, -0.5f * lr, 1.f);           // w_out -= lr * h^T (dz / 2)
    if (h_layers[NH].bias)
      tr_colsum(st, 1, dz, 1, nullptr, B, 1, 0, const_cast<float*>(h_layers[NH].bias), -0.5f * lr, 1.f);
    float* tmp = g; g = gn; gn = tmp;
  }
  for (int l = NH - 1; l >= 0; --l) {
    const int N = h_layers[l].out_dim, Kd = h_layers[l].in_dim;
    const float* h = ws + w.h[l];
    const float* hin = l > 0 ? ws + w.h[l - 1] : x;
    float* W = const_cast<float*>(h_layers[l].weight);
    if (!(l < NH - 1 && tr_m32_ok(M, h_layers[l + 1].out_dim)))   // (else the product above applied this mask in its epilogue)
      tr_relu_drop_bwd_kernel<<<tr_grid(B * N, 256), 256, 0, st>>>(g, h, B * N, 1, dropout, seed, (uint32_t)l, ctl);
    // gradient w.r.t. this layer's input with the OLD weights: [B,N] x [N,Kd]  (B(k, n) = W[n * N + k])
    float* dst = l > 0 ? gn : dx;
    tr_dense_bwd_input(st, g, N, W, dst, Kd, M, l > 0 ? ws + w.h[l - 1] : nullptr, dropout, seed, (uint32_t)(l - 1), ctl);
    tr_gemm(st, hin, 1, Kd, g, N, 1, W, N, Kd, N, M, -lr, 1.f, const_cast<float*>(h_layers[l].bias), -lr);                           // W -= lr * in^T g
    if (l > 0) { float* tmp = g; g = gn; gn = tmp; }
  }
  // ---- FM backward (old w1, v), then its updates
  tr_fm_update(st, x, n, B, dz, 0.5f, fm_w0, fm_w1, fm_v, k_fm, reg_w, reg_v, lr, xv, t, s2, dx);
  // ---- embedding rows
  if (ids_kind == CTRHOT_IDS_I32)
    tr_embed_sgd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * F * k, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, D, B, dx, n, lr, ctl);
  else
    tr_embed_sgd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * F * k, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, D, B, dx, n, lr, ctl);
  if (ctl) tr_advance_kernel<<<1, 1, 0, st>>>(ctl, (long long)B);
  CTRHOT_LAUNCH_OK("ctrhot_deepfm_train_step_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_deepfm_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                                            int64_t ids_ld, const float* dense, int64_t dense_ld, int D, const float* y,
                                            int64_t B, float* fm_w0, float* fm_w1, float* fm_v, int k_fm, float reg_w,
                                            float reg_v, const ctrhot_dense_desc* h_layers, int n_layers, float lr,
                                            float dropout, uint64_t seed, float* loss_out, void* workspace,
                                            int64_t workspace_bytes, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_deepfm_train_step_f32");
  return tr_step(arena, row_off, F, k, ids, ids_kind, ids_ld, dense, dense_ld, D, y, B, fm_w0, fm_w1, fm_v, k_fm, reg_w, reg_v,
                 h_layers, n_layers, lr, dropout, seed, loss_out, workspace, workspace_bytes, err_flag, nullptr, stream);
}

extern "C" int ctrhot_deepfm_train_step_ctl_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids,
                                                int ids_kind, int64_t ids_ld, const float* dense, int64_t dense_ld, int D,
                                                const float* y, int64_t B, float* fm_w0, float* fm_w1, float* fm_v, int k_fm,
                                                float reg_w, float reg_v, const ctrhot_dense_desc* h_layers, int n_layers,
                                                float lr, float dropout, uint64_t seed, float* loss_out, void* workspace,
                                                int64_t workspace_bytes, int* err_flag, void* d_ctl, void* stream) {
  CTRHOT_RANGE("ctrhot_deepfm_train_step_ctl_f32");
  if (!d_ctl || (reinterpret_cast<uintptr_t>(d_ctl) & 7u)) return CTRHOT_EARG;
  return tr_step(arena, row_off, F, k, ids, ids_kind, ids_ld, dense, dense_ld, D, y, B, fm_w0, fm_w1, fm_v, k_fm, reg_w, reg_v,
                 h_layers, n_layers, lr, dropout, seed, loss_out, workspace, workspace_bytes, err_flag,
                 static_cast<TrCtl*>(d_ctl), stream);
}


// ---- FM (DL/model/fm.py:14-23: p = sigmoid(FMLayer(x)) on the dense one-hot matrix) under compile_fit -----------------------
extern "C" int64_t ctrhot_fm_train_workspace_bytes(int64_t B, int n, int k_fm) {
  if (B < 0 || n <= 0 || k_fm < 1 || k_fm > 32) return CTRHOT_ESHAPE;
  return (2 * (B * k_fm + 3) / 4 * 4 + 2 * ((B + 3) / 4 * 4) + (n + 3) / 4 * 4) * (int64_t)sizeof(float);
}

extern "C" int ctrhot_fm_train_step_f32(const float* x, int64_t x_ld, int n, const float* y, int64_t B, float* fm_w0,
                                        float* fm_w1, float* fm_v, int k_fm, float reg_w, float reg_v, float lr,
                                        float* loss_out, void* workspace, int64_t workspace_bytes, void* stream) {
  CTRHOT_RANGE("ctrhot_fm_train_step_f32");
  if (!x || !y || !fm_w0 || !fm_w1 || !fm_v || !loss_out || !workspace || B <= 0 || n <= 0 || x_ld != n || !(lr >= 0.f))
    return CTRHOT_EARG;                                     // (dense rows only: x_ld == n)
  if (k_fm < 1 || k_fm > 32) return CTRHOT_ESHAPE;
  if (workspace_bytes < ctrhot_fm_train_workspace_bytes(B, n, k_fm) || !aligned16(workspace)) return CTRHOT_EARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* ws = static_cast<float*>(workspace);
  const int64_t bk = (B * k_fm + 3) / 4 * 4, b4 = (B + 3) / 4 * 4;
  float *xv = ws, *t = ws + bk, *fm = ws + 2 * bk, *dz = fm + b4, *s2 = dz + b4;
  tr_fm_fwd_kernel<<<(int)((B + 7) / 8), 256, 0, st>>>(x, n, B, fm_w0, fm_w1, fm_v, k_fm, xv, fm);
  tr_loss_kernel<<<1, 256, 0, st>>>(fm, nullptr, nullptr, y, B, dz, loss_out, nullptr, 1.0f);
  tr_fm_update(st, x, n, B, dz, 1.0f, fm_w0, fm_w1, fm_v, k_fm, reg_w, reg_v, lr, xv, t, s2, nullptr);
  CTRHOT_LAUNCH_OK("ctrhot_fm_train_step_f32");
  return CTRHOT_OK;
}

// ---------------------------------------------------------------- DCN training step (DL/model/dcn.py under compile_fit)
namespace {
struct TrDcnLayout { int64_t x, XL, S, G, C, zc, zt, d, dz, gd, dx, g0, g1, h[TR_MAXL], total; int maxw; };
bool tr_dcn_layout(int64_t B, int n, int L, const ctrhot_dense_desc* Ld, int n_layers, TrDcnLayout& w) {
  if (!Ld || n_layers < 2 || n_layers > TR_MAXL || B < 0 || n <= 0 || n > 32 * TR_NJ || L < 1 || L > TR_MAXL) return false;
  if (Ld[0].in_dim != n || Ld[n_layers - 1].out_dim != 1) return false;
  int64_t o = 0;
  auto take = [&](int64_t cnt) { const int64_t at = o; o += (cnt + 3) / 4 * 4; return at; };
  w.maxw = n;
  for (int l = 0; l < n_layers; ++l) {
    if (l > 0 && Ld[l].in_dim != Ld[l - 1].out_dim) return false;
    if (l < n_layers - 1 && Ld[l].act != CTRHOT_ACT_RELU) return false;
    if (!Ld[l].weight) return false;
    if (Ld[l].out_dim > w.maxw) w.maxw = Ld[l].out_dim;
  }
  w.x = take(B * n); w.XL = take((int64_t)L * B * n); w.S = take((int64_t)L * B); w.G = take((int64_t)L * B * n);
  w.C = take((int64_t)L * B); w.zc = take(B); w.zt = take(B); w.d = take(B); w.dz = take(B); w.gd = take(B); w.dx = take(B * n);
  w.g0 = take(B * (int64_t)w.maxw); w.g1 = take(B * (int64_t)w.maxw);
  for (int l = 0; l < n_layers - 1; ++l) w.h[l] = take(B * (int64_t)Ld[l].out_dim);
  w.total = o;
  return true;
}
}  // namespace

extern "C" int64_t ctrhot_dcn_train_workspace_bytes(int64_t B, int n, int L, const ctrhot_dense_desc* h_layers, int n_layers) {
  TrDcnLayout w;
  if (!tr_dcn_layout(B, n, L, h_layers, n_layers, w)) return CTRHOT_ESHAPE;
  return w.total * (int64_t)sizeof(float);
}

extern "C" int ctrhot_dcn_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                                         int64_t ids_ld, const float* dense, int64_t dense_ld, int D, const float* y, int64_t B,
                                         const int64_t* h_cross_w, const int64_t* h_cross_b, int L, float reg_w, float reg_b,
                                         const ctrhot_dense_desc* h_layers, int n_layers, float* final_k, float* final_b,
                                         float lr, float dropout, uint64_t seed, float* loss_out, void* workspace,
                                         int64_t workspace_bytes, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_dcn_train_step_f32");
  if (!arena || !row_off || !ids || !y || !h_cross_w || !h_cross_b || !final_k || !final_b || !loss_out || !workspace || F <= 0 ||
      k <= 0 || D < 0 || (D > 0 && !dense) || B <= 0 || ids_ld < F || (D > 0 && dense_ld < D)) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (!(dropout >= 0.f && dropout < 1.f) || !(lr >= 0.f)) return CTRHOT_EARG;
  const int n = D + F * k;
  TrDcnLayout w;
  if (!tr_dcn_layout(B, n, L, h_layers, n_layers, w)) return CTRHOT_ESHAPE;
  if (workspace_bytes < w.total * (int64_t)sizeof(float) || !aligned16(workspace)) return CTRHOT_EARG;
  TrCross cw;
  for (int l = 0; l < L; ++l) {
    cw.w[l] = reinterpret_cast<const float*>((uintptr_t)h_cross_w[l]);
    cw.b[l] = reinterpret_cast<const float*>((uintptr_t)h_cross_b[l]);
    if (!cw.w[l] || !cw.b[l]) return CTRHOT_EARG;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* ws = static_cast<float*>(workspace);
  float *x = ws + w.x, *XL = ws + w.XL, *S = ws + w.S, *G = ws + w.G, *C = ws + w.C, *zc = ws + w.zc, *zt = ws + w.zt,
        *d = ws + w.d, *dz = ws + w.dz, *gd = ws + w.gd, *dx = ws + w.dx;
  const int NH = n_layers - 1, M = (int)B;
  const int wgrid = (int)((B + 7) / 8);
  // ---- forward: x, cross layers, tower, logit = x_L . fk[0:n] + d fk[n] + fb
  if (ids_kind == CTRHOT_IDS_I32)
    tr_gather_kernel<CTRHOT_IDS_I32><<<tr_grid(B * n, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, dense, dense_ld, D, B, x, n, err_flag, nullptr);
  else
    tr_gather_kernel<CTRHOT_IDS_F32><<<tr_grid(B * n, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, dense, dense_ld, D, B, x, n, err_flag, nullptr);
  tr_cross_fwd_kernel<<<wgrid, 256, 0, st>>>(x, n, B, L, cw, final_k, XL, S, zc);
  const float* in = x;
  int in_dim = n;
  for (int l = 0; l < NH; ++l) {
    float* h = ws + w.h[l];
    const int N = h_layers[l].out_dim;
    tr_dense_fwd(st, in, in_dim, h_layers[l].weight, h_layers[l].bias, N, h, M, 1, dropout, seed, (uint32_t)l, nullptr);
    in = h; in_dim = N;
  }
  tr_gemm(st, in, in_dim, 1, h_layers[NH].weight, 1, 1, d, 1, M, 1, in_dim, 1.f, 0.f);
  {
    // d[b] += b_d  (one element per sample; the tower's last bias), then zt = zc + d * fk[n]
    if (h_layers[NH].bias) tr_bias_relu_drop_kernel<<<tr_grid(B, 256), 256, 0, st>>>(d, h_layers[NH].bias, 1, B, 0, 0.f, 0, 0, nullptr);
    tr_axpy_scalar_kernel<<<tr_grid(B, 256), 256, 0, st>>>(zc, d, final_k + n, B, zt);
  }
  tr_loss_kernel<<<1, 256, 0, st>>>(zt, nullptr, final_b, y, B, dz, loss_out, nullptr, 1.0f);
  tr_axpy_scalar_kernel<<<tr_grid(B, 256), 256, 0, st>>>(nullptr, dz, final_k + n, B, gd);      // d loss / d d = dz fk[n] (old fk)
  // ---- backward through the tower from gd
  float* g = ws + w.g0;
  float* gn = ws + w.g1;
  {
    float* wo = const_cast<float*>(h_layers[NH].weight);
    tr_gemm(st, gd, 1, 1, wo, 0, 1, gn, in_dim, M, in_dim, 1, 1.f, 0.f);
    tr_gemm(st, in, 1, in_dim, gd, 1, 1, wo, 1, in_dim, 1, M, -lr, 1.f);
    if (h_layers[NH].bias)
      tr_colsum(st, 1, gd, 1, nullptr, B, 1, 0, const_cast<float*>(h_layers[NH].bias), -lr, 1.f);
    float* tmp = g; g = gn; gn = tmp;
  }
  for (int l = NH - 1; l >= 0; --l) {
    const int N = h_layers[l].out_dim, Kd = h_layers[l].in_dim;
    const float* h = ws + w.h[l];
    const float* hin = l > 0 ? ws + w.h[l - 1] : x;
    float* W = const_cast<float*>(h_layers[l].weight);
    if (!(l < NH - 1 && tr_m32_ok(M, h_layers[l + 1].out_dim)))   // (else the product above applied this mask in its epilogue)
      tr_relu_drop_bwd_kernel<<<tr_grid(B * N, 256), 256, 0, st>>>(g, h, B * N, 1, dropout, seed, (uint32_t)l, nullptr);
    float* dst = l > 0 ? gn : dx;
    tr_dense_bwd_input(st, g, N, W, dst, Kd, M, l > 0 ? ws + w.h[l - 1] : nullptr, dropout, seed, (uint32_t)(l - 1), nullptr);
    tr_gemm(st, hin, 1, Kd, g, N, 1, W, N, Kd, N, M, -lr, 1.f, const_cast<float*>(h_layers[l].bias), -lr);
    if (l > 0) { float* tmp = g; g = gn; gn = tmp; }
  }
  // ---- cross layers backward (old w_l, old fk), dx += its share; then the updates of w_l, b_l and of the final Dense
  tr_cross_bwd_kernel<<<wgrid, 256, 0, st>>>(x, n, B, L, cw, final_k, dz, S, G, C, dx);
  for (int l = 0; l < L; ++l) {
    const float* xin = l > 0 ? XL + (int64_t)(l - 1) * B * n : x;
    tr_gemm(st, xin, 1, n, C + (int64_t)l * B, 1, 1, const_cast<float*>(cw.w[l]), 1, n, 1, M, -lr, 1.f - 2.f * lr * reg_w);
    tr_colsum(st, (n + 31) / 32, G + (int64_t)l * B * n, n, nullptr, B, n, 0, const_cast<float*>(cw.b[l]), -lr,
                                                     1.f - 2.f * lr * reg_b);
  }
  tr_gemm(st, XL + (int64_t)(L - 1) * B * n, 1, n, dz, 1, 1, final_k, 1, n, 1, M, -lr, 1.f);   // fk[0:n] -= lr x_L^T dz
  tr_colsum(st, 1, d, 1, dz, B, 1, 0, final_k + n, -lr, 1.f);              // fk[n]   -= lr sum_b d dz
  tr_colsum(st, 1, dz, 1, nullptr, B, 1, 0, final_b, -lr, 1.f);
  // ---- embedding rows
  if (ids_kind == CTRHOT_IDS_I32)
    tr_embed_sgd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * F * k, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, D, B, dx, n, lr, nullptr);
  else
    tr_embed_sgd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * F * k, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, D, B, dx, n, lr, nullptr);
  CTRHOT_LAUNCH_OK("ctrhot_dcn_train_step_f32");
  return CTRHOT_OK;
}

// ---------------------------------------------------------------- PNN training step (DL/model/pnn.py:74-85)
namespace {
struct TrPnnLayout { int64_t x, z, dz, dx, g0, g1, h[TR_MAXL], total; int maxw, n, P; };
bool tr_pnn_layout(int64_t B, int F, int k, int mode, const ctrhot_dense_desc* Ld, int n_layers, TrPnnLayout& w) {
  if (!Ld || n_layers < 2 || n_layers > TR_MAXL || B < 0 || F < 2 || k < 1 || mode < 1 || mode > 3) return false;
  w.P = F * (F - 1) / 2;
  w.n = F * k + w.P * (mode == 3 ? 2 : 1);
  if (Ld[0].in_dim != w.n || Ld[n_layers - 1].out_dim != 1) return false;
  int64_t o = 0;
  auto take = [&](int64_t cnt) { const int64_t at = o; o += (cnt + 3) / 4 * 4; return at; };
  w.maxw = w.n;
  for (int l = 0; l < n_layers; ++l) {
    if (l > 0 && Ld[l].in_dim != Ld[l - 1].out_dim) return false;
    if (l < n_layers - 1 && Ld[l].act != CTRHOT_ACT_RELU) return false;
    if (!Ld[l].weight) return false;
    if (Ld[l].out_dim > w.maxw) w.maxw = Ld[l].out_dim;
  }
  w.x = take(B * w.n); w.z = take(B); w.dz = take(B); w.dx = take(B * w.n);
  w.g0 = take(B * (int64_t)w.maxw); w.g1 = take(B * (int64_t)w.maxw);
  for (int l = 0; l < n_layers - 1; ++l) w.h[l] = take(B * (int64_t)Ld[l].out_dim);
  w.total = o;
  return true;
}
}  // namespace

extern "C" int64_t ctrhot_pnn_train_workspace_bytes(int64_t B, int F, int k, int mode, const ctrhot_dense_desc* h_layers,
                                                    int n_layers) {
  TrPnnLayout w;
  if (!tr_pnn_layout(B, F, k, mode, h_layers, n_layers, w)) return CTRHOT_ESHAPE;
  return w.total * (int64_t)sizeof(float);
}

extern "C" int ctrhot_pnn_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                                         int64_t ids_ld, const float* y, int64_t B, float* outer_w, int mode,
                                         const ctrhot_dense_desc* h_layers, int n_layers, float lr, float dropout,
                                         uint64_t seed, float* loss_out, void* workspace, int64_t workspace_bytes,
                                         int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_pnn_train_step_f32");
  if (!arena || !row_off || !ids || !y || !loss_out || !workspace || B <= 0 || ids_ld < F) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (!(dropout >= 0.f && dropout < 1.f) || !(lr >= 0.f) || ((mode & 2) && !outer_w)) return CTRHOT_EARG;
  TrPnnLayout w;
  if (!tr_pnn_layout(B, F, k, mode, h_layers, n_layers, w)) return CTRHOT_ESHAPE;
  if (workspace_bytes < w.total * (int64_t)sizeof(float) || !aligned16(workspace)) return CTRHOT_EARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* ws = static_cast<float*>(workspace);
  float *x = ws + w.x, *z = ws + w.z, *dz = ws + w.dz, *dx = ws + w.dx;
  const int n = w.n, P = w.P, fk = F * k, NH = n_layers - 1, M = (int)B;
  const int col_in = (mode & 1) ? fk : -1, col_out = (mode & 2) ? fk + ((mode & 1) ? P : 0) : -1;
  // ---- forward: x = [rows | inner | outer] (exact fp32 product kernels), tower, raw output
  if (ids_kind == CTRHOT_IDS_I32)
    tr_gather_kernel<CTRHOT_IDS_I32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, nullptr, 0, 0, B, x, fk, err_flag, nullptr, n);
  else
    tr_gather_kernel<CTRHOT_IDS_F32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, nullptr, 0, 0, B, x, fk, err_flag, nullptr, n);
  int rc = CTRHOT_OK;
  if (mode & 1) rc = ctrhot_pnn_inner_f32(x, n, F, k, B, x + col_in, n, stream);
  if (rc != CTRHOT_OK) return rc;
  if (mode & 2) rc = ctrhot_pnn_outer_f32(x, n, F, k, B, outer_w, x + col_out, n, stream);
  if (rc != CTRHOT_OK) return rc;
  const float* in = x;
  int in_dim = n;
  for (int l = 0; l < NH; ++l) {
    float* h = ws + w.h[l];
    const int N = h_layers[l].out_dim;
    tr_dense_fwd(st, in, in_dim, h_layers[l].weight, h_layers[l].bias, N, h, M, 1, dropout, seed, (uint32_t)l, nullptr);
    in = h; in_dim = N;
  }
  tr_gemm(st, in, in_dim, 1, h_layers[NH].weight, 1, 1, z, 1, M, 1, in_dim, 1.f, 0.f);
  tr_raw_bce_kernel<<<1, 256, 0, st>>>(z, h_layers[NH].bias, y, B, dz, loss_out);
  // ---- backward through the tower
  float* g = ws + w.g0;
  float* gn = ws + w.g1;
  {
    float* wo = const_cast<float*>(h_layers[NH].weight);
    tr_gemm(st, dz, 1, 1, wo, 0, 1, gn, in_dim, M, in_dim, 1, 1.f, 0.f);
    tr_gemm(st, in, 1, in_dim, dz, 1, 1, wo, 1, in_dim, 1, M, -lr, 1.f);
    if (h_layers[NH].bias)
      tr_colsum(st, 1, dz, 1, nullptr, B, 1, 0, const_cast<float*>(h_layers[NH].bias), -lr, 1.f);
    float* tmp = g; g = gn; gn = tmp;
  }
  for (int l = NH - 1; l >= 0; --l) {
    const int N = h_layers[l].out_dim, Kd = h_layers[l].in_dim;
    const float* h = ws + w.h[l];
    const float* hin = l > 0 ? ws + w.h[l - 1] : x;
    float* W = const_cast<float*>(h_layers[l].weight);
    if (!(l < NH - 1 && tr_m32_ok(M, h_layers[l + 1].out_dim)))   // (else the product above applied this mask in its epilogue)
      tr_relu_drop_bwd_kernel<<<tr_grid(B * N, 256), 256, 0, st>>>(g, h, B * N, 1, dropout, seed, (uint32_t)l, nullptr);
    float* dst = l > 0 ? gn : dx;
    tr_dense_bwd_input(st, g, N, W, dst, Kd, M, l > 0 ? ws + w.h[l - 1] : nullptr, dropout, seed, (uint32_t)(l - 1), nullptr);
    tr_gemm(st, hin, 1, Kd, g, N, 1, W, N, Kd, N, M, -lr, 1.f, const_cast<float*>(h_layers[l].bias), -lr);
    if (l > 0) { float* tmp = g; g = gn; gn = tmp; }
  }
  // ---- product layers backward into dx's flat columns (old W), then W, then the embedding rows
  tr_pnn_prod_bwd_kernel<<<tr_grid(B * fk, 256), 256, 0, st>>>(x, n, B, F, k, P, outer_w, col_in, col_out, dx);
  if (mode & 2) tr_pnn_dw_kernel<<<tr_grid((int64_t)k * P * k, 128), 128, 0, st>>>(x, n, B, F, k, P, dx, col_out, outer_w, lr);
  if (ids_kind == CTRHOT_IDS_I32)
    tr_embed_sgd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, 0, B, dx, n, lr, nullptr);
  else
    tr_embed_sgd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, 0, B, dx, n, lr, nullptr);
  CTRHOT_LAUNCH_OK("ctrhot_pnn_train_step_f32");
  return CTRHOT_OK;
}

// ---------------------------------------------------------------- Wide&Deep training step
namespace {
struct TrWdLayout { int64_t x, wide, dnn, dz, dx, g0, g1, h[TR_MAXL], total; int maxw; };
bool tr_wd_layout(int64_t B, int n, const ctrhot_dense_desc* Ld, int n_layers, TrWdLayout& w) {
  if (!Ld || n_layers < 2 || n_layers > TR_MAXL || B < 0 || n <= 0) return false;
  if (Ld[0].in_dim != n || Ld[n_layers - 1].out_dim != 1) return false;
  int64_t o = 0;
  auto take = [&](int64_t cnt) { const int64_t at = o; o += (cnt + 3) / 4 * 4; return at; };
  w.maxw = n;
  for (int l = 0; l < n_layers; ++l) {
    if (l > 0 && Ld[l].in_dim != Ld[l - 1].out_dim) return false;
    if (l < n_layers - 1 && Ld[l].act != CTRHOT_ACT_RELU) return false;
    if (!Ld[l].weight) return false;
    if (Ld[l].out_dim > w.maxw) w.maxw = Ld[l].out_dim;
  }
  w.x = take(B * n); w.wide = take(B); w.dnn = take(B); w.dz = take(B); w.dx = take(B * n);
  w.g0 = take(B * (int64_t)w.maxw); w.g1 = take(B * (int64_t)w.maxw);
  for (int l = 0; l < n_layers - 1; ++l) w.h[l] = take(B * (int64_t)Ld[l].out_dim);
  w.total = o;
  return true;
}
}  // namespace

extern "C" int64_t ctrhot_widedeep_train_workspace_bytes(int64_t B, int n, const ctrhot_dense_desc* h_layers, int n_layers) {
  TrWdLayout w;
  if (!tr_wd_layout(B, n, h_layers, n_layers, w)) return CTRHOT_ESHAPE;
  return w.total * (int64_t)sizeof(float);
}

extern "C" int ctrhot_widedeep_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                                              int64_t ids_ld, const float* dense, int64_t dense_ld, int D, const float* y,
                                              int64_t B, float* wide_w0, float* wide_w, const int64_t* col_off, int64_t n_wide,
                                              float reg_w, const ctrhot_dense_desc* h_layers, int n_layers, float lr,
                                              float dropout, uint64_t seed, float* loss_out, void* workspace,
                                              int64_t workspace_bytes, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_widedeep_train_step_f32");
  if (!arena || !row_off || !ids || !y || !wide_w0 || !wide_w || !col_off || !loss_out || !workspace || F <= 0 || k <= 0 || D < 0 ||
      (D > 0 && !dense) || B <= 0 || ids_ld < F || (D > 0 && dense_ld < D) || n_wide < 2 * (int64_t)D) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (!(dropout >= 0.f && dropout < 1.f) || !(lr >= 0.f)) return CTRHOT_EARG;
  const int n = F * k;                                           // the deep side sees the embeddings only
  TrWdLayout w;
  if (!tr_wd_layout(B, n, h_layers, n_layers, w)) return CTRHOT_ESHAPE;
  if (workspace_bytes < w.total * (int64_t)sizeof(float) || !aligned16(workspace)) return CTRHOT_EARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* ws = static_cast<float*>(workspace);
  float *x = ws + w.x, *wide = ws + w.wide, *dnn = ws + w.dnn, *dz = ws + w.dz, *dx = ws + w.dx;
  const int NH = n_layers - 1, M = (int)B;
  // ---- forward
  if (ids_kind == CTRHOT_IDS_I32) {
    tr_gather_kernel<CTRHOT_IDS_I32><<<tr_grid(B * n, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, nullptr, 0, 0, B, x, n, err_flag, nullptr);
    tr_wide_fwd_kernel<CTRHOT_IDS_I32><<<tr_grid(B, 128), 128, 0, st>>>(dense, dense_ld, D, ids, ids_ld, F, col_off, wide_w0, wide_w, B, wide);
  } else {
    tr_gather_kernel<CTRHOT_IDS_F32><<<tr_grid(B * n, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, nullptr, 0, 0, B, x, n, err_flag, nullptr);
    tr_wide_fwd_kernel<CTRHOT_IDS_F32><<<tr_grid(B, 128), 128, 0, st>>>(dense, dense_ld, D, ids, ids_ld, F, col_off, wide_w0, wide_w, B, wide);
  }
  const float* in = x;
  int in_dim = n;
  for (int l = 0; l < NH; ++l) {
    float* h = ws + w.h[l];
    const int N = h_layers[l].out_dim;
    tr_dense_fwd(st, in, in_dim, h_layers[l].weight, h_layers[l].bias, N, h, M, 1, dropout, seed, (uint32_t)l, nullptr);
    in = h; in_dim = N;
  }
  tr_gemm(st, in, in_dim, 1, h_layers[NH].weight, 1, 1, dnn, 1, M, 1, in_dim, 1.f, 0.f);
  tr_loss_kernel<<<1, 256, 0, st>>>(wide, dnn, h_layers[NH].bias, y, B, dz, loss_out, nullptr, 0.5f);
  // ---- backward through the tower (d deep = dz / 2)
  float* g = ws + w.g0;
  float* gn = ws + w.g1;
  {
    float* wo = const_cast<float*>(h_layers[NH].weight);
    tr_gemm(st, dz, 1, 1, wo, 0, 1, gn, in_dim, M, in_dim, 1, 0.5f, 0.f);
    tr_gemm(st, in, 1, in_dim, dz, 1, 1, wo, 1, in_dim, 1, M, -0.5f * lr, 1.f);
    if (h_layers[NH].bias)
      tr_colsum(st, 1, dz, 1, nullptr, B, 1, 0, const_cast<float*>(h_layers[NH].bias), -0.5f * lr, 1.f);
    float* tmp = g; g = gn; gn = tmp;
  }
  for (int l = NH - 1; l >= 0; --l) {
    const int N = h_layers[l].out_dim, Kd = h_layers[l].in_dim;
    const float* h = ws + w.h[l];
    const float* hin = l > 0 ? ws + w.h[l - 1] : x;
    float* W = const_cast<float*>(h_layers[l].weight);
    if (!(l < NH - 1 && tr_m32_ok(M, h_layers[l + 1].out_dim)))   // (else the product above applied this mask in its epilogue)
      tr_relu_drop_bwd_kernel<<<tr_grid(B * N, 256), 256, 0, st>>>(g, h, B * N, 1, dropout, seed, (uint32_t)l, nullptr);
    float* dst = l > 0 ? gn : dx;
    tr_dense_bwd_input(st, g, N, W, dst, Kd, M, l > 0 ? ws + w.h[l - 1] : nullptr, dropout, seed, (uint32_t)(l - 1), nullptr);
    tr_gemm(st, hin, 1, Kd, g, N, 1, W, N, Kd, N, M, -lr, 1.f, const_cast<float*>(h_layers[l].bias), -lr);
    if (l > 0) { float* tmp = g; g = gn; gn = tmp; }
  }
  // ---- wide side: w <- (1 - 2 lr reg) w - lr x_w^T (dz / 2) over [dense | dense | one-hot], w0
  tr_scale_kernel<<<tr_grid(n_wide, 256), 256, 0, st>>>(wide_w, n_wide, 1.f - 2.f * lr * reg_w);
  if (D > 0) {
    tr_colsum(st, (D + 31) / 32, dense, dense_ld, dz, B, D, 0, wide_w, -0.5f * lr, 1.f);
    tr_colsum(st, (D + 31) / 32, dense, dense_ld, dz, B, D, 0, wide_w + D, -0.5f * lr, 1.f);
  }
  tr_colsum(st, 1, dz, 1, nullptr, B, 1, 0, wide_w0, -0.5f * lr, 1.f);
  if (ids_kind == CTRHOT_IDS_I32) {
    tr_wide_onehot_sgd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * F, 256), 256, 0, st>>>(ids, ids_ld, F, col_off, dz, 0.5f * lr, B, wide_w);
    tr_embed_sgd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * n, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, 0, B, dx, n, lr, nullptr);
  } else {
    tr_wide_onehot_sgd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * F, 256), 256, 0, st>>>(ids, ids_ld, F, col_off, dz, 0.5f * lr, B, wide_w);
    tr_embed_sgd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * n, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, 0, B, dx, n, lr, nullptr);
  }
  CTRHOT_LAUNCH_OK("ctrhot_widedeep_train_step_f32");
  return CTRHOT_OK;
}

// ---------------------------------------------------------------- DeepCrossing training step
// h_layers = n_res groups of per_res descriptors: per_res - 1 ReLU Dense layers, then the unit's Dense back to width n (act none)
namespace {
constexpr int TR_MAXRES = 8;
struct TrDxLayout { int64_t xs, z, dz, g, gm, ga, gb, act[TR_MAXRES][TR_MAXL], total; int maxw; };
bool tr_dx_layout(int64_t B, int n, const ctrhot_dense_desc* Ld, int n_res, int per_res, TrDxLayout& w) {
  if (!Ld || n_res < 1 || n_res > TR_MAXRES || per_res < 2 || per_res > TR_MAXL || B < 0 || n <= 0) return false;
  int64_t o = 0;
  auto take = [&](int64_t cnt) { const int64_t at = o; o += (cnt + 3) / 4 * 4; return at; };
  w.maxw = n;
  for (int r = 0; r < n_res; ++r)
    for (int l = 0; l < per_res; ++l) {
      const ctrhot_dense_desc& d = Ld[r * per_res + l];
      if (!d.weight || !d.bias) return false;
      if (d.in_dim != (l == 0 ? n : Ld[r * per_res + l - 1].out_dim)) return false;
      if (l < per_res - 1 ? d.act != CTRHOT_ACT_RELU : (d.act != CTRHOT_ACT_NONE || d.out_dim != n)) return false;
      if (d.out_dim > w.maxw) w.maxw = d.out_dim;
    }
  w.xs = take((int64_t)(n_res + 1) * B * n); w.z = take(B); w.dz = take(B);
  w.g = take(B * n); w.gm = take(B * n); w.ga = take(B * (int64_t)w.maxw); w.gb = take(B * (int64_t)w.maxw);
  for (int r = 0; r < n_res; ++r)
    for (int l = 0; l < per_res; ++l) w.act[r][l] = take(B * (int64_t)Ld[r * per_res + l].out_dim);
  w.total = o;
  return true;
}
}  // namespace

extern "C" int64_t ctrhot_deepcrossing_train_workspace_bytes(int64_t B, int n, const ctrhot_dense_desc* h_layers, int n_res,
                                                             int per_res) {
  TrDxLayout w;
  if (!tr_dx_layout(B, n, h_layers, n_res, per_res, w)) return CTRHOT_ESHAPE;
  return w.total * (int64_t)sizeof(float);
}

extern "C" int ctrhot_deepcrossing_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                                                  int64_t ids_ld, const float* dense, int64_t dense_ld, int D, const float* y,
                                                  int64_t B, const ctrhot_dense_desc* h_layers, int n_res, int per_res,
                                                  float* final_k, float* final_b, float lr, float* loss_out, void* workspace,
                                                  int64_t workspace_bytes, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_deepcrossing_train_step_f32");
  if (!arena || !row_off || !ids || !y || !final_k || !final_b || !loss_out || !workspace || F <= 0 || k <= 0 || D < 0 ||
      (D > 0 && !dense) || B <= 0 || ids_ld < F || (D > 0 && dense_ld < D) || !(lr >= 0.f)) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  const int n = D + F * k;
  TrDxLayout w;
  if (!tr_dx_layout(B, n, h_layers, n_res, per_res, w)) return CTRHOT_ESHAPE;
  if (workspace_bytes < w.total * (int64_t)sizeof(float) || !aligned16(workspace)) return CTRHOT_EARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* ws = static_cast<float*>(workspace);
  float *xs = ws + w.xs, *z = ws + w.z, *dz = ws + w.dz, *g = ws + w.g, *gm = ws + w.gm;
  const int M = (int)B;
  const int64_t Bn = B * n;
  // ---- forward: x_0 = [dense | rows]; unit r: relu chain, Dense back to n, x_{r+1} = relu(x_r + h)
  if (ids_kind == CTRHOT_IDS_I32)
    tr_gather_kernel<CTRHOT_IDS_I32><<<tr_grid(Bn, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, dense, dense_ld, D, B, xs, n, err_flag, nullptr);
  else
    tr_gather_kernel<CTRHOT_IDS_F32><<<tr_grid(Bn, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, dense, dense_ld, D, B, xs, n, err_flag, nullptr);
  for (int r = 0; r < n_res; ++r) {
    const float* in = xs + (int64_t)r * Bn;
    int in_dim = n;
    for (int l = 0; l < per_res; ++l) {
      const ctrhot_dense_desc& d = h_layers[r * per_res + l];
      float* h = ws + w.act[r][l];
      tr_dense_fwd(st, in, in_dim, d.weight, d.bias, d.out_dim, h, M, l < per_res - 1, 0.f, 0, 0, nullptr);
      in = h; in_dim = d.out_dim;
    }
    tr_residual_relu_kernel<<<tr_grid(Bn, 256), 256, 0, st>>>(xs + (int64_t)r * Bn, in, Bn, xs + (int64_t)(r + 1) * Bn);
  }
  const float* xR = xs + (int64_t)n_res * Bn;
  tr_gemm(st, xR, n, 1, final_k, 1, 1, z, 1, M, 1, n, 1.f, 0.f);
  tr_loss_kernel<<<1, 256, 0, st>>>(z, nullptr, final_b, y, B, dz, loss_out, nullptr, 1.0f);
  // ---- backward: g = dL/dx_R = dz fk^T (old fk), then the final Dense
  tr_gemm(st, dz, 1, 1, final_k, 0, 1, g, n, M, n, 1, 1.f, 0.f);
  tr_gemm(st, xR, 1, n, dz, 1, 1, final_k, 1, n, 1, M, -lr, 1.f);
  tr_colsum(st, 1, dz, 1, nullptr, B, 1, 0, final_b, -lr, 1.f);
  for (int r = n_res - 1; r >= 0; --r) {
    tr_mask_copy_kernel<<<tr_grid(Bn, 256), 256, 0, st>>>(g, xs + (int64_t)(r + 1) * Bn, Bn, gm);   // through the unit's relu
    const float* gcur = gm;                                  // gradient at the current layer's pre-activation
    float* ga = ws + w.ga;
    float* gb = ws + w.gb;
    for (int l = per_res - 1; l >= 0; --l) {
      const ctrhot_dense_desc& d = h_layers[r * per_res + l];
      const int N = d.out_dim, Kd = d.in_dim;
      const float* hin = l > 0 ? ws + w.act[r][l - 1] : xs + (int64_t)r * Bn;
      float* W = const_cast<float*>(d.weight);
      if (l < per_res - 1) {                                 // relu of this hidden layer (gcur is writable here: ga / gb)
        tr_relu_drop_bwd_kernel<<<tr_grid(B * N, 256), 256, 0, st>>>(const_cast<float*>(gcur), ws + w.act[r][l], B * N, 1, 0.f, 0, 0, nullptr);
      }
      float* dst = l > 0 ? (gcur == ga ? gb : ga) : g;       // the unit's input gradient lands in g, then the skip path is added
      tr_gemm(st, gcur, N, 1, W, 1, N, dst, Kd, M, Kd, N, 1.f, 0.f);
      tr_gemm(st, hin, 1, Kd, gcur, N, 1, W, N, Kd, N, M, -lr, 1.f, const_cast<float*>(d.bias), -lr);
      gcur = dst;
    }
    tr_add_kernel<<<tr_grid(Bn, 256), 256, 0, st>>>(g, gm, Bn);
  }
  if (ids_kind == CTRHOT_IDS_I32)
    tr_embed_sgd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * F * k, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, D, B, g, n, lr, nullptr);
  else
    tr_embed_sgd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * F * k, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, D, B, g, n, lr, nullptr);
  CTRHOT_LAUNCH_OK("ctrhot_deepcrossing_train_step_f32");
  return CTRHOT_OK;
}

// ---------------------------------------------------------------- NFM training step (SURVEY 8(f) ranks 1 + 3)
// h_layers: the DNNLayer's ReLU hidden layers (Dropout after each), then its Dense(output_dim) (act none); final_k [output_dim, 1]
namespace {
struct TrNfmLayout { int64_t x, S, xhat, y, rstd, d, z, dz, gd, gy, gx, dE, g0, g1, h[TR_MAXL], total; int maxw; };
bool tr_nfm_layout(int64_t B, int n, int fk, int k, const ctrhot_dense_desc* Ld, int n_layers, TrNfmLayout& w) {
  if (!Ld || n_layers < 2 || n_layers > TR_MAXL || B < 0 || n <= 0) return false;
  if (Ld[0].in_dim != n) return false;
  int64_t o = 0;
  auto take = [&](int64_t cnt) { const int64_t at = o; o += (cnt + 3) / 4 * 4; return at; };
  w.maxw = n;
  for (int l = 0; l < n_layers; ++l) {
    if (l > 0 && Ld[l].in_dim != Ld[l - 1].out_dim) return false;
    if (Ld[l].act != (l < n_layers - 1 ? CTRHOT_ACT_RELU : CTRHOT_ACT_NONE)) return false;
    if (!Ld[l].weight || !Ld[l].bias) return false;
    if (Ld[l].out_dim > w.maxw) w.maxw = Ld[l].out_dim;
  }
  const int od = Ld[n_layers - 1].out_dim;
  w.x = take(B * n); w.S = take(B * k); w.xhat = take(B * n); w.y = take(B * n); w.rstd = take(n); w.d = take(B * od);
  w.z = take(B); w.dz = take(B); w.gd = take(B * od); w.gy = take(B * n); w.gx = take(B * n); w.dE = take(B * (int64_t)fk);
  w.g0 = take(B * (int64_t)w.maxw); w.g1 = take(B * (int64_t)w.maxw);
  for (int l = 0; l < n_layers - 1; ++l) w.h[l] = take(B * (int64_t)Ld[l].out_dim);
  w.total = o;
  return true;
}
}  // namespace

extern "C" int64_t ctrhot_nfm_train_workspace_bytes(int64_t B, int D, int F, int k, const ctrhot_dense_desc* h_layers, int n_layers) {
  TrNfmLayout w;
  if (!tr_nfm_layout(B, D + k, F * k, k, h_layers, n_layers, w)) return CTRHOT_ESHAPE;
  return w.total * (int64_t)sizeof(float);
}

extern "C" int ctrhot_nfm_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                                         int64_t ids_ld, const float* dense, int64_t dense_ld, int D, const float* y, int64_t B,
                                         float* bn_gamma, float* bn_beta, float* bn_mean, float* bn_var, float momentum, float eps,
                                         const ctrhot_dense_desc* h_layers, int n_layers, float* final_k, float* final_b,
                                         float lr, float dropout, uint64_t seed, float* loss_out, void* workspace,
                                         int64_t workspace_bytes, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_nfm_train_step_f32");
  if (!arena || !row_off || !ids || !y || !bn_gamma || !bn_beta || !bn_mean || !bn_var || !final_k || !final_b || !loss_out ||
      !workspace || F <= 0 || k <= 0 || D < 0 || (D > 0 && !dense) || B <= 0 || ids_ld < F || (D > 0 && dense_ld < D))
    return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (!(dropout >= 0.f && dropout < 1.f) || !(lr >= 0.f)) return CTRHOT_EARG;
  const int n = D + k, fk = F * k;
  TrNfmLayout w;
  if (!tr_nfm_layout(B, n, fk, k, h_layers, n_layers, w)) return CTRHOT_ESHAPE;
  if (workspace_bytes < w.total * (int64_t)sizeof(float) || !aligned16(workspace)) return CTRHOT_EARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* ws = static_cast<float*>(workspace);
  float *x = ws + w.x, *S = ws + w.S, *xhat = ws + w.xhat, *yb = ws + w.y, *rstd = ws + w.rstd, *d = ws + w.d, *z = ws + w.z,
        *dz = ws + w.dz, *gd = ws + w.gd, *gy = ws + w.gy, *gx = ws + w.gx, *dE = ws + w.dE;
  const int NH = n_layers - 1, M = (int)B, od = h_layers[NH].out_dim;
  // ---- forward: Bi-Interaction, BatchNormalization on the batch's statistics, tower, Dense(1)
  if (ids_kind == CTRHOT_IDS_I32)
    tr_nfm_bi_fwd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * n, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, dense, dense_ld, D, B, x, S, err_flag);
  else
    tr_nfm_bi_fwd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * n, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, dense, dense_ld, D, B, x, S, err_flag);
  tr_bn_fwd_kernel<<<n, 256, 0, st>>>(x, n, B, bn_gamma, bn_beta, bn_mean, bn_var, momentum, eps, xhat, yb, rstd);
  const float* in = yb;
  int in_dim = n;
  for (int l = 0; l < NH; ++l) {
    float* h = ws + w.h[l];
    const int N = h_layers[l].out_dim;
    tr_dense_fwd(st, in, in_dim, h_layers[l].weight, h_layers[l].bias, N, h, M, 1, dropout, seed, (uint32_t)l, nullptr);
    in = h; in_dim = N;
  }
  tr_dense_fwd(st, in, in_dim, h_layers[NH].weight, h_layers[NH].bias, od, d, M, 0, 0.f, 0, 0, nullptr);
  tr_gemm(st, d, od, 1, final_k, 1, 1, z, 1, M, 1, od, 1.f, 0.f);
  tr_loss_kernel<<<1, 256, 0, st>>>(z, nullptr, final_b, y, B, dz, loss_out, nullptr, 1.0f);
  // ---- backward: final Dense(1), the DNNLayer's output layer, its hidden layers
  tr_gemm(st, dz, 1, 1, final_k, 0, 1, gd, od, M, od, 1, 1.f, 0.f);                         // gd = dz fk^T (old fk)
  tr_gemm(st, d, 1, od, dz, 1, 1, final_k, 1, od, 1, M, -lr, 1.f);
  tr_colsum(st, 1, dz, 1, nullptr, B, 1, 0, final_b, -lr, 1.f);
  float* g = ws + w.g0;
  float* gn = ws + w.g1;
  {
    float* wo = const_cast<float*>(h_layers[NH].weight);
    tr_gemm(st, gd, od, 1, wo, 1, od, g, in_dim, M, in_dim, od, 1.f, 0.f);                  // [B, od] x [od, h]: B(k, n) = wo[n * od + k]
    tr_gemm(st, in, 1, in_dim, gd, od, 1, wo, od, in_dim, od, M, -lr, 1.f, const_cast<float*>(h_layers[NH].bias), -lr);
  }
  for (int l = NH - 1; l >= 0; --l) {
    const int N = h_layers[l].out_dim, Kd = h_layers[l].in_dim;
    const float* h = ws + w.h[l];
    const float* hin = l > 0 ? ws + w.h[l - 1] : yb;
    float* W = const_cast<float*>(h_layers[l].weight);
    if (!(l < NH - 1 && tr_m32_ok(M, h_layers[l + 1].out_dim)))   // (else the product above applied this mask in its epilogue)
      tr_relu_drop_bwd_kernel<<<tr_grid(B * N, 256), 256, 0, st>>>(g, h, B * N, 1, dropout, seed, (uint32_t)l, nullptr);
    float* dst = l > 0 ? gn : gy;
    tr_dense_bwd_input(st, g, N, W, dst, Kd, M, l > 0 ? ws + w.h[l - 1] : nullptr, dropout, seed, (uint32_t)(l - 1), nullptr);
    tr_gemm(st, hin, 1, Kd, g, N, 1, W, N, Kd, N, M, -lr, 1.f, const_cast<float*>(h_layers[l].bias), -lr);
    if (l > 0) { float* tmp = g; g = gn; gn = tmp; }
  }
  // ---- BatchNormalization (gamma, beta move), Bi-Interaction, embedding rows
  tr_bn_bwd_kernel<<<n, 256, 0, st>>>(gy, xhat, n, B, bn_gamma, bn_beta, rstd, lr, gx);
  if (ids_kind == CTRHOT_IDS_I32) {
    tr_nfm_bi_bwd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, D, B, gx, S, dE);
    tr_embed_sgd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, 0, B, dE, fk, lr, nullptr);
  } else {
    tr_nfm_bi_bwd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, D, B, gx, S, dE);
    tr_embed_sgd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, 0, B, dE, fk, lr, nullptr);
  }
  CTRHOT_LAUNCH_OK("ctrhot_nfm_train_step_f32");
  return CTRHOT_OK;
}

// ---------------------------------------------------------------- AFM training step (SURVEY 8(f) ranks 1 + 3)
extern "C" int64_t ctrhot_afm_train_workspace_bytes(int64_t B, int F, int k) {
  if (B < 0 || F < 2 || k < 1) return CTRHOT_ESHAPE;
  auto r4 = [](int64_t c) { return (c + 3) / 4 * 4; };
  return (4 * r4(B * k) + 2 * r4(B) + r4(B * (int64_t)F * k)) * (int64_t)sizeof(float);
}

extern "C" int ctrhot_afm_train_step_f32(float* arena, const int64_t* row_off, int F, int k, const void* ids, int ids_kind,
                                         int64_t ids_ld, const float* y, int64_t B, int mode, int n_sigmoid, float* out_k,
                                         float* out_b, float lr, float* loss_out, void* workspace, int64_t workspace_bytes,
                                         int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_afm_train_step_f32");
  if (!arena || !row_off || !ids || !y || !out_k || !out_b || !loss_out || !workspace || B <= 0 || ids_ld < F || !(lr >= 0.f))
    return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (F < 2 || F > 0xFFFF || k < 1 || mode < 0 || mode > 2 || (n_sigmoid != 1 && n_sigmoid != 2)) return CTRHOT_ESHAPE;
  if (workspace_bytes < ctrhot_afm_train_workspace_bytes(B, F, k) || !aligned16(workspace)) return CTRHOT_EARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto r4 = [](int64_t c) { return (c + 3) / 4 * 4; };
  float* ws = static_cast<float*>(workspace);
  float* x = ws; float* S = x + r4(B * k); float* gx = S + r4(B * k);
  int32_t* amax = reinterpret_cast<int32_t*>(gx + r4(B * k));
  float* z = reinterpret_cast<float*>(amax) + r4(B * k); float* dz = z + r4(B); float* dE = dz + r4(B);
  const int M = (int)B, fk = F * k, P = F * (F - 1) / 2;
  const bool i32 = ids_kind == CTRHOT_IDS_I32;
  if (i32) tr_afm_fwd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * k, 128), 128, 0, st>>>(arena, row_off, F, k, ids, ids_ld, B, mode, x, S, amax, err_flag);
  else tr_afm_fwd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * k, 128), 128, 0, st>>>(arena, row_off, F, k, ids, ids_ld, B, mode, x, S, amax, err_flag);
  tr_gemm(st, x, k, 1, out_k, 1, 1, z, 1, M, 1, k, 1.f, 0.f);
  tr_afm_loss_kernel<<<1, 256, 0, st>>>(z, out_b, y, B, n_sigmoid, dz, loss_out);
  // gx = dz w^T (old w; the mean's 1 / P folded in), then w, b
  tr_gemm(st, dz, 1, 1, out_k, 0, 1, gx, k, M, k, 1, mode == 0 ? 1.f / (float)P : 1.f, 0.f);
  tr_gemm(st, x, 1, k, dz, 1, 1, out_k, 1, k, 1, M, -lr, 1.f);
  tr_colsum(st, 1, dz, 1, nullptr, B, 1, 0, out_b, -lr, 1.f);
  if (mode == 2) {
    if (i32) tr_afm_max_bwd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * k, 128), 128, 0, st>>>(arena, row_off, F, k, ids, ids_ld, B, gx, amax, dE);
    else tr_afm_max_bwd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * k, 128), 128, 0, st>>>(arena, row_off, F, k, ids, ids_ld, B, gx, amax, dE);
  } else {                                                       // sum of pairs: d / d e_fc = S_c - e_fc, as NFM's Bi-Interaction
    if (i32) tr_nfm_bi_bwd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, 0, B, gx, S, dE);
    else tr_nfm_bi_bwd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, 0, B, gx, S, dE);
  }
  if (i32) tr_embed_sgd_kernel<CTRHOT_IDS_I32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, 0, B, dE, fk, lr, nullptr);
  else tr_embed_sgd_kernel<CTRHOT_IDS_F32><<<tr_grid(B * fk, 256), 256, 0, st>>>(arena, row_off, F, k, ids, ids_ld, 0, B, dE, fk, lr, nullptr);
  CTRHOT_LAUNCH_OK("ctrhot_afm_train_step_f32");
  return CTRHOT_OK;
}

// ---------------------------------------------------------------- FFM training step (DL/model/ffm.py under compile_fit)
extern "C" int64_t ctrhot_ffm_train_workspace_bytes(int64_t B, int field_num, int k) {
  if (B < 0 || field_num < 1 || k < 1) return CTRHOT_ESHAPE;
  auto r4 = [](int64_t c) { return (c + 3) / 4 * 4; };
  return (r4(B * (int64_t)field_num * k) + r4(B * k) + 2 * r4(B)) * (int64_t)sizeof(float);
}

extern "C" int ctrhot_ffm_train_step_f32(const float* dense, int64_t dense_ld, int D, const void* ids, int ids_kind,
                                         int64_t ids_ld, int F, const int64_t* col_off, const float* y, int64_t B, float* w0,
                                         float* w, float* v, int64_t n_features, int field_num, int k, float reg_w, float reg_v,
                                         float lr, float* loss_out, void* workspace, int64_t workspace_bytes, void* stream) {
  CTRHOT_RANGE("ctrhot_ffm_train_step_f32");
  if (!y || !w0 || !w || !v || !loss_out || !workspace || B <= 0 || D < 0 || F < 0 || n_features < D || !(lr >= 0.f))
    return CTRHOT_EARG;
  if ((D > 0 && (!dense || dense_ld < D)) || (F > 0 && (!ids || !col_off || ids_ld < F))) return CTRHOT_EARG;
  if (ids_kind != CTRHOT_IDS_I32 && ids_kind != CTRHOT_IDS_F32) return CTRHOT_EARG;
  if (field_num < 1 || k < 1 || (int64_t)field_num * k > (1 << 20)) return CTRHOT_ESHAPE;
  if (workspace_bytes < ctrhot_ffm_train_workspace_bytes(B, field_num, k) || !aligned16(workspace)) return CTRHOT_EARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto r4 = [](int64_t c) { return (c + 3) / 4 * 4; };
  const int len = field_num * k;
  float* ff = static_cast<float*>(workspace); float* S = ff + r4(B * len); float* z = S + r4(B * k); float* dz = z + r4(B);
  const bool i32 = ids_kind == CTRHOT_IDS_I32;
  const int g1 = tr_grid(B * 32, 128);
  if (i32) tr_ffm_fwd_kernel<CTRHOT_IDS_I32><<<g1, 128, 0, st>>>(dense, dense_ld, D, ids, ids_ld, F, col_off, B, w0, w, v, len, k, ff, S, z);
  else tr_ffm_fwd_kernel<CTRHOT_IDS_F32><<<g1, 128, 0, st>>>(dense, dense_ld, D, ids, ids_ld, F, col_off, B, w0, w, v, len, k, ff, S, z);
  tr_loss_kernel<<<1, 256, 0, st>>>(z, nullptr, nullptr, y, B, dz, loss_out, nullptr, 1.0f);
  // l2(reg)(w) adds 2 reg w to EVERY entry's gradient (FFMLayer.build, interaction.py:131-138): decay first, then the batch's rows
  if (reg_w != 0.f) tr_scale_kernel<<<tr_grid(n_features, 256), 256, 0, st>>>(w, n_features, 1.f - 2.f * lr * reg_w);
  if (reg_v != 0.f) tr_scale_kernel<<<tr_grid(n_features * len, 256), 256, 0, st>>>(v, n_features * len, 1.f - 2.f * lr * reg_v);
  tr_colsum(st, 1, dz, 1, nullptr, B, 1, 0, w0, -lr, 1.f);
  const int g2 = tr_grid(B * (D + F) * 32, 128);
  if (i32) tr_ffm_sgd_kernel<CTRHOT_IDS_I32><<<g2, 128, 0, st>>>(dense, dense_ld, D, ids, ids_ld, F, col_off, B, dz, lr, len, k, ff, S, w, v);
  else tr_ffm_sgd_kernel<CTRHOT_IDS_F32><<<g2, 128, 0, st>>>(dense, dense_ld, D, ids, ids_ld, F, col_off, B, dz, lr, len, k, ff, S, w, v);
  CTRHOT_LAUNCH_OK("ctrhot_ffm_train_step_f32");
  return CTRHOT_OK;
}

// ---------------------------------------------------------------- DIN training step (DL/model/din.py under compile_fit)
namespace {
struct TrDinLayout {
  int64_t kbuf, qbuf, A, mask, s, pw, emb, xhat, yb, rstd, z, dz, g0, g1, gx, gs, gk, total;
  int64_t apre[8], ax[8], tpre[8], tx[8];
};
bool tr_din_layout(int64_t B, int T, int K, int n, const ctrhot_dense_desc* att, int n_att, const ctrhot_dense_desc* tw, int n_tw,
                   TrDinLayout& w) {
  if (n_att < 1 || n_att > 8 || n_tw < 1 || n_tw > 8 || !att || !tw) return false;
  int64_t o = 0;
  auto take = [&](int64_t cnt) { const int64_t at = o; o += (cnt + 3) / 4 * 4; return at; };
  const int64_t BT = B * T;
  int64_t wide_a = 4 * K, wide_t = n;
  for (int l = 0; l < n_att; ++l) {
    if (att[l].in_dim != (l ? att[l - 1].out_dim : 4 * K) || !att[l].weight || !att[l].bias || !att[l].act_param ||
        att[l].alpha_rows != T) return false;
    if (att[l].out_dim > wide_a) wide_a = att[l].out_dim;
  }
  for (int l = 0; l < n_tw; ++l) {
    if (tw[l].in_dim != (l ? tw[l - 1].out_dim : n) || !tw[l].weight || !tw[l].bias || !tw[l].act_param || tw[l].alpha_rows != 1)
      return false;
    if (tw[l].out_dim > wide_t) wide_t = tw[l].out_dim;
  }
  w.kbuf = take(BT * K); w.qbuf = take(B * K); w.A = take(BT * 4 * K); w.mask = take(BT); w.s = take(BT); w.pw = take(BT);
  w.emb = take(B * n); w.xhat = take(B * n); w.yb = take(B * n); w.rstd = take(n); w.z = take(B); w.dz = take(B);
  const int64_t gw = BT * wide_a > B * wide_t ? BT * wide_a : B * wide_t;
  w.g0 = take(gw); w.g1 = take(gw); w.gx = take(B * n); w.gs = take(BT); w.gk = take(BT * K);
  for (int l = 0; l < n_att; ++l) { w.apre[l] = take(BT * att[l].out_dim); w.ax[l] = take(BT * att[l].out_dim); }
  for (int l = 0; l < n_tw; ++l) { w.tpre[l] = take(B * tw[l].out_dim); w.tx[l] = take(B * tw[l].out_dim); }
  w.total = o;
  return true;
}
}  // namespace

extern "C" int64_t ctrhot_din_train_workspace_bytes(int64_t B, int T, int K, int n, const ctrhot_dense_desc* att_layers, int n_att,
                                                    const ctrhot_dense_desc* tower_layers, int n_tower) {
  TrDinLayout w;
  if (B < 0 || T < 1 || K < 1 || n < 2 * K) return CTRHOT_ESHAPE;
  if (!tr_din_layout(B, T, K, n, att_layers, n_att, tower_layers, n_tower, w)) return CTRHOT_ESHAPE;
  return w.total * (int64_t)sizeof(float);
}

extern "C" int ctrhot_din_train_step_f32(float* seq_table, int64_t seq_rows, int K, const int32_t* hist, int64_t hist_ld, int T,
                                         const int32_t* cand, float* const* other_tables, const int64_t* other_rows,
                                         const int32_t* other_dims, int n_other, const int32_t* other_ids, int64_t other_ids_ld,
                                         const float* dense, int64_t dense_ld, int D, const float* y, int64_t B,
                                         const ctrhot_dense_desc* att_layers, int n_att, float* att_out_k, float* att_out_b,
                                         float* bn_gamma, float* bn_beta, float* bn_mean, float* bn_var, float momentum, float eps,
                                         const ctrhot_dense_desc* tower_layers, int n_tower, float* out_k, float* out_b, float lr,
                                         float* loss_out, void* workspace, int64_t workspace_bytes, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_din_train_step_f32");
  if (!seq_table || !hist || !cand || !y || !att_out_k || !att_out_b || !bn_gamma || !bn_beta || !bn_mean || !bn_var || !out_k ||
      !out_b || !loss_out || !workspace || B <= 0 || T < 1 || K < 1 || hist_ld < T || seq_rows < 1 || n_other < 0 || D < 0 ||
      !(lr >= 0.f))
    return CTRHOT_EARG;
  if ((n_other > 0 && (!other_tables || !other_rows || !other_dims || !other_ids || other_ids_ld < n_other)) ||
      (D > 0 && (!dense || dense_ld < D)))
    return CTRHOT_EARG;
  int n = 2 * K + D;
  for (int i = 0; i < n_other; ++i) {
    if (!other_tables[i] || other_rows[i] < 1 || other_dims[i] < 1) return CTRHOT_EARG;
    n += other_dims[i];
  }
  TrDinLayout w;
  if (B * (int64_t)T > INT32_MAX || !tr_din_layout(B, T, K, n, att_layers, n_att, tower_layers, n_tower, w)) return CTRHOT_ESHAPE;
  if (workspace_bytes < w.total * (int64_t)sizeof(float) || !aligned16(workspace)) return CTRHOT_EARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* ws = static_cast<float*>(workspace);
  float *kbuf = ws + w.kbuf, *qbuf = ws + w.qbuf, *A = ws + w.A, *mask = ws + w.mask, *s = ws + w.s, *pw = ws + w.pw,
        *emb = ws + w.emb, *xhat = ws + w.xhat, *yb = ws + w.yb, *rstd = ws + w.rstd, *z = ws + w.z, *dz = ws + w.dz,
        *gx = ws + w.gx, *gs = ws + w.gs, *gk = ws + w.gk;
  const int M = (int)B, BT = (int)(B * T);
  // ---- forward: rows, the attention unit over every (sample, position), masked softmax pooling
  tr_din_gather_kernel<<<tr_grid(B * T * K, 256), 256, 0, st>>>(seq_table, seq_rows, K, hist, hist_ld, T, cand, B, kbuf, qbuf, A, mask,
                                                               emb, n, err_flag);
  const float* in = A;
  int in_dim = 4 * K;
  for (int l = 0; l < n_att; ++l) {
    const int N = att_layers[l].out_dim;
    float* pre = ws + w.apre[l];
    float* x = ws + w.ax[l];
    tr_gemm(st, in, in_dim, 1, att_layers[l].weight, N, 1, pre, N, BT, N, in_dim, 1.f, 0.f);
    tr_bias_prelu_kernel<<<tr_grid((int64_t)BT * N, 256), 256, 0, st>>>(pre, att_layers[l].bias, att_layers[l].act_param, BT, N, T, x);
    in = x; in_dim = N;
  }
  const int ha = in_dim;
  tr_gemm(st, in, ha, 1, att_out_k, 1, 1, s, 1, BT, 1, ha, 1.f, 0.f);
  tr_din_softmax_att_kernel<<<tr_grid(B * 32, 128), 128, 0, st>>>(s, att_out_b, mask, kbuf, T, K, B, pw, emb, n);
  int col = 2 * K;
  for (int i = 0; i < n_other; ++i) {
    tr_din_other_gather_kernel<<<tr_grid(B * other_dims[i], 256), 256, 0, st>>>(other_tables[i], other_rows[i], other_dims[i],
                                                                               other_ids + i, other_ids_ld, B, emb + col, n, err_flag);
    col += other_dims[i];
  }
  if (D > 0) tr_copy_cols_kernel<<<tr_grid(B * D, 256), 256, 0, st>>>(dense, dense_ld, D, B, emb + col, n);
  // ---- BatchNormalization on the batch's statistics, PReLU tower, Dense(1), loss
  tr_bn_fwd_kernel<<<n, 256, 0, st>>>(emb, n, B, bn_gamma, bn_beta, bn_mean, bn_var, momentum, eps, xhat, yb, rstd);
  in = yb; in_dim = n;
  for (int l = 0; l < n_tower; ++l) {
    const int N = tower_layers[l].out_dim;
    float* pre = ws + w.tpre[l];
    float* x = ws + w.tx[l];
    tr_gemm(st, in, in_dim, 1, tower_layers[l].weight, N, 1, pre, N, M, N, in_dim, 1.f, 0.f);
    tr_bias_prelu_kernel<<<tr_grid(B * N, 256), 256, 0, st>>>(pre, tower_layers[l].bias, tower_layers[l].act_param, B, N, 1, x);
    in = x; in_dim = N;
  }
  const int ht = in_dim;
  tr_gemm(st, in, ht, 1, out_k, 1, 1, z, 1, M, 1, ht, 1.f, 0.f);
  tr_loss_kernel<<<1, 256, 0, st>>>(z, nullptr, out_b, y, B, dz, loss_out, nullptr, 1.0f);
  // ---- backward: Dense(1), tower
  float* g = ws + w.g0;
  float* gn = ws + w.g1;
  tr_gemm(st, dz, 1, 1, out_k, 0, 1, g, ht, M, ht, 1, 1.f, 0.f);                            // g = dz out_k^T (old out_k)
  tr_gemm(st, in, 1, ht, dz, 1, 1, out_k, 1, ht, 1, M, -lr, 1.f);
  tr_colsum(st, 1, dz, 1, nullptr, B, 1, 0, out_b, -lr, 1.f);
  for (int l = n_tower - 1; l >= 0; --l) {
    const int N = tower_layers[l].out_dim, Kd = tower_layers[l].in_dim;
    const float* hin = l > 0 ? ws + w.tx[l - 1] : yb;
    float* W = const_cast<float*>(tower_layers[l].weight);
    tr_prelu_bwd_kernel<<<(N + 127) / 128, 128, 0, st>>>(g, ws + w.tpre[l], const_cast<float*>(tower_layers[l].act_param), B, N, 1, lr);
    tr_gemm(st, g, N, 1, W, 1, N, gn, Kd, M, Kd, N, 1.f, 0.f);                              // through the old W
    tr_gemm(st, hin, 1, Kd, g, N, 1, W, N, Kd, N, M, -lr, 1.f, const_cast<float*>(tower_layers[l].bias), -lr);
    float* tmp = g; g = gn; gn = tmp;
  }
  tr_bn_bwd_kernel<<<n, 256, 0, st>>>(g, xhat, n, B, bn_gamma, bn_beta, rstd, lr, gx);
  // ---- attention pooling, the unit's out layer and its Dense + PReLU layers
  tr_din_att_bwd_kernel<<<tr_grid(B * 32, 128), 128, 0, st>>>(gx, n, kbuf, pw, mask, T, K, B, gs, gk);
  const float* xl = ws + w.ax[n_att - 1];
  tr_gemm(st, gs, 1, 1, att_out_k, 0, 1, g, ha, BT, ha, 1, 1.f, 0.f);                       // g = gs w3^T (old w3)
  tr_gemm(st, xl, 1, ha, gs, 1, 1, att_out_k, 1, ha, 1, BT, -lr, 1.f);
  tr_colsum(st, 1, gs, 1, nullptr, BT, 1, 0, att_out_b, -lr, 1.f);
  for (int l = n_att - 1; l >= 0; --l) {
    const int N = att_layers[l].out_dim, Kd = att_layers[l].in_dim;
    const float* hin = l > 0 ? ws + w.ax[l - 1] : A;
    float* W = const_cast<float*>(att_layers[l].weight);
    tr_prelu_bwd_kernel<<<(T * N + 127) / 128, 128, 0, st>>>(g, ws + w.apre[l], const_cast<float*>(att_layers[l].act_param), BT, N, T, lr);
    tr_gemm(st, g, N, 1, W, 1, N, gn, Kd, BT, Kd, N, 1.f, 0.f);
    tr_gemm(st, hin, 1, Kd, g, N, 1, W, N, Kd, N, BT, -lr, 1.f, const_cast<float*>(att_layers[l].bias), -lr);
    float* tmp = g; g = gn; gn = tmp;
  }
  // ---- the concat [q, k, q - k, q * k], then every table (all reads of the tables happened in the forward pass)
  tr_din_concat_bwd_sgd_kernel<<<tr_grid(B * K, 128), 128, 0, st>>>(g, gk, gx, n, kbuf, qbuf, hist, hist_ld, cand, T, K, B, seq_rows, lr,
                                                                   seq_table);
  col = 2 * K;
  for (int i = 0; i < n_other; ++i) {
    tr_din_other_sgd_kernel<<<tr_grid(B * other_dims[i], 256), 256, 0, st>>>(other_tables[i], other_rows[i], other_dims[i], other_ids + i,
                                                                            other_ids_ld, B, gx + col, n, lr);
    col += other_dims[i];
  }
  CTRHOT_LAUNCH_OK("ctrhot_din_train_step_f32");
  return CTRHOT_OK;
}
