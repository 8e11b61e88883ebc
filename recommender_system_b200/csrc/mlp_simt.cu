// R4 exact-fp32 Dense tower on CUDA cores: y = act(x @ W + b), layers chained through a
// ping-pong workspace.  This is the bit-faithful fp32 path (FFMA, fp32 accumulate, the
// reference's precision); the tcgen05/TMEM TF32 path in mlp_tc.cu is the fast one.
#include <cstdio>
#include <mutex>
#include "common.cuh"
#include "epilogue.cuh"
#include "mlp.cuh"

namespace ctrhot {

constexpr int BM = 64, BN = 64, BK = 16;

__global__ void __launch_bounds__(256)
dense_f32_kernel(const float* __restrict__ X, int64_t x_ld, int64_t M, int K, int N,
                 const float* __restrict__ W, const float* __restrict__ bias, Epilogue ep,
                 float* __restrict__ Y, int64_t y_ld) {
  __shared__ float sA[BK][BM + 4];
  __shared__ float sB[BK][BN + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t m0 = (int64_t)blockIdx.y * BM;
  const int n0 = blockIdx.x * BN;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < K; k0 += BK) {
    for (int t = threadIdx.x; t < BM * BK; t += 256) {
      const int m = t / BK, kk = t - m * BK;
      const int64_t gm = m0 + m;
      const int gk = k0 + kk;
      sA[kk][m] = (gm < M && gk < K) ? X[gm * x_ld + gk] : 0.f;
    }
    for (int t = threadIdx.x; t < BK * BN; t += 256) {
      const int kk = t / BN, n = t - kk * BN;
      const int gk = k0 + kk, gn = n0 + n;
      sB[kk][n] = (gk < K && gn < N) ? __ldg(W + (size_t)gk * N + gn) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = sA[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = sB[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gn = n0 + tx * 4 + j;
      if (gn >= N) continue;
      float t = acc[i][j] + (bias ? __ldg(bias + gn) : 0.f);
      Y[gm * y_ld + gn] = ep.apply(t, gm, gn, N);
    }
  }
}

int launch_dense_f32(const float* X, int64_t x_ld, int64_t M, const ctrhot_dense_desc& d, float* Y, int64_t y_ld,
                     cudaStream_t st) {
  dim3 grid((d.out_dim + BN - 1) / BN, (unsigned)((M + BM - 1) / BM));
  dense_f32_kernel<<<grid, 256, 0, st>>>(X, x_ld, M, d.in_dim, d.out_dim, d.weight, d.bias, Epilogue::from(d), Y, y_ld);
  CTRHOT_LAUNCH_OK("dense_f32");
  return CTRHOT_OK;
}

}  // namespace ctrhot

using namespace ctrhot;

extern "C" int64_t ctrhot_mlp_workspace_bytes(int64_t M, const ctrhot_dense_desc* h_layers, int n_layers) {
  if (M < 0 || !h_layers || n_layers <= 0) return CTRHOT_EARG;
  int64_t mx = 0;
  for (int l = 0; l + 1 < n_layers; ++l) {
    int64_t w = (h_layers[l].out_dim + 31) / 32 * 32;
    if (w > mx) mx = w;
  }
  return 2 * M * mx * (int64_t)sizeof(float) + 256;
}

extern "C" int ctrhot_mlp_f32(const float* x, int64_t x_ld, int64_t M, const ctrhot_dense_desc* h_layers,
                              int n_layers, float* out, int64_t out_ld, void* workspace,
                              int64_t workspace_bytes, void* stream) {
  CTRHOT_RANGE("ctrhot_mlp_f32");
  if (!x || !h_layers || n_layers <= 0 || !out || M < 0) return CTRHOT_EARG;
  if (x_ld < h_layers[0].in_dim || out_ld < h_layers[n_layers - 1].out_dim) return CTRHOT_EARG;
  for (int l = 0; l < n_layers; ++l) {
    const ctrhot_dense_desc& d = h_layers[l];
    if (!d.weight || d.in_dim <= 0 || d.out_dim <= 0) return CTRHOT_EARG;
    if (l > 0 && d.in_dim != h_layers[l - 1].out_dim) return CTRHOT_EARG;
    if ((d.act == CTRHOT_ACT_PRELU || d.act == CTRHOT_ACT_DICE) && !d.act_param) return CTRHOT_EARG;
    if (d.act < CTRHOT_ACT_NONE || d.act > CTRHOT_ACT_TANH) return CTRHOT_EARG;
  }
  if (M == 0) return CTRHOT_OK;
  const int64_t need = ctrhot_mlp_workspace_bytes(M, h_layers, n_layers);
  if (n_layers > 1 && (!workspace || workspace_bytes < need)) return CTRHOT_EARG;
  int64_t mx = 0;
  for (int l = 0; l + 1 < n_layers; ++l) {
    int64_t w = (h_layers[l].out_dim + 31) / 32 * 32;
    if (w > mx) mx = w;
  }
  float* buf[2] = {static_cast<float*>(workspace), static_cast<float*>(workspace) + M * mx};
  const float* cur = x;
  int64_t cur_ld = x_ld;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  for (int l = 0; l < n_layers; ++l) {
    const ctrhot_dense_desc& d = h_layers[l];
    const bool last = (l == n_layers - 1);
    float* y = last ? out : buf[l & 1];
    const int64_t y_ld = last ? out_ld : mx;
    const int rc = launch_dense_f32(cur, cur_ld, M, d, y, y_ld, st);
    if (rc != CTRHOT_OK) return rc;
    cur = y;
    cur_ld = y_ld;
  }
  return CTRHOT_OK;
}

// ---------------------------------------------------------------- ctrhot_mlp_f32_static: the same tower, NO allocation anywhere
// The ping-pong buffers of ctrhot_mlp_f32 live in a file-scope device array instead of a caller workspace, for callers that
// cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Every GPU that loads the library
// holds these CTRHOT_MLP_STATIC_SCRATCH_BYTES until exit.
namespace ctrhot {
static __device__ __align__(256) unsigned char g_mlp_static_scratch[CTRHOT_MLP_STATIC_SCRATCH_BYTES];

// In HOST code `g_mlp_static_scratch` is only a stand-in for the symbol, not a device address: the address a kernel may be
// given comes from cudaGetSymbolAddress, asked once per device with that device current (each device has its own copy).
static int mlp_static_scratch_ptr(float** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static float* ptr[kMaxDev] = {nullptr};
  int dev = -1;
  CTRHOT_CUDA_OK(cudaGetDevice(&dev), "ctrhot_mlp_f32_static: cudaGetDevice");
  if (dev < 0 || dev >= kMaxDev) { set_error_text("ctrhot_mlp_f32_static: device ordinal beyond 63"); return CTRHOT_ESHAPE; }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    void* p = nullptr;
    CTRHOT_CUDA_OK(cudaGetSymbolAddress(&p, g_mlp_static_scratch), "ctrhot_mlp_f32_static: cudaGetSymbolAddress");
    if (!p) { set_error_text("ctrhot_mlp_f32_static: cudaGetSymbolAddress returned a null address"); return CTRHOT_ECUDA; }
    ptr[dev] = static_cast<float*>(p);
  }
  *out = ptr[dev];
  return CTRHOT_OK;
}
}  // namespace ctrhot

extern "C" int ctrhot_mlp_f32_static(const float* x, int64_t x_ld, int64_t M, const ctrhot_dense_desc* h_layers,
                                     int n_layers, float* out, int64_t out_ld, void* stream) {
  CTRHOT_RANGE("ctrhot_mlp_f32_static");
  if (!x || !h_layers || n_layers <= 0 || !out || M < 0) return CTRHOT_EARG;
  if (x_ld < h_layers[0].in_dim || out_ld < h_layers[n_layers - 1].out_dim) return CTRHOT_EARG;
  for (int l = 0; l < n_layers; ++l) {
    const ctrhot_dense_desc& d = h_layers[l];
    if (!d.weight || d.in_dim <= 0 || d.out_dim <= 0) return CTRHOT_EARG;
    if (l > 0 && d.in_dim != h_layers[l - 1].out_dim) return CTRHOT_EARG;
    if ((d.act == CTRHOT_ACT_PRELU || d.act == CTRHOT_ACT_DICE) && !d.act_param) return CTRHOT_EARG;
    if (d.act < CTRHOT_ACT_NONE || d.act > CTRHOT_ACT_TANH) return CTRHOT_EARG;
  }
  if (M == 0) return CTRHOT_OK;
  int64_t mx = 0;
  for (int l = 0; l + 1 < n_layers; ++l) {
    int64_t w = (h_layers[l].out_dim + 31) / 32 * 32;
    if (w > mx) mx = w;
  }
  // two buffers of M x mx floats; checked without forming 2 * M * mx (M is caller-controlled)
  const int64_t cap_floats = (int64_t)CTRHOT_MLP_STATIC_SCRATCH_BYTES / (int64_t)sizeof(float);
  if (mx > 0 && M > cap_floats / (2 * mx)) {
    char msg[256];
    snprintf(msg, sizeof(msg),
             "ctrhot_mlp_f32_static: hidden activations need 2 x %lld rows x %lld floats = %lld bytes, the static scratch "
             "holds %lld (use ctrhot_mlp_f32 with a caller workspace, or fewer rows per call)",
             (long long)M, (long long)mx, (long long)(2 * M * mx * (int64_t)sizeof(float)),
             (long long)CTRHOT_MLP_STATIC_SCRATCH_BYTES);
    set_error_text(msg);
    return CTRHOT_ESHAPE;
  }
  float* scratch = nullptr;
  const int rc0 = mlp_static_scratch_ptr(&scratch);
  if (rc0 != CTRHOT_OK) return rc0;
  // Nothing in the scratch has to start at zero: layer l writes columns [0, out_dim) of rows [0, M) of its buffer and layer
  // l + 1 reads exactly those, so whatever an earlier call left behind is overwritten before it is read.
  float* buf[2] = {scratch, scratch + M * mx};
  const float* cur = x;
  int64_t cur_ld = x_ld;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  for (int l = 0; l < n_layers; ++l) {
    const ctrhot_dense_desc& d = h_layers[l];
    const bool last = (l == n_layers - 1);
    float* y = last ? out : buf[l & 1];
    const int64_t y_ld = last ? out_ld : mx;
    const int rc = launch_dense_f32(cur, cur_ld, M, d, y, y_ld, st);
    if (rc != CTRHOT_OK) return rc;
    cur = y;
    cur_ld = y_ld;
  }
  return CTRHOT_OK;
}
