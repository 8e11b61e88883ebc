// K8', config 5 over NVSwitch peer memory (SURVEY 8(e) "stretch"): the requester's gather kernel reads the rows it needs
// straight out of the owner's HBM (P2P loads over NVLink 5), so route + all-to-all(ids) + gather_pack + all-to-all(rows) +
// unpack collapse into ONE kernel with no staging buffer: x[b, D + f*k : ...] = arena_{id mod G}[row_off_{id mod G}[f] + id div G].
// The arenas are cudaMalloc'd by ctrhot_peer_alloc so that a cudaIpcMemHandle names exactly their base address.
#include <cstring>
#include "common.cuh"

namespace ctrhot {

constexpr int kMaxPeers = 16;

struct PeerArenas {
  const float* base[kMaxPeers];
};

// One sub-warp of tpr = k/4 lanes per (sample, field) slot (k = 128 -> a full warp reads one 512-B row with 16-B loads).
// x rows start at an arbitrary float offset (D = 13 in DeepFM), so the destination is usually NOT 16-byte aligned: the
// row is rotated through one shuffle so that every lane but the ends still issues an aligned 128-bit store.
template <int UNROLL>
__global__ void __launch_bounds__(256)
shard_lookup_peer_kernel(const __grid_constant__ PeerArenas arenas, const int64_t* __restrict__ row_off /* [G, F+1] */, int F, int k, int G,
                         const int32_t* __restrict__ ids, int64_t n_slots, int D, const float* __restrict__ dense,
                         int64_t dense_ld, float* __restrict__ x, int64_t x_ld, int tpr_log2, int* __restrict__ err_flag) {
  const int tpr = 1 << tpr_log2;                               // lanes per row
  const int c = threadIdx.x & (tpr - 1);
  const int64_t rows_per_pass = ((int64_t)gridDim.x * blockDim.x) >> tpr_log2;
  const int sub = (threadIdx.x & 31) >> tpr_log2;              // which row of the warp's 32/tpr rows
  // the trip count is decided on the warp's FIRST row so that all 32 lanes stay in the loop for the full-mask shuffles
  int64_t rw = (((int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31))) >> tpr_log2;
  const int k4 = k >> 2;                                       // 16-B pieces per row; k4 = tpr * pieces_per_lane
  const int ppl = k4 >> tpr_log2;                              // pieces per lane (1 unless k > 128)
  for (; rw < n_slots; rw += UNROLL * rows_per_pass) {
    const int64_t r0 = rw + sub;
    float4 v[UNROLL];
    const float* src[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t r = r0 + u * rows_per_pass;
      src[u] = nullptr;
      if (r < n_slots) {
        const int32_t id = ids[r];
        const int f = (int)(r % F);
        const int o = (int)((unsigned)id % (unsigned)G);
        const int64_t lr = (unsigned)id / (unsigned)G;
        const int64_t* ro = row_off + (int64_t)o * (F + 1) + f;
        const int64_t base = ro[0];
        if (id >= 0 && lr < ro[1] - base) src[u] = arenas.base[o] + (base + lr) * k;
        else if (err_flag && c == 0) *err_flag = 1;
      }
    }
    for (int p = 0; p < ppl; ++p) {
#pragma unroll
      for (int u = 0; u < UNROLL; ++u)
        v[u] = src[u] ? ldg_nc_f4(src[u] + (p * tpr + c) * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int64_t r = r0 + u * rows_per_pass;
        const bool live = r < n_slots;
        const int64_t b = live ? r / F : 0;
        const int f = live ? (int)(r - b * F) : 0;
        float* dst = x + b * x_ld + D + (int64_t)f * k + (p * tpr + c) * 4;
        const int m = (int)((reinterpret_cast<uintptr_t>(dst) >> 2) & 3);     // floats past a 16-B boundary (warp-uniform
        // when x_ld % 4 == 0 and k % 4 == 0; otherwise per-row -- the code below only needs it uniform per sub-warp)
        // neighbour's piece: lane c+1 of the same row (the sub-warp is tpr lanes wide)
        float4 nx;
        nx.x = __shfl_down_sync(0xffffffffu, v[u].x, 1, tpr);
        nx.y = __shfl_down_sync(0xffffffffu, v[u].y, 1, tpr);
        nx.z = __shfl_down_sync(0xffffffffu, v[u].z, 1, tpr);
        nx.w = __shfl_down_sync(0xffffffffu, v[u].w, 1, tpr);
        if (!live) continue;
        if (m == 0) { stg_na_f4(dst, v[u]); continue; }
        // aligned chunk starting at dst + (4 - m): my last m floats... spelled out per m to keep everything in registers
        const bool last = (c == tpr - 1);
        float* al = dst + (4 - m);
        if (m == 1) {
          if (c == 0) { dst[0] = v[u].x; dst[1] = v[u].y; dst[2] = v[u].z; }
          if (!last) stg_na_f4(al, make_float4(v[u].w, nx.x, nx.y, nx.z));
          else al[0] = v[u].w;
        } else if (m == 2) {
          if (c == 0) { dst[0] = v[u].x; dst[1] = v[u].y; }
          if (!last) stg_na_f4(al, make_float4(v[u].z, v[u].w, nx.x, nx.y));
          else { al[0] = v[u].z; al[1] = v[u].w; }
        } else {
          if (c == 0) dst[0] = v[u].x;
          if (!last) stg_na_f4(al, make_float4(v[u].y, v[u].z, v[u].w, nx.x));
          else { al[0] = v[u].y; al[1] = v[u].z; al[2] = v[u].w; }
        }
      }
    }
  }
  // dense columns x[b, 0:D] (deepFM.py:26 column order [dense | embeddings]) and the zero pad behind column D + F*k
  if (dense) {
    const int64_t B = n_slots / F;
    const int n = D + F * k;
    const int w = D + (int)(x_ld - n);
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < B * w; t += (int64_t)gridDim.x * blockDim.x) {
      const int64_t b = t / w;
      const int j = (int)(t - b * w);
      if (j < D) x[b * x_ld + j] = dense[b * dense_ld + j];
      else x[b * x_ld + n + (j - D)] = 0.f;
    }
  }
}

}  // namespace ctrhot

using namespace ctrhot;

extern "C" int ctrhot_peer_handle_bytes(void) {
  CTRHOT_RANGE("ctrhot_peer_handle_bytes"); return (int)sizeof(cudaIpcMemHandle_t); }

extern "C" int ctrhot_peer_alloc(int64_t bytes, void** ptr) {
  CTRHOT_RANGE("ctrhot_peer_alloc");
  if (!ptr || bytes < 0) return CTRHOT_EARG;
  *ptr = nullptr;
  CTRHOT_CUDA_OK(cudaMalloc(ptr, (size_t)(bytes > 0 ? bytes : 256)), "ctrhot_peer_alloc");
  return CTRHOT_OK;
}

extern "C" int ctrhot_peer_free(void* ptr) {
  CTRHOT_RANGE("ctrhot_peer_free");
  if (ptr) CTRHOT_CUDA_OK(cudaFree(ptr), "ctrhot_peer_free");
  return CTRHOT_OK;
}

extern "C" int ctrhot_peer_export(void* ptr, void* handle_out) {
  CTRHOT_RANGE("ctrhot_peer_export");
  if (!ptr || !handle_out) return CTRHOT_EARG;
  cudaIpcMemHandle_t h;
  CTRHOT_CUDA_OK(cudaIpcGetMemHandle(&h, ptr), "ctrhot_peer_export");
  memcpy(handle_out, &h, sizeof(h));
  return CTRHOT_OK;
}

extern "C" int ctrhot_peer_open(const void* handle, void** ptr) {
  CTRHOT_RANGE("ctrhot_peer_open");
  if (!handle || !ptr) return CTRHOT_EARG;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  *ptr = nullptr;
  CTRHOT_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess), "ctrhot_peer_open");
  return CTRHOT_OK;
}

extern "C" int ctrhot_peer_enable(int peer_device) {
  CTRHOT_RANGE("ctrhot_peer_enable");
  // one process driving several GPUs: let kernels on the current device dereference peer_device's allocations
  const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { (void)cudaGetLastError(); return CTRHOT_OK; }
  CTRHOT_CUDA_OK(e, "ctrhot_peer_enable");
  return CTRHOT_OK;
}

extern "C" int ctrhot_peer_close(void* ptr) {
  CTRHOT_RANGE("ctrhot_peer_close");
  if (ptr) CTRHOT_CUDA_OK(cudaIpcCloseMemHandle(ptr), "ctrhot_peer_close");
  return CTRHOT_OK;
}

extern "C" int ctrhot_shard_lookup_peer_f32(const void* const* arenas, const int64_t* row_off, int F, int k, int G,
                                            const int32_t* ids, int64_t B, int D, const float* dense, int64_t dense_ld,
                                            float* x, int64_t x_ld, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_shard_lookup_peer_f32");
  if (!arenas || !row_off || !ids || !x || F <= 0 || G <= 0 || G > kMaxPeers || k <= 0 || (k & 3) || D < 0 || B < 0)
    return CTRHOT_EARG;
  if (x_ld < D + (int64_t)F * k || (dense && dense_ld < D)) return CTRHOT_ESHAPE;
  if ((reinterpret_cast<uintptr_t>(x) & 3) != 0) return CTRHOT_EARG;
  if (B == 0) return CTRHOT_OK;
  // lanes per row: the largest power of two <= 32 dividing k/4
  int k4 = k >> 2, tl = 0;
  while (tl < 5 && (k4 % (2 << tl)) == 0) ++tl;
  PeerArenas pa;
  for (int g = 0; g < kMaxPeers; ++g) pa.base[g] = g < G ? static_cast<const float*>(arenas[g]) : nullptr;
  for (int g = 0; g < G; ++g) if (!pa.base[g]) return CTRHOT_EARG;
  const int64_t n_slots = B * F;
  const int64_t thr = n_slots << tl;
  int64_t grid = (thr + 255) / 256;
  const int64_t cap = (int64_t)kNumSMs * 8 * 4;               // 8 CTAs of 256 threads per SM, 4 waves
  if (grid > cap) grid = cap;
  shard_lookup_peer_kernel<2><<<(int)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      pa, row_off, F, k, G, ids, n_slots, D, dense, dense_ld, x, x_ld, tl, err_flag);
  CTRHOT_LAUNCH_OK("ctrhot_shard_lookup_peer_f32");
  return CTRHOT_OK;
}
