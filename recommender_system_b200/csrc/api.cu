// libctrhot: version / status / error plumbing and the config-5 shard helpers (K8).
#include <cstdio>
#include <cstring>
#include "common.cuh"

namespace ctrhot {

static thread_local char g_err[512] = "";

void set_cuda_error(cudaError_t e, const char* where) {
  snprintf(g_err, sizeof(g_err), "%s: %s (%s)", where, cudaGetErrorString(e), cudaGetErrorName(e));
}

// same thread-local text, for a refusal that is not a CUDA error (read back through ctrhot_last_cuda_error())
void set_error_text(const char* text) { snprintf(g_err, sizeof(g_err), "%s", text); }

// ---------------------------------------------------------------- K8 shard route / pack / unpack
// Row-sharded tables: owner(id) = id mod G, local row = id div G  (SURVEY 8(e)).
// route: counting sort of the B*F (sample, field) slots by owner, stable inside an owner.
__global__ void route_count_kernel(const int32_t* __restrict__ ids, int64_t n, int G,
                                   int32_t* __restrict__ counts) {
  extern __shared__ int32_t sc[];
  for (int g = threadIdx.x; g < G; g += blockDim.x) sc[g] = 0;
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t id = ids[i];
    const int o = (int)(((id % G) + G) % G);
    atomicAdd(&sc[o], 1);
  }
  __syncthreads();
  for (int g = threadIdx.x; g < G; g += blockDim.x) if (sc[g]) atomicAdd(&counts[g], sc[g]);
}

__global__ void route_scan_kernel(const int32_t* __restrict__ counts, int G, int32_t* __restrict__ cursor) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    int32_t acc = 0;
    for (int g = 0; g < G; ++g) { cursor[g] = acc; acc += counts[g]; }
  }
}

// Positions are reserved hierarchically: lanes with the same owner elect a leader (__match_any_sync) that bumps a
// shared-memory counter once for the group; the block then reserves one global range per owner with a single atomicAdd
// each.  (One global atomicAdd per slot serialised on G addresses: 4.5 ms for 6.8 M slots at G = 1, profiles/r01d.)
__global__ void __launch_bounds__(256)
route_scatter_kernel(const int32_t* __restrict__ ids, int64_t n, int F, int G,
                     int32_t* __restrict__ cursor, int32_t* __restrict__ local_row,
                     int32_t* __restrict__ field, int32_t* __restrict__ slot) {
  extern __shared__ int32_t sm[];                            // [G] counts of this block's chunk, then [G] global bases
  int32_t* cnt = sm;
  int32_t* gbase = sm + G;
  const int lane = threadIdx.x & 31;
  for (int64_t c0 = (int64_t)blockIdx.x * blockDim.x; c0 < n; c0 += (int64_t)gridDim.x * blockDim.x) {
    for (int g = threadIdx.x; g < G; g += blockDim.x) cnt[g] = 0;
    __syncthreads();
    const int64_t i = c0 + threadIdx.x;
    const bool live = i < n;
    const int32_t id = live ? ids[i] : 0;
    const int o = live ? (int)(((id % G) + G) % G) : -1 - lane;      // dead lanes match nobody
    const unsigned peers = __match_any_sync(0xffffffffu, o);
    const int leader = __ffs(peers) - 1;
    const int rank_in = __popc(peers & ((1u << lane) - 1u));
    int32_t base = 0;
    if (live && lane == leader) base = atomicAdd(&cnt[o], __popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    __syncthreads();
    for (int g = threadIdx.x; g < G; g += blockDim.x) gbase[g] = cnt[g] ? atomicAdd(&cursor[g], cnt[g]) : 0;
    __syncthreads();
    if (live) {
      const int32_t pos = gbase[o] + base + rank_in;
      local_row[pos] = (id >= 0) ? id / G : -1;
      field[pos] = (int32_t)(i % F);
      slot[pos] = (int32_t)i;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256)
shard_gather_pack_kernel(const float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k,
                         const int32_t* __restrict__ local_row, const int32_t* __restrict__ field, int64_t n,
                         float* __restrict__ rows, int* __restrict__ err_flag) {
  const int k4 = k >> 2;
  const int64_t total = n * k4;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = t / k4;
    const int c = (int)(t - r * k4);
    const int f = field[r];
    const int64_t lr = local_row[r];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (f >= 0 && f < F && lr >= 0 && lr < row_off[f + 1] - row_off[f]) v = ldg_nc_f4(arena + (row_off[f] + lr) * k + c * 4);
    else if (err_flag) *err_flag = 1;
    stg_na_f4(rows + r * k + c * 4, v);
  }
}

__global__ void __launch_bounds__(256)
shard_unpack_kernel(const float* __restrict__ rows, const int32_t* __restrict__ slot, int64_t n, int F, int k,
                    int D, float* __restrict__ x, int64_t x_ld) {
  const int k4 = k >> 2;
  const int64_t total = n * k4;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = t / k4;
    const int c = (int)(t - r * k4);
    const int32_t s = slot[r];
    const int64_t b = s / F;
    const int f = s - (int)(b * F);
    const float4 v = ldg_nc_f4(rows + r * k + c * 4);
    float* dst = x + b * x_ld + D + (int64_t)f * k + c * 4;
    // (the aligned store is inline asm on purpose: a float4 cast here lets the compiler assume alignment and turn the
    //  scalar branch into a 128-bit store too -- found by tests/test_gpu_sharded.py with D = 13)
    if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) stg_na_f4(dst, v);
    else { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w; }
  }
}

// Synthetic tables for config 5: value(f, global row, col) = U(-0.05, 0.05) from a counter-based hash, so any
// rank can regenerate any row without a 133-GB host copy (oracle/synth.py::hash_table_value is the same function in NumPy).
__host__ __device__ inline float shard_hash_value(uint64_t seed, int f, int64_t grow, int col) {
  uint64_t z = seed + 0x9E3779B97F4A7C15ull * ((uint64_t)f + 1u) + 0xBF58476D1CE4E5B9ull * (uint64_t)grow +
               0x94D049BB133111EBull * ((uint64_t)col + 1u);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  const float u = (float)(uint32_t)(z >> 40) * (1.0f / 16777216.0f);      // 24 bits -> [0, 1)
  return (u - 0.5f) * 0.1f;
}

__global__ void __launch_bounds__(256)
shard_init_hash_kernel(float* __restrict__ arena, const int64_t* __restrict__ row_off, int F, int k, int G, int rank,
                       uint64_t seed) {
  const int64_t total = row_off[F] * k;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t ar = t / k;
    const int c = (int)(t - ar * k);
    int f = 0;
    while (f + 1 < F && row_off[f + 1] <= ar) ++f;
    const int64_t lr = ar - row_off[f];
    arena[t] = shard_hash_value(seed, f, lr * G + rank, c);
  }
}

}  // namespace ctrhot

using namespace ctrhot;

extern "C" int ctrhot_shard_init_hash_f32(float* arena, const int64_t* row_off, int64_t total_rows, int F, int k,
                                          int G, int rank, uint64_t seed, void* stream) {
  CTRHOT_RANGE("ctrhot_shard_init_hash_f32");
  if (!arena || !row_off || F <= 0 || k <= 0 || G <= 0 || rank < 0 || rank >= G || total_rows < 0) return CTRHOT_EARG;
  if (total_rows == 0) return CTRHOT_OK;
  int64_t g = (total_rows * k + 255) / 256;
  if (g > kNumSMs * 32) g = kNumSMs * 32;
  shard_init_hash_kernel<<<(int)g, 256, 0, static_cast<cudaStream_t>(stream)>>>(arena, row_off, F, k, G, rank, seed);
  CTRHOT_LAUNCH_OK("ctrhot_shard_init_hash_f32");
  return CTRHOT_OK;
}

extern "C" int ctrhot_version(void) { return CTRHOT_VERSION; }

extern "C" const char* ctrhot_status_str(int status) {
  switch (status) {
    case CTRHOT_OK: return "ok";
    case CTRHOT_EARG: return "bad argument";
    case CTRHOT_ESHAPE: return "unsupported shape";
    case CTRHOT_ECUDA: return "CUDA error";
    case CTRHOT_EID: return "embedding id out of range";
    default: return "unknown status";
  }
}

extern "C" const char* ctrhot_last_cuda_error(void) { return g_err; }

extern "C" int ctrhot_check_flag(int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_check_flag");
  if (!err_flag) return CTRHOT_EARG;
  int h = 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CTRHOT_CUDA_OK(cudaMemcpyAsync(&h, err_flag, sizeof(int), cudaMemcpyDeviceToHost, st), "check_flag copy");
  CTRHOT_CUDA_OK(cudaStreamSynchronize(st), "check_flag sync");
  if (h != 0) {
    CTRHOT_CUDA_OK(cudaMemsetAsync(err_flag, 0, sizeof(int), st), "check_flag clear");
    return CTRHOT_EID;
  }
  return CTRHOT_OK;
}

extern "C" int64_t ctrhot_shard_route_workspace_bytes(int64_t B, int F, int G) {
  if (B < 0 || F <= 0 || G <= 0) return CTRHOT_EARG;
  return (int64_t)G * sizeof(int32_t) + 256;
}

extern "C" int ctrhot_shard_route_ids(const int32_t* ids, int64_t B, int F, int G,
                                      int32_t* send_local_row, int32_t* send_field, int32_t* send_slot,
                                      int32_t* counts, void* workspace, int64_t workspace_bytes, void* stream) {
  CTRHOT_RANGE("ctrhot_shard_route_ids");
  if (!ids || !send_local_row || !send_field || !send_slot || !counts || !workspace) return CTRHOT_EARG;
  if (B < 0 || F <= 0 || G <= 0 || G > 1024 || B * F > 2147483647LL) return CTRHOT_EARG;
  if (workspace_bytes < ctrhot_shard_route_workspace_bytes(B, F, G)) return CTRHOT_EARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t n = B * F;
  int32_t* cursor = static_cast<int32_t*>(workspace);
  CTRHOT_CUDA_OK(cudaMemsetAsync(counts, 0, G * sizeof(int32_t), st), "route memset");
  if (n > 0) {
    int64_t g = (n + 255) / 256;
    if (g > kNumSMs * 8) g = kNumSMs * 8;
    route_count_kernel<<<(int)g, 256, G * sizeof(int32_t), st>>>(ids, n, G, counts);
    CTRHOT_LAUNCH_OK("route_count");
  }
  route_scan_kernel<<<1, 32, 0, st>>>(counts, G, cursor);
  CTRHOT_LAUNCH_OK("route_scan");
  if (n > 0) {
    int64_t g = (n + 255) / 256;
    if (g > kNumSMs * 8) g = kNumSMs * 8;
    route_scatter_kernel<<<(int)g, 256, 2 * G * sizeof(int32_t), st>>>(ids, n, F, G, cursor, send_local_row, send_field, send_slot);
    CTRHOT_LAUNCH_OK("route_scatter");
  }
  return CTRHOT_OK;
}

extern "C" int ctrhot_shard_gather_pack(const float* arena, const int64_t* row_off, int F, int k,
                                        const int32_t* local_row, const int32_t* field, int64_t n,
                                        float* rows, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_shard_gather_pack");
  if (!arena || !row_off || !local_row || !field || !rows || F <= 0 || k <= 0 || n < 0) return CTRHOT_EARG;
  if (k % 4 != 0 || !aligned16(arena) || !aligned16(rows)) return CTRHOT_ESHAPE;
  if (n == 0) return CTRHOT_OK;
  int64_t g = (n * (k / 4) + 255) / 256;
  if (g > kNumSMs * 16) g = kNumSMs * 16;
  shard_gather_pack_kernel<<<(int)g, 256, 0, static_cast<cudaStream_t>(stream)>>>(arena, row_off, F, k, local_row,
                                                                               field, n, rows, err_flag);
  CTRHOT_LAUNCH_OK("ctrhot_shard_gather_pack");
  return CTRHOT_OK;
}

extern "C" int ctrhot_shard_unpack(const float* rows, const int32_t* slot, int64_t n, int F, int k, int D,
                                   float* x, int64_t x_ld, void* stream) {
  CTRHOT_RANGE("ctrhot_shard_unpack");
  if (!rows || !slot || !x || F <= 0 || k <= 0 || n < 0 || D < 0 || x_ld < (int64_t)D + (int64_t)F * k) return CTRHOT_EARG;
  if (k % 4 != 0 || !aligned16(rows)) return CTRHOT_ESHAPE;
  if (n == 0) return CTRHOT_OK;
  int64_t g = (n * (k / 4) + 255) / 256;
  if (g > kNumSMs * 16) g = kNumSMs * 16;
  shard_unpack_kernel<<<(int)g, 256, 0, static_cast<cudaStream_t>(stream)>>>(rows, slot, n, F, k, D, x, x_ld);
  CTRHOT_LAUNCH_OK("ctrhot_shard_unpack");
  return CTRHOT_OK;
}
