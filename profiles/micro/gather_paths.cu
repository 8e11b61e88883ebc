// Microbenchmark: how fast can ONE SM (and all 148 together) pull random 64-byte rows out of a table that is far larger than
// L2, through each of the paths a gather can use?   One persistent CTA per SM, rows land in a shared-memory ring.
//   mode 0  cp.async (LDGSTS) 16 B per thread, 4 threads per row        -- what ctr_fused_kernel's gather warps do
//   mode 1  cp.async.bulk (TMA engine) 64 B per row, one thread per row
//   mode 2  ld.global.nc.v4 -> registers -> st.shared (8 loads in flight per thread)
//   mode 3  half the rows by mode 0, half by mode 1, concurrently
// ids come from a per-thread LCG (no id loads), nothing consumes the rows: this is the row-fetch ceiling only.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I recommender_system_b200/csrc -o profiles/micro/gather_paths profiles/micro/gather_paths.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "tc05.cuh"
using namespace ctrhot::tc;

constexpr int ROWS = 256;            // rows of 64 B per stage = 16 KB
constexpr int STAGE = ROWS * 64;

__device__ __forceinline__ uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s; }

// RB-byte rows, mode 0 (cp.async, RB/16 threads per row) and mode 1 (one bulk copy per row) only
template <int MODE>
__global__ void __launch_bounds__(1024, 1) fetch_rb_kernel(const float* __restrict__ table, uint32_t n_rows, int iters, int stages,
                                                           int n_thr, int RB, long long* cyc, float* sink) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  const uint32_t bars = base + (uint32_t)stages * STAGE;
  auto full = [&](int s) { return bars + 8u * s; };
  const int tid = threadIdx.x;
  const int rows = STAGE / RB, ppr = RB / 16;
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(full(s), MODE == 0 ? n_thr : 1);
    mbar_fence_init();
  }
  __syncthreads();
  const uint32_t seed = (blockIdx.x * 1024u + 7u) * 2654435761u + 12345u;
  const long long t0 = clock64();
  if (tid < n_thr) {
    int s = 0; uint32_t ph = 0;
    for (int it = 0; it < iters; ++it) {
      if (it >= stages) mbar_wait(full(s), ph ^ 1u);
      const uint32_t dst = base + (uint32_t)s * STAGE;
      if (MODE == 0) {
        for (int piece = tid; piece < rows * ppr; piece += n_thr) {
          const int r = piece / ppr, c = piece - r * ppr;
          uint32_t sd = (seed + (uint32_t)(it * 977 + r) * 2246822519u);
          const uint32_t id = lcg(sd) % n_rows;
          cp_async16(dst + (uint32_t)piece * 16u, table + (size_t)id * (RB / 4) + c * 4, 16u);
        }
        cp_async_mbar_arrive_noinc(full(s));
      } else {
        if (tid == 0) mbar_arrive_expect_tx(full(s), (uint32_t)STAGE);
        for (int r = tid; r < rows; r += n_thr) {
          uint32_t sd = (seed + (uint32_t)(it * 977 + r) * 2246822519u);
          const uint32_t id = lcg(sd) % n_rows;
          bulk_g2s(dst + (uint32_t)r * RB, table + (size_t)id * (RB / 4), (uint32_t)RB, full(s));
        }
      }
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
    for (int k = 0; k < stages && k < iters; ++k) {
      mbar_wait(full(s), ph ^ 1u);
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
  }
  __syncthreads();
  if (tid == 0) {
    cyc[blockIdx.x] = clock64() - t0;
    sink[blockIdx.x] = *reinterpret_cast<float*>(smem_raw + (base - smem_u32(smem_raw)) + 64);
  }
}

template <int MODE>
static void run_rb(const char* name, const float* table, size_t table_bytes, int RB, int n_thr, int stages, long long* cyc, float* sink) {
  const int iters = 400, grid = 148;
  const uint32_t n_rows = (uint32_t)(table_bytes / RB);
  const size_t smem = 128 + (size_t)stages * STAGE + 8 * stages + 16;
  cudaFuncSetAttribute(fetch_rb_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  fetch_rb_kernel<MODE><<<grid, 1024, smem>>>(table, n_rows, 20, stages, n_thr, RB, cyc, sink);
  cudaEventRecord(e0);
  fetch_rb_kernel<MODE><<<grid, 1024, smem>>>(table, n_rows, iters, stages, n_thr, RB, cyc, sink);
  cudaEventRecord(e1);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s: error %s\n", name, cudaGetErrorString(cudaGetLastError())); exit(1); }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  std::vector<long long> h(grid); cudaMemcpy(h.data(), cyc, grid * 8, cudaMemcpyDeviceToHost);
  double c = 0; for (auto v : h) c += (double)v; c /= grid;
  const double bytes = (double)iters * STAGE;
  printf("%-12s rows of %3d B, table %6.0f MB, thr=%4d stages=%2d : %6.2f B/clk/SM  %7.1f cyc per 16 KB  %6.2f TB/s chip\n", name, RB,
         table_bytes / 1e6, n_thr, stages, bytes / c, c / iters, bytes * grid / (ms * 1e-3) / 1e12);
}

template <int MODE>
__global__ void __launch_bounds__(512, 1) fetch_kernel(const float* __restrict__ table, uint32_t n_rows, int iters, int stages,
                                                       int n_thr, long long* cyc, float* sink) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  const uint32_t bars = base + (uint32_t)stages * STAGE;
  auto full = [&](int s) { return bars + 8u * s; };
  const int tid = threadIdx.x;
  // arrivals per stage: mode 0: every fetch thread (noinc arrive); mode 1: one expect_tx arrive; mode 2: every thread; mode 3: half + 1
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      const uint32_t cnt = MODE == 0 ? n_thr : MODE == 1 ? 1 : MODE == 2 ? n_thr : n_thr / 2 + 1;
      mbar_init(full(s), cnt);
    }
    mbar_fence_init();
  }
  __syncthreads();
  uint32_t seed = (blockIdx.x * 1024u + tid) * 2654435761u + 12345u;
  const long long t0 = clock64();
  if (tid < n_thr) {
    int s = 0; uint32_t ph = 0;
    // a stage is reused only after its previous fill completed: the issuing thread waits for `full` of the stage it is about to
    // overwrite (stages - 1 fills stay in flight)
    for (int it = 0; it < iters; ++it) {
      if (it >= stages) mbar_wait(full(s), ph ^ 1u);
      const uint32_t dst = base + (uint32_t)s * STAGE;
      if (MODE == 0 || (MODE == 3 && tid < n_thr / 2)) {
        const int nt = MODE == 0 ? n_thr : n_thr / 2;
        const int rows_here = MODE == 0 ? ROWS : ROWS / 2;
        // 4 threads per row; each thread: rows_here * 4 / nt pieces
        const int per = rows_here * 4 / nt;
        for (int j = 0; j < per; ++j) {
          const int piece = j * nt + tid;
          const int r = piece >> 2, c = piece & 3;
          uint32_t sd = (seed + (uint32_t)(it * 977 + r) * 2246822519u);
          const uint32_t id = lcg(sd) % n_rows;
          cp_async16(dst + (uint32_t)r * 64u + (uint32_t)c * 16u, table + (size_t)id * 16 + c * 4, 16u);
        }
        cp_async_mbar_arrive_noinc(full(s));
      } else if (MODE == 1 || MODE == 3) {
        const int t1 = MODE == 1 ? tid : tid - n_thr / 2;
        const int nt = MODE == 1 ? n_thr : n_thr / 2;
        const int rows_here = MODE == 1 ? ROWS : ROWS / 2;
        const uint32_t off = MODE == 1 ? 0u : (uint32_t)(ROWS / 2) * 64u;
        if (t1 == 0) mbar_arrive_expect_tx(full(s), (uint32_t)rows_here * 64u);
        for (int r = t1; r < rows_here; r += nt) {
          uint32_t sd = (seed + (uint32_t)(it * 977 + r) * 2246822519u);
          const uint32_t id = lcg(sd) % n_rows;
          bulk_g2s(dst + off + (uint32_t)r * 64u, table + (size_t)id * 16, 64u, full(s));
        }
      } else {
        const int per = ROWS * 4 / n_thr;
        float4 v[8];
        for (int j0 = 0; j0 < per; j0 += 8) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (j0 + j < per) {
              const int piece = (j0 + j) * n_thr + tid;
              const int r = piece >> 2, c = piece & 3;
              uint32_t sd = (seed + (uint32_t)(it * 977 + r) * 2246822519u);
              const uint32_t id = lcg(sd) % n_rows;
              asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                           : "=f"(v[j].x), "=f"(v[j].y), "=f"(v[j].z), "=f"(v[j].w) : "l"(table + (size_t)id * 16 + c * 4));
            }
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (j0 + j < per) {
              const int piece = (j0 + j) * n_thr + tid;
              asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" :: "r"(dst + (uint32_t)piece * 16u), "f"(v[j].x), "f"(v[j].y),
                           "f"(v[j].z), "f"(v[j].w) : "memory");
            }
          }
        }
        mbar_arrive(full(s));
      }
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
    // drain
    for (int k = 0; k < stages && k < iters; ++k) {
      mbar_wait(full(s), ph ^ 1u);
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
  }
  __syncthreads();
  if (tid == 0) {
    cyc[blockIdx.x] = clock64() - t0;
    sink[blockIdx.x] = *reinterpret_cast<float*>(smem_raw + (base - smem_u32(smem_raw)) + 64);
  }
}

template <int MODE>
static void run(const char* name, const float* table, uint32_t n_rows, int n_thr, int stages, int grid, long long* cyc, float* sink) {
  const int iters = 400;
  const size_t smem = 128 + (size_t)stages * STAGE + 8 * stages + 16;
  cudaFuncSetAttribute(fetch_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  fetch_kernel<MODE><<<grid, 512, smem>>>(table, n_rows, 20, stages, n_thr, cyc, sink);
  cudaEventRecord(e0);
  fetch_kernel<MODE><<<grid, 512, smem>>>(table, n_rows, iters, stages, n_thr, cyc, sink);
  cudaEventRecord(e1);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s: error %s\n", name, cudaGetErrorString(cudaGetLastError())); exit(1); }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  std::vector<long long> h(grid); cudaMemcpy(h.data(), cyc, grid * 8, cudaMemcpyDeviceToHost);
  double c = 0; for (auto v : h) c += (double)v; c /= grid;
  const double bytes = (double)iters * STAGE;
  printf("%-34s thr=%3d stages=%2d grid=%3d : %6.2f B/clk/SM  %7.1f cyc per 16 KB  %6.2f TB/s chip\n", name, n_thr, stages, grid,
         bytes / c, c / iters, bytes * grid / (ms * 1e-3) / 1e12);
}

int main(int argc, char** argv) {
  const size_t big = (size_t)3 << 30;
  float* table; cudaMalloc(&table, big); cudaMemset(table, 0, big);
  long long* cyc; cudaMalloc(&cyc, 148 * 8);
  float* sink; cudaMalloc(&sink, 148 * 4);
  printf("-- issue scaling of the bulk path, 64 B rows, 3 GB table\n");
  for (int thr : {256, 512, 1024}) for (int st : {4, 8}) run_rb<1>("bulk", table, big, 64, thr, st, cyc, sink);
  printf("-- table-size sweep, 64 B rows (TLB reach is 256 MB by the guide; L2 126 MB)\n");
  for (size_t mb : {16, 64, 128, 256, 512, 1024, 3072}) {
    run_rb<0>("cp.async", table, mb << 20, 64, 256, 8, cyc, sink);
    run_rb<1>("bulk", table, mb << 20, 64, 1024, 8, cyc, sink);
  }
  printf("-- row-size sweep, 3 GB table\n");
  for (int rb : {64, 128, 256, 512}) {
    run_rb<0>("cp.async", table, big, rb, 256, 8, cyc, sink);
    run_rb<0>("cp.async", table, big, rb, 1024, 8, cyc, sink);
    run_rb<1>("bulk", table, big, rb, 256, 8, cyc, sink);
    run_rb<1>("bulk", table, big, rb, 1024, 8, cyc, sink);
  }
  if (argc > 1) {
    printf("-- first round (paths side by side, 3 GB table, then 16 MB)\n");
    const uint32_t n_rows = 48u << 20;
    for (int grid : {148, 1}) for (int stages : {4, 8}) {
      for (int thr : {128, 256, 512}) run<0>("cp.async 16 B x 4 per row", table, n_rows, thr, stages, grid, cyc, sink);
      for (int thr : {32, 128, 256}) run<1>("cp.async.bulk 64 B per row", table, n_rows, thr, stages, grid, cyc, sink);
      for (int thr : {128, 256, 512}) run<2>("ld.global.nc.v4 -> st.shared", table, n_rows, thr, stages, grid, cyc, sink);
      for (int thr : {256, 512}) run<3>("half cp.async + half bulk", table, n_rows, thr, stages, grid, cyc, sink);
    }
  }
  return 0;
}
