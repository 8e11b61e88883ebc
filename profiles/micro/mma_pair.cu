// Microbenchmark + correctness probe for tcgen05.mma.cta_group::2 in TS form (A from TMEM, B from shared memory),
// next to cta_group::1, for the issue patterns ctr_fused_kernel uses (wide main MMA + narrow side MMAs per K step).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I recommender_system_b200/csrc -o profiles/micro/mma_pair profiles/micro/mma_pair.cu
// Prints cycles per K step (8 tf32 / 16 bf16 columns) of each pattern, per 128 rows, and checks D = A B^T once per CTA group.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include "tc05.cuh"
using namespace ctrhot::tc;

template <int CG> __device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t db, uint32_t idesc, uint32_t acc, int bf16) {
  if (CG == 1) {
    if (bf16) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                           :: "r"(d), "r"(a), "l"(db), "r"(idesc), "r"(acc) : "memory");
    else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
                      :: "r"(d), "r"(a), "l"(db), "r"(idesc), "r"(acc) : "memory");
  } else {
    if (bf16) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                           :: "r"(d), "r"(a), "l"(db), "r"(idesc), "r"(acc) : "memory");
    else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
                      :: "r"(d), "r"(a), "l"(db), "r"(idesc), "r"(acc) : "memory");
  }
}
template <int CG> __device__ __forceinline__ void commit_all(uint32_t bar) {
  if (CG == 1) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
  else asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                    :: "r"(bar), "h"((uint16_t)3) : "memory");
}
template <int CG> __device__ __forceinline__ void alloc512(uint32_t slot) {
  if (CG == 1) { asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(slot), "r"(512u) : "memory");
                 asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory"); }
  else { asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(slot), "r"(512u) : "memory");
         asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory"); }
}
template <int CG> __device__ __forceinline__ void dealloc512(uint32_t t) {
  if (CG == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(t), "r"(512u) : "memory");
  else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" :: "r"(t), "r"(512u) : "memory");
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xFFFFFFFF;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint32_t ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void csync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__host__ __device__ constexpr uint32_t idesc_of(int bf16, uint32_t M, uint32_t N) {
  return bf16 ? ((1u << 4) | (1u << 7) | (1u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24)) : make_idesc_tf32(M, N);
}
__device__ __forceinline__ void st_cols(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}

// TMEM columns: D main at 0 (256), D side at 256 + 32 * alt index, A at 416 (32 columns = one K chunk of 32 tf32 / 64 bf16)
constexpr uint32_t A_COL = 416, DS_COL = 256;
// shared memory per CTA: B main [256/CG rows][128 B] at 0, B side [32/CG rows][128 B] at 32 KB
// pattern: 0 main only; 1 side only (N = nside), alt accumulators; 2 interleaved (main, side) x4 then lo x4 (the fused kernel's
// first layer); 3 = 4 main, then 4 side, then 4 lo;  4 = main + one N=48 side
// CONV = 0: `if (tid == 0)` issues (divergent branch, as ctr_fused_kernel does today); 1: warp 0 stays converged and the
// MMAs sit under an elect.sync predicate (no per-instruction ELECT / BRA.U.ANY waterfall in the SASS)
template <int CG, int CONV>
__global__ void __launch_bounds__(128, 1) pair_kernel(int pattern, int bf16, int nmain, int nside, int alt, int iters,
                                                      long long* cyc, float* dout, int check) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* bp = smem_raw + (base - raw);
  const uint32_t bm_s = base, bs_s = base + 32768, bars = base + 32768 + 4096, slot = bars + 16;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t rank = (CG == 2) ? ctarank() : 0u;
  if (tid == 0) { mbar_init(bars, 1); mbar_fence_init(); }
  // B values: B[n][k] = ((n * 7 + k * 3) % 11 - 5) / 8  (exact in bf16 and tf32); this CTA holds rows n of its half
  const int nm_loc = nmain / CG, ns_loc = nside / CG;
  const int kelem = bf16 ? 64 : 32;
  for (int i = tid; i < nm_loc * kelem; i += 128) {
    const int n = i / kelem, k = i - n * kelem, gn = n + (int)rank * nm_loc;
    const float v = (float)((gn * 7 + k * 3) % 11 - 5) * 0.125f;
    const uint32_t esz = bf16 ? 2 : 4;
    const uint32_t off = sw128_offset((uint32_t)n, (uint32_t)(k * esz) >> 4) + ((k * esz) & 15);
    if (bf16) *reinterpret_cast<uint16_t*>(bp + off) = (uint16_t)(__float_as_uint(v) >> 16);
    else *reinterpret_cast<float*>(bp + off) = v;
  }
  for (int i = tid; i < ns_loc * kelem; i += 128) {
    const int n = i / kelem, k = i - n * kelem, gn = n + (int)rank * ns_loc;
    const float v = (float)((gn * 5 + k) % 7 - 3) * 0.25f;
    const uint32_t esz = bf16 ? 2 : 4;
    const uint32_t off = 32768 + sw128_offset((uint32_t)n, (uint32_t)(k * esz) >> 4) + ((k * esz) & 15);
    if (bf16) *reinterpret_cast<uint16_t*>(bp + off) = (uint16_t)(__float_as_uint(v) >> 16);
    else *reinterpret_cast<float*>(bp + off) = v;
  }
  fence_proxy_async_smem();
  if (warp == 0) alloc512<CG>(slot);
  tc_fence_before();
  __syncthreads();
  if (CG == 2) csync();
  tc_fence_after();
  const uint32_t tm = *reinterpret_cast<volatile uint32_t*>(bp + (slot - base));
  // A[m][k] = ((m * 3 + k * 5) % 13 - 6) / 4, row m = rank * 128 + tid = TMEM lane tid; 32 columns of 32 bits
  {
    const int m = (int)rank * 128 + tid;
    const uint32_t t_lane = tm + ((uint32_t)(warp * 32) << 16) + A_COL;
    for (int c0 = 0; c0 < 32; c0 += 8) {
      uint32_t v[8];
      for (int j = 0; j < 8; ++j) {
        if (bf16) {
          const int k0 = 2 * (c0 + j);
          const float a0 = (float)((m * 3 + k0 * 5) % 13 - 6) * 0.25f, a1 = (float)((m * 3 + (k0 + 1) * 5) % 13 - 6) * 0.25f;
          v[j] = (__float_as_uint(a0) >> 16) | (__float_as_uint(a1) & 0xFFFF0000u);
        } else {
          v[j] = __float_as_uint((float)((m * 3 + (c0 + j) * 5) % 13 - 6) * 0.25f);
        }
      }
      st_cols(t_lane + c0, v);
    }
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) csync();
  tc_fence_after();
  const uint32_t M = 128u * CG;
  if (rank == 0 && (CONV ? warp == 0 : tid == 0)) {
    const uint32_t id_m = idesc_of(bf16, M, (uint32_t)nmain), id_s = idesc_of(bf16, M, (uint32_t)nside),
                   id_l = idesc_of(bf16, M, (uint32_t)((CG == 2 && nside / 2 < 32) ? 32 : nside / 2));   // 2-CTA TS form: N % 32 == 0
    const uint64_t dbm = make_desc_sw128(bm_s), dbs = make_desc_sw128(bs_s);
    const uint32_t kcols = 8;                      // TMEM columns of A per K step (8 tf32 or 16 bf16 = 32 bytes either way)
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint32_t acc = (check && it == 0) ? 0u : 1u;
      if (CONV && !elect_one()) continue;
      if (pattern == 0) {
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_ts<CG>(tm, tm + A_COL + kcols * k, dbm + 2 * k, id_m, acc | (uint32_t)k, bf16);
      } else if (pattern == 1) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          mma_ts<CG>(tm + DS_COL + 32u * (uint32_t)((it * 4 + k) & (alt - 1)), tm + A_COL + kcols * k, dbs + 2 * k, id_s, acc | (uint32_t)k, bf16);
      } else if (pattern == 2) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          mma_ts<CG>(tm, tm + A_COL + kcols * k, dbm + 2 * k, id_m, acc | (uint32_t)k, bf16);
          mma_ts<CG>(tm + DS_COL, tm + A_COL + kcols * k, dbs + 2 * k, id_s, acc | (uint32_t)k, bf16);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_ts<CG>(tm + DS_COL, tm + A_COL + kcols * k, dbs + 2 * k, id_l, 1u, bf16);
      } else if (pattern == 3) {
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_ts<CG>(tm, tm + A_COL + kcols * k, dbm + 2 * k, id_m, acc | (uint32_t)k, bf16);
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_ts<CG>(tm + DS_COL, tm + A_COL + kcols * k, dbs + 2 * k, id_s, acc | (uint32_t)k, bf16);
#pragma unroll
        for (int k = 0; k < 4; ++k) mma_ts<CG>(tm + DS_COL + 64u, tm + A_COL + kcols * k, dbs + 2 * k, id_l, acc | (uint32_t)k, bf16);
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          mma_ts<CG>(tm, tm + A_COL + kcols * k, dbm + 2 * k, id_m, acc | (uint32_t)k, bf16);
          mma_ts<CG>(tm + DS_COL, tm + A_COL + kcols * k, dbs + 2 * k, id_s, acc | (uint32_t)k, bf16);
        }
      }
    }
    if (CONV) __syncwarp();
    if (!CONV || elect_one()) commit_all<CG>(bars);
    mbar_wait(bars, 0);
    if (tid == 0) cyc[blockIdx.x / CG] = clock64() - t0;
  }
  if (!(rank == 0 && (CONV ? warp == 0 : tid == 0))) mbar_wait(bars, 0);
  tc_fence_after();
  if (check) {
    // D main [M][nmain] and D side [M][nside] of this CTA's 128 rows -> dout[(cluster, rank*128 + tid)][nmain + nside]
    const int m = (int)rank * 128 + tid;
    float* o = dout + ((size_t)(blockIdx.x / CG) * M + m) * (size_t)(nmain + nside);
    const uint32_t t_lane = tm + ((uint32_t)(warp * 32) << 16);
    for (int c0 = 0; c0 < nmain + nside; c0 += 16) {
      float v[16];
      const uint32_t col = c0 < nmain ? (uint32_t)c0 : DS_COL + (uint32_t)(c0 - nmain);
      tmem_ld16(t_lane + col, v);
      for (int j = 0; j < 16; ++j) o[c0 + j] = v[j];
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) csync();
  if (warp == 0) { tc_fence_after(); dealloc512<CG>(tm); }
}

template <int CG, int CONV = 0>
static double run(int pattern, int bf16, int nmain, int nside, int alt, int iters, long long* cyc, float* dout, int check) {
  const size_t smem = 1024 + 32768 + 4096 + 64;
  cudaFuncSetAttribute(pair_kernel<CG, CONV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(148); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, pair_kernel<CG, CONV>, pattern, bf16, nmain, nside, alt, iters, cyc, dout, check);
  if (e != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); exit(1); }
  std::vector<long long> h(148 / CG);
  cudaMemcpy(h.data(), cyc, h.size() * 8, cudaMemcpyDeviceToHost);
  double s = 0; for (auto v : h) s += (double)v;
  return s / h.size() / iters;                 // cycles per iteration (= one 32-column K chunk: 4 K steps)
}

template <int CG>
static void check_once(int bf16, long long* cyc, float* dout) {
  const int nmain = 256, nside = 32, M = 128 * CG, kel = bf16 ? 64 : 32;
  cudaMemset(dout, 0, (size_t)148 * 128 * 288 * 4);
  run<CG>(4, bf16, nmain, nside, 1, 1, cyc, dout, 1);
  std::vector<float> h((size_t)M * (nmain + nside));
  cudaMemcpy(h.data(), dout, h.size() * 4, cudaMemcpyDeviceToHost);   // cluster 0
  double worst = 0;
  for (int m = 0; m < M; ++m) for (int n = 0; n < nmain + nside; ++n) {
    double want = 0;
    for (int k = 0; k < kel; ++k) {
      const double a = (double)((m * 3 + k * 5) % 13 - 6) * 0.25;
      const double b = n < nmain ? (double)((n * 7 + k * 3) % 11 - 5) * 0.125 : (double)(((n - nmain) * 5 + k) % 7 - 3) * 0.25;
      want += a * b;
    }
    worst = fmax(worst, fabs(want - (double)h[(size_t)m * (nmain + nside) + n]));
  }
  printf("check cta_group::%d %s M=%d: max |D - A B^T| = %.3g %s\n", CG, bf16 ? "bf16" : "tf32", M, worst, worst == 0 ? "OK" : "MISMATCH");
}

int main() {
  long long* cyc; cudaMalloc(&cyc, 148 * 8);
  float* dout; cudaMalloc(&dout, (size_t)148 * 128 * 288 * 4);
  for (int bf16 : {0, 1}) { check_once<1>(bf16, cyc, dout); check_once<2>(bf16, cyc, dout); }
  const int iters = 2000;
#define BOTH(CG, ...) run<CG, 0>(__VA_ARGS__), run<CG, 1>(__VA_ARGS__)
  for (int bf16 : {0, 1}) {
    const char* nm = bf16 ? "bf16" : "tf32";
    printf("%s: cycles per 32-column K chunk (4 MMAs of each kind)   cg1 divergent / converged   cg2 divergent / converged\n", nm);
    for (int N : {256, 208, 128, 64})
      printf("  main N=%3d                  : %7.1f %7.1f   %7.1f %7.1f\n", N, BOTH(1, 0, bf16, N, 32, 1, iters, cyc, dout, 0),
             BOTH(2, 0, bf16, N, 32, 1, iters, cyc, dout, 0));
    for (int N : {16, 32}) for (int alt : {1, 4}) {
      printf("  side N=%3d accs=%d           : %7.1f %7.1f", N, alt, BOTH(1, 1, bf16, 256, N, alt, iters, cyc, dout, 0));
      if (N >= 32) printf("   %7.1f %7.1f", BOTH(2, 1, bf16, 256, N, alt, iters, cyc, dout, 0));
      printf("\n");
    }
    printf("  (main256, side32) x4, lo x4 : %7.1f %7.1f   %7.1f %7.1f\n", BOTH(1, 2, bf16, 256, 32, 1, iters, cyc, dout, 0),
           BOTH(2, 2, bf16, 256, 32, 1, iters, cyc, dout, 0));
    printf("  4 main, 4 side32, 4 lo      : %7.1f %7.1f   %7.1f %7.1f\n", BOTH(1, 3, bf16, 256, 32, 1, iters, cyc, dout, 0),
           BOTH(2, 3, bf16, 256, 32, 1, iters, cyc, dout, 0));
    printf("  (main256, side32) x4        : %7.1f %7.1f   %7.1f %7.1f\n", BOTH(1, 4, bf16, 256, 32, 1, iters, cyc, dout, 0),
           BOTH(2, 4, bf16, 256, 32, 1, iters, cyc, dout, 0));
    printf("  (main128, side32) x4        : %7.1f %7.1f   %7.1f %7.1f\n", BOTH(1, 4, bf16, 128, 32, 1, iters, cyc, dout, 0),
           BOTH(2, 4, bf16, 128, 32, 1, iters, cyc, dout, 0));
  }
  return 0;
}
