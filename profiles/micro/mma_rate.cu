// Microbenchmark: cycles per tcgen05.mma kind::tf32 at M=128 for several N, A from smem (SS) or TMEM (TS),
// with and without a concurrent bulk-copy stream into shared memory.  One CTA per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I recommender_system_b200/csrc -o profiles/micro/mma_rate profiles/micro/mma_rate.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "tc05.cuh"
using namespace ctrhot::tc;

__device__ __forceinline__ void mma_bf16_ss(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               :: "r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_bf16_ts(uint32_t d, uint32_t a, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
               :: "r"(d), "r"(a), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
// mode: 0 tf32 SS, 1 tf32 TS, 2 bf16 SS, 3 bf16 TS;  alt: number of distinct accumulators cycled through
__global__ void __launch_bounds__(128, 1) rate_kernel(int N, int ts, int iters, int copy_bytes, const float* src,
                                                      long long* out, int alt, int M) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t a_s = base, b_s = base + 16384, c_s = b_s + 40960, bars = c_s + 65536;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) { mbar_init(bars, 1); mbar_init(bars + 8, 1); mbar_fence_init(); }
  for (int i = tid; i < (16384 + 40960) / 4; i += 128) reinterpret_cast<float*>(smem_raw + (base - raw))[i] = 1.0f;
  fence_proxy_async_smem();
  if (warp == 0) { tmem_alloc(smem_u32(&slot), 512u); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (tid == 0) {
    const uint32_t idesc = (ts < 2) ? make_idesc_tf32((uint32_t)M, (uint32_t)N)
                                    : ((1u << 4) | (1u << 7) | (1u << 10) | (((uint32_t)N >> 3) << 17) | (((uint32_t)M >> 4) << 24));
    const uint64_t da = make_desc_sw128(a_s), db = make_desc_sw128(b_s);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t d = tm + (uint32_t)(((it * 4 + k) % alt) * 64);
        if (ts == 1) mma_tf32_ts(d, tm + 288 + 8 * k, db + 2 * k, idesc, 1u);
        else if (ts == 0) mma_tf32_ss(d, da + 2 * k, db + 2 * k, idesc, 1u);
        else if (ts == 2) mma_bf16_ss(d, da + 2 * k, db + 2 * k, idesc, 1u);
        else mma_bf16_ts(d, tm + 288 + 8 * k, db + 2 * k, idesc, 1u);
      }
    }
    mma_commit(bars);
    mbar_wait(bars, 0);
    out[blockIdx.x] = clock64() - t0;
  } else if (tid == 32 && copy_bytes > 0) {
    // concurrent stream of bulk copies into smem (like the W producer); not synchronised with the MMAs
    for (int it = 0; it < iters; ++it) {
      mbar_arrive_expect_tx(bars + 8, (uint32_t)copy_bytes);
      bulk_g2s(c_s, src + (size_t)(it & 7) * 16384, (uint32_t)copy_bytes, bars + 8);
      mbar_wait(bars + 8, it & 1);
    }
  }
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tm, 512u); }
}

int main() {
  float* src; cudaMalloc(&src, 8 * 65536); cudaMemset(src, 0, 8 * 65536);
  long long* out; cudaMalloc(&out, 148 * 8);
  const size_t smem = 1024 + 16384 + 40960 + 65536 + 64;
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int iters = 2000;
  const char* names[4] = {"tf32 SS", "tf32 TS", "bf16 SS", "bf16 TS"};
  for (int M : {128, 64}) for (int alt : {1, 2, 4}) for (int mode : {0, 1, 2, 3}) for (int N : {16, 32, 64, 128, 208, 256}) {
    if (alt * 64 > 256 && N > 64) continue;
    if (alt > 1 && N > 64) continue;
    if (M == 64 && alt > 1) continue;
    rate_kernel<<<148, 128, smem>>>(N, mode, iters, 0, src, out, alt, M);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
    long long h[148]; cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    double s = 0; for (int i = 0; i < 148; ++i) s += h[i];
    printf("M=%3d accs=%d %s N=%3d : %.1f cycles per MMA\n", M, alt, names[mode], N, s / 148 / (iters * 4.0));
  }
  return 0;
}
