// Shared device/host helpers for libctrhot (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
#include <stdint.h>
#include "ctrhot.h"

namespace ctrhot {

constexpr int kWarp = 32;
constexpr int kNumSMs = 148;   // B200

// host: record + translate CUDA errors (thread-local text, no global mutable state shared across threads)
void set_cuda_error(cudaError_t e, const char* where);
void set_error_text(const char* text);
#define CTRHOT_CUDA_OK(expr, where)                         \
  do {                                                      \
    cudaError_t _e = (expr);                                \
    if (_e != cudaSuccess) {                                \
      ::ctrhot::set_cuda_error(_e, where);                  \
      return CTRHOT_ECUDA;                                  \
    }                                                       \
  } while (0)
#define CTRHOT_LAUNCH_OK(where) CTRHOT_CUDA_OK(cudaGetLastError(), where)

// NVTX range around every compute entry point (SURVEY 5 build note): named like the symbol, so a timeline reads in the C-ABI's
// vocabulary.  Header-only nvtx3: a few nanoseconds when no tool is attached.
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};
#define CTRHOT_RANGE(name) ::ctrhot::NvtxRange _ctrhot_nvtx_range(name)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Transport-record field descriptors (csrc/transport.cu): byte-granular  (byte offset) | (width in bytes) << 16,  or, with
// bit 31 set, bit-granular  (bit offset, 24 bits) | (width in bits, <= 32) << 24.  -> bit offset inside the record, width in bits.
__host__ __device__ __forceinline__ void rows_field(uint32_t d, uint32_t& bit_off, uint32_t& bits) {
  if (d & 0x80000000u) { bit_off = d & 0xFFFFFFu; bits = (d >> 24) & 0x7Fu; }
  else { bit_off = (d & 0xFFFFu) * 8u; bits = (d >> 16) * 8u; }
}

// ---------------------------------------------------------------- device
__device__ __forceinline__ float4 ldg_nc_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ float ldg_nc_f1(const float* p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_na_f4(float* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + __expf(-x)); }

// ids arrive as int32 or as the reference's float columns (keras Embedding casts to int32)
template <int KIND> struct IdLoad;
template <> struct IdLoad<CTRHOT_IDS_I32> {
  static __device__ __forceinline__ int64_t at(const void* p, int64_t i) {
    return static_cast<int64_t>(__ldg(static_cast<const int32_t*>(p) + i));
  }
};
template <> struct IdLoad<CTRHOT_IDS_F32> {
  static __device__ __forceinline__ int64_t at(const void* p, int64_t i) {
    return static_cast<int64_t>(__float2int_rz(__ldg(static_cast<const float*>(p) + i)));
  }
};

}  // namespace ctrhot
