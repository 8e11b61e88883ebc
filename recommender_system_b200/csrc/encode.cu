// The input step on the left of the hot path: Criteo TSV text -> MinMax-scaled dense features and
// label-encoded ids, on the device.  Reference: create_criteo_dataset, DL/utils/dataset.py:36-66 -- pandas read_csv (:38),
// fillna 0 / '-1' (:42-43), sklearn MinMaxScaler on the 13 counts (:46), sklearn LabelEncoder per categorical column (:51-52).
//
//   parse : one thread per text line: label, 13 integer counts (empty -> missing), 26 tokens of 8 lowercase hex digits (empty ->
//           missing).  A token is kept as the 32-bit value of its digits: for fixed-width lowercase hex the string order
//           LabelEncoder sorts by IS the numeric order, and '-1' (the fill value) sorts before every digit string.
//   encode: dense = fp32( x * scale + min_ ) evaluated in float64 exactly as MinMaxScaler.transform does (scale = 1 / (max - min),
//           1 where the column is constant; min_ = -min * scale); id = rank of the token in the column's sorted distinct values
//           (missing = the fill value '-1' = rank 0 when the column has any), by binary search.  A token that is not in the
//           vocabulary sets *err_flag and gets id -1 (LabelEncoder.transform raises on unseen labels).
// The vocabularies and min / max are FIT once per data set (host side of this file's callers: utils/dataset.py); encode is
// what runs per batch.
#include "common.cuh"

namespace ctrhot {

constexpr int ENC_D = 13, ENC_F = 26;

__global__ void criteo_parse_kernel(const uint8_t* __restrict__ text, const int64_t* __restrict__ line_start, int64_t N,
                                    int has_label, int8_t* __restrict__ label, double* __restrict__ dense_raw,
                                    int64_t* __restrict__ tokens, int* __restrict__ err_flag) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < N; r += (int64_t)gridDim.x * blockDim.x) {
    const uint8_t* p = text + line_start[r];
    const uint8_t* end = text + line_start[r + 1];             // one past the line's '\n' (or the end of the buffer)
    while (end > p && (end[-1] == '\n' || end[-1] == '\r')) --end;
    bool bad = false;
    const int n_fields = (has_label ? 1 : 0) + ENC_D + ENC_F;
    for (int c = 0; c < n_fields; ++c) {
      const uint8_t* q = p;
      while (q < end && *q != '\t') ++q;                        // field = [p, q)
      const int col = c - (has_label ? 1 : 0);
      if (col < 0) {
        label[r] = (q > p && *p == '1') ? 1 : 0;
      } else if (col < ENC_D) {
        if (q == p) {
          dense_raw[r * ENC_D + col] = __longlong_as_double(0x7FF8000000000000LL);     // NaN = missing (pandas' NaN)
        } else {
          const uint8_t* s = p;
          const bool neg = (*s == '-');
          if (neg) ++s;
          long long v = 0;
          if (s == q) bad = true;
          for (; s < q; ++s) {
            if (*s < '0' || *s > '9') { bad = true; break; }
            v = v * 10 + (*s - '0');
          }
          dense_raw[r * ENC_D + col] = (double)(neg ? -v : v);
        }
      } else {
        const int f = col - ENC_D;
        if (q == p) {
          tokens[r * ENC_F + f] = -1;                           // missing -> the fill value '-1', which sorts first
        } else {
          long long v = 0;
          if (q - p != 8) bad = true;
          for (const uint8_t* s = p; s < q; ++s) {
            const int ch = *s;
            const int d = (ch >= '0' && ch <= '9') ? ch - '0' : (ch >= 'a' && ch <= 'f') ? ch - 'a' + 10 : -1;
            if (d < 0) { bad = true; break; }
            v = (v << 4) | d;
          }
          tokens[r * ENC_F + f] = v;
        }
      }
      p = (q < end) ? q + 1 : end;
      if (q >= end && c + 1 < n_fields) {
        // a short line: the remaining fields are missing (pandas fills trailing columns with NaN)
        for (int c2 = c + 1; c2 < n_fields; ++c2) {
          const int col2 = c2 - (has_label ? 1 : 0);
          if (col2 < ENC_D) dense_raw[r * ENC_D + col2] = __longlong_as_double(0x7FF8000000000000LL);
          else tokens[r * ENC_F + (col2 - ENC_D)] = -1;
        }
        break;
      }
    }
    if (bad && err_flag) *err_flag = 1;
  }
}

__global__ void criteo_encode_kernel(const double* __restrict__ dense_raw, const int64_t* __restrict__ tokens, int64_t N,
                                     const double* __restrict__ d_scale, const double* __restrict__ d_min,
                                     const int64_t* __restrict__ vocab, const int64_t* __restrict__ vocab_off,
                                     float* __restrict__ dense, int64_t dense_ld, int32_t* __restrict__ ids, int64_t ids_ld,
                                     int* __restrict__ err_flag) {
  const int C = ENC_D + ENC_F;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N * C; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / C;
    const int c = (int)(i - r * C);
    if (c < ENC_D) {
      double x = dense_raw[r * ENC_D + c];
      if (x != x) x = 0.0;                                                 // fillna(0), dataset.py:42
      // MinMaxScaler.transform: X *= scale_; X += min_  -- two rounded float64 operations, not an fma
      const double y = __dadd_rn(__dmul_rn(x, d_scale[c]), d_min[c]);
      dense[r * dense_ld + c] = (float)y;
    } else {
      const int f = c - ENC_D;
      const int64_t t = tokens[r * ENC_F + f];
      const int64_t* v = vocab + vocab_off[f];
      int64_t lo = 0, hi = vocab_off[f + 1] - vocab_off[f];                // first index with v[idx] >= t
      while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (__ldg(v + mid) < t) lo = mid + 1; else hi = mid;
      }
      const bool found = lo < vocab_off[f + 1] - vocab_off[f] && __ldg(v + lo) == t;
      if (!found && err_flag) *err_flag = 1;
      ids[r * ids_ld + f] = found ? (int32_t)lo : -1;
    }
  }
}

// ---------------------------------------------------------------- Hash (DL/layer/utils.py:21-56)
// tf.strings.to_hash_bucket_fast = FarmHash Fingerprint64 (farmhashna::Hash64) % num_buckets of the value's string form
// (tf.as_string: decimal for integers).  The string sits in 64-bit words w[] (little endian, zero past the end); Fetch64 at a
// byte offset is a funnel shift of two words.  Lengths <= 64 (HashLen0to16 / 17to32 / 33to64): every decimal integer is <= 20.
namespace fh {
constexpr uint64_t k0 = 0xc3a5c85c97cb3127ull, k1 = 0xb492b66fbe98f273ull, k2 = 0x9ae16a3b2f90404full;
__device__ __forceinline__ uint64_t rot(uint64_t v, int s) { return (v >> s) | (v << (64 - s)); }
__device__ __forceinline__ uint64_t f64(const uint64_t* w, int off) {
  const int i = off >> 3, sh = (off & 7) * 8;
  const uint64_t lo = w[i] >> sh;
  return sh ? (lo | (w[i + 1] << (64 - sh))) : lo;
}
__device__ __forceinline__ uint64_t f32(const uint64_t* w, int off) { return f64(w, off) & 0xffffffffull; }
__device__ __forceinline__ uint64_t byte_at(const uint64_t* w, int off) { return (w[off >> 3] >> ((off & 7) * 8)) & 0xffull; }
__device__ __forceinline__ uint64_t hl16(uint64_t u, uint64_t v, uint64_t mul) {
  uint64_t a = (u ^ v) * mul;
  a ^= a >> 47;
  uint64_t b = (v ^ a) * mul;
  b ^= b >> 47;
  return b * mul;
}
__device__ uint64_t fingerprint64(const uint64_t* w, int n) {
  const uint64_t mul = k2 + 2ull * (uint64_t)n;
  if (n == 0) return k2;
  if (n < 4) {
    const uint32_t y = (uint32_t)byte_at(w, 0) + ((uint32_t)byte_at(w, n >> 1) << 8);
    const uint32_t z = (uint32_t)n + ((uint32_t)byte_at(w, n - 1) << 2);
    const uint64_t v = ((uint64_t)y * k2) ^ ((uint64_t)z * k0);
    return (v ^ (v >> 47)) * k2;
  }
  if (n < 8) return hl16((uint64_t)n + (f32(w, 0) << 3), f32(w, n - 4), mul);
  if (n <= 16) {
    const uint64_t a = f64(w, 0) + k2, b = f64(w, n - 8);
    return hl16(rot(b, 37) * mul + a, (rot(a, 25) + b) * mul, mul);
  }
  if (n <= 32) {
    const uint64_t a = f64(w, 0) * k1, b = f64(w, 8), c = f64(w, n - 8) * mul, d = f64(w, n - 16) * k2;
    return hl16(rot(a + b, 43) + rot(c, 30) + d, a + rot(b + k2, 18) + c, mul);
  }
  const uint64_t a = f64(w, 0) * k2, b = f64(w, 8), c = f64(w, n - 8) * mul, d = f64(w, n - 16) * k2;
  const uint64_t y = rot(a + b, 43) + rot(c, 30) + d;
  const uint64_t z = hl16(y, a + rot(b + k2, 18) + c, mul);
  const uint64_t e = f64(w, 16) * mul, f = f64(w, 24), g = (y + f64(w, n - 32)) * mul, h = (z + f64(w, n - 24)) * mul;
  return hl16(rot(e + f, 43) + rot(g, 30) + h, e + rot(f + a, 18) + g, mul);
}
}  // namespace fh

template <typename T>
__global__ void hash_bucket_int_kernel(const T* __restrict__ x, int64_t N, uint64_t nb, int mask_zero, int64_t* __restrict__ out) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < N; r += (int64_t)gridDim.x * blockDim.x) {
    const long long v = (long long)x[r];
    unsigned long long mag = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
    // decimal digits, most significant first, after an optional '-': tf.as_string
    int nd = 1;
    for (unsigned long long t = mag; t >= 10ull; t /= 10ull) ++nd;
    const int n = nd + (v < 0 ? 1 : 0);
    uint64_t w[4] = {0ull, 0ull, 0ull, 0ull};
    for (int pos = n - 1; pos >= (v < 0 ? 1 : 0); --pos) {
      w[pos >> 3] |= (uint64_t)(48 + (int)(mag % 10ull)) << ((pos & 7) * 8);
      mag /= 10ull;
    }
    if (v < 0) w[0] |= (uint64_t)'-';
    const uint64_t h = fh::fingerprint64(w, n) % nb;
    out[r] = mask_zero ? (v == 0 ? 0ll : (long long)(h + 1ull)) : (long long)h;
  }
}

__global__ void hash_bucket_bytes_kernel(const uint8_t* __restrict__ s, int64_t stride, const int32_t* __restrict__ len, int64_t N,
                                         uint64_t nb, int mask_zero, int64_t* __restrict__ out, int* __restrict__ err_flag) {
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < N; r += (int64_t)gridDim.x * blockDim.x) {
    const int n = len[r];
    if (n < 0 || n > 64 || n > stride) {
      if (err_flag) *err_flag = 1;
      out[r] = -1;
      continue;
    }
    uint64_t w[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) w[i] = 0ull;
    const uint8_t* p = s + r * stride;
    for (int i = 0; i < n; ++i) w[i >> 3] |= (uint64_t)p[i] << ((i & 7) * 8);
    const uint64_t h = fh::fingerprint64(w, n) % nb;
    const bool is_zero = (n == 1 && p[0] == (uint8_t)'0');
    out[r] = mask_zero ? (is_zero ? 0ll : (long long)(h + 1ull)) : (long long)h;
  }
}

}  // namespace ctrhot

using namespace ctrhot;

extern "C" int ctrhot_criteo_parse_tsv(const void* text, const int64_t* line_start, int64_t N, int has_label, int8_t* label,
                                       double* dense_raw, int64_t* tokens, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_criteo_parse_tsv");
  if (!text || !line_start || N < 0 || !dense_raw || !tokens || (has_label && !label)) return CTRHOT_EARG;
  if (N == 0) return CTRHOT_OK;
  int64_t grid = (N + 127) / 128;
  if (grid > (int64_t)kNumSMs * 16) grid = (int64_t)kNumSMs * 16;
  criteo_parse_kernel<<<(int)grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(text), line_start, N, has_label, label, dense_raw, tokens, err_flag);
  CTRHOT_LAUNCH_OK("ctrhot_criteo_parse_tsv");
  return CTRHOT_OK;
}

extern "C" int ctrhot_criteo_encode(const double* dense_raw, const int64_t* tokens, int64_t N, const double* d_scale,
                                    const double* d_min, const int64_t* vocab, const int64_t* vocab_off, float* dense,
                                    int64_t dense_ld, int32_t* ids, int64_t ids_ld, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_criteo_encode");
  if (!dense_raw || !tokens || N < 0 || !d_scale || !d_min || !vocab || !vocab_off || !dense || !ids || dense_ld < ENC_D ||
      ids_ld < ENC_F) return CTRHOT_EARG;
  if (N == 0) return CTRHOT_OK;
  const int64_t n = N * (ENC_D + ENC_F);
  int64_t grid = (n + 255) / 256;
  if (grid > (int64_t)kNumSMs * 16) grid = (int64_t)kNumSMs * 16;
  criteo_encode_kernel<<<(int)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(dense_raw, tokens, N, d_scale, d_min, vocab,
                                                                                 vocab_off, dense, dense_ld, ids, ids_ld, err_flag);
  CTRHOT_LAUNCH_OK("ctrhot_criteo_encode");
  return CTRHOT_OK;
}

extern "C" int ctrhot_hash_bucket_int(const void* x, int is_i64, int64_t N, int64_t num_buckets, int mask_zero, int64_t* out,
                                      void* stream) {
  CTRHOT_RANGE("ctrhot_hash_bucket_int");
  if (!x || !out || N < 0) return CTRHOT_EARG;
  const int64_t nb = num_buckets - (mask_zero ? 1 : 0);
  if (nb < 1) return CTRHOT_EARG;
  if (N == 0) return CTRHOT_OK;
  int64_t grid = (N + 255) / 256;
  if (grid > (int64_t)kNumSMs * 16) grid = (int64_t)kNumSMs * 16;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (is_i64) hash_bucket_int_kernel<int64_t><<<(int)grid, 256, 0, st>>>(static_cast<const int64_t*>(x), N, (uint64_t)nb, mask_zero, out);
  else hash_bucket_int_kernel<int32_t><<<(int)grid, 256, 0, st>>>(static_cast<const int32_t*>(x), N, (uint64_t)nb, mask_zero, out);
  CTRHOT_LAUNCH_OK("ctrhot_hash_bucket_int");
  return CTRHOT_OK;
}

extern "C" int ctrhot_hash_bucket_bytes(const uint8_t* s, int64_t stride, const int32_t* len, int64_t N, int64_t num_buckets,
                                        int mask_zero, int64_t* out, int* err_flag, void* stream) {
  CTRHOT_RANGE("ctrhot_hash_bucket_bytes");
  if (!s || !len || !out || N < 0 || stride < 1) return CTRHOT_EARG;
  const int64_t nb = num_buckets - (mask_zero ? 1 : 0);
  if (nb < 1) return CTRHOT_EARG;
  if (N == 0) return CTRHOT_OK;
  int64_t grid = (N + 127) / 128;
  if (grid > (int64_t)kNumSMs * 16) grid = (int64_t)kNumSMs * 16;
  hash_bucket_bytes_kernel<<<(int)grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(s, stride, len, N, (uint64_t)nb, mask_zero,
                                                                                     out, err_flag);
  CTRHOT_LAUNCH_OK("ctrhot_hash_bucket_bytes");
  return CTRHOT_OK;
}
