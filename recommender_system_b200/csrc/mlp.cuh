// Internal interfaces shared by the Dense-tower translation units.
#pragma once
#include "common.cuh"

namespace ctrhot {

struct TcGeom { int KC, n_tiles, BN, stages, tmem_cols; };
TcGeom tc_geom(int K, int N);

// one Dense layer, exact fp32 on CUDA cores (mlp_simt.cu)
int launch_dense_f32(const float* X, int64_t x_ld, int64_t M, const ctrhot_dense_desc& d, float* Y, int64_t y_ld,
                     cudaStream_t st);
// one Dense layer on tcgen05 (mlp_tc.cu); dot_w != NULL fuses a trailing Dense(N->1) into the epilogue
int launch_dense_tc(const float* X, int64_t x_ld, int64_t M, const ctrhot_dense_desc& d, float* Y, int64_t y_ld,
                    const float* dot_w, const float* dot_b, float* y_dot, cudaStream_t st);

}  // namespace ctrhot
