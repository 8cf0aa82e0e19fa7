// common.cuh -- shared declarations of the rave_b200 engine (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/rave_b200.h"

namespace rb {

// thread-local error message behind rb_last_error()
void set_error(const char* fmt, ...);

#define RB_CUDA(call)                                                                  \
  do {                                                                                 \
    cudaError_t e__ = (call);                                                          \
    if (e__ != cudaSuccess) {                                                          \
      rb::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
                    __LINE__);                                                         \
      return RB_ERR_CUDA;                                                              \
    }                                                                                  \
  } while (0)

#define RB_CHECK(cond, code, ...)   \
  do {                              \
    if (!(cond)) {                  \
      rb::set_error(__VA_ARGS__);   \
      return (code);                \
    }                               \
  } while (0)

// A view of one activation tensor in the fp32 engine: [B][C][ld], time innermost.
// Element (b, c, t) lives at p[(b*C + c)*ld + hist + t], t in [-hist, n).
struct View {
  float* p = nullptr;
  int C = 0;
  int ld = 0;
  int hist = 0;
  int n = 0;       // valid new samples per row for this call
};

struct ConvArgs {
  View x, y, res;            // res.p == nullptr when there is no residual
  const float* w;            // packed [c_in][k][c_out]
  const float* bias;         // [c_out] or nullptr
  const float* alpha;        // Snake alpha [c_in] or nullptr
  int B, c_in, c_out, k, stride, dil, off, ileave, pre_act, res_delay;
  int lo;                    // lowest valid input index (0 offline, -hist streaming)
};

// fp32 CUDA-core kernels (kernels_f32.cu)
int launch_conv_f32(const ConvArgs& a, cudaStream_t s);
int launch_pqmf_fwd_f32(const View& x, const View& y, const float* w /*[k][8]*/, int B, int n_keep,
                        int n_band, int k, int off, int lo, cudaStream_t s);
int launch_pqmf_inv_f32(const View& x, const View& y, const float* w /*[c][k][m]*/, int B, int n_band,
                        int k, int off, int lo, cudaStream_t s);
int launch_affine_f32(const View& x, const float* a, const float* b, int B, cudaStream_t s);
// AdaIN running statistics (blocks.py:870-873,897-914): per (stream, channel) mean and unbiased std of
// x over time, folded into rows [0, B) of the tables as target += (value - target) / (count + 1)
int launch_adain_stats_f32(const View& x, int B, float* mean_t, float* std_t, int count, cudaStream_t s);
// a = std_y / (std_x + 1e-5), b = mean_y - mean_x * a  (identity when the transfer is inactive)
int launch_adain_ab(const float* stats, int rows, int C, int active, float* a, float* b, cudaStream_t s);
int launch_head_v2_f32(const View& w, const View& noise, const View& y, int B, int data, int am,
                       cudaStream_t s);
int launch_head_v1_f32(const View& wave, const View& loud, const View& y, int B, int data,
                       cudaStream_t s);
int launch_noise_v2_f32(const View& amp, const float* noise, int64_t noise_b_stride, const View& y,
                        int B, int data, unsigned long long seed, cudaStream_t s);
// Conditioning channels CONCAT_COND appends behind the latent (model.py:615-620, scripts/export.py:393-397):
// c_spk speaker channels, constant over time -- row b0 + b of the table, or its only row (stride 0) --
// then f0_C pitch channels read from f0 [streams][f0_C][f0_F] at the frame index (nullptr: zeros).
struct CondArgs {
  const float* spk;
  long long spk_stride;
  const float* f0;
  int c_spk, f0_C, f0_F;
  int b0;                   // first stream of the running microbatch
};
int launch_concat_f32(const View& z, const View& y, const CondArgs& ca, int B, int c_z, int c_out,
                      cudaStream_t s);
int launch_copy_rows_f32(const float* src, int src_ld, float* dst, int dst_ld, int rows, int n,
                         cudaStream_t s);
int launch_roll_f32(float* p, int rows, int ld, int hist, int n, cudaStream_t s);

// RVQ (rvq.cu)
int launch_rvq_encode(const float* z, int B, int D, int F, const float* codebooks, const float* e2,
                      int Q, int K, int64_t* idx, float* residual_ws, cudaStream_t s);
int launch_rvq_decode(const int64_t* idx, int B, int D, int F, const float* codebooks, int Q, int K,
                      float* zq, cudaStream_t s);
int launch_rvq_norms(const float* codebooks, int rows, int D, float* e2, cudaStream_t s);

extern thread_local long long g_launches;   // counted by every launch_* above

}  // namespace rb
