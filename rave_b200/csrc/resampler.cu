// resampler.cu -- `rave.resampler.Resampler` (rave/resampler.py:9-66), the Kaiser FIR pair the exported
// model wraps around encode / decode (scripts/export.py:377-378 to_model_sampling_rate, :407-408
// from_model_sampling_rate).  Both directions are one continuous FIR over [stream][time] audio:
//
//   to_model    y[t]              = sum_j down[j]   * x[t*ratio + j - off_d]          (resampler.py:31-41, :58-59)
//   from_model  y[t*ratio + ph]   = sum_j up[ph][j] * x[t + j - off_u]                (resampler.py:43-56, :61-66)
//
// Offline (cached_conv off): off = left half of the centred padding, zeros outside [0, n).
// Streaming (cached_conv.CachedConv1d): off = total padding (+ the stride-alignment delay), and
// x[i < 0] is the tail of the previous chunks, kept per stream in a small history buffer.
// HBM-bound: every input sample is read once from HBM (window staged in shared memory), every
// output written once; fp32 or bf16 samples, fp32 arithmetic.
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "common.cuh"

namespace rb {
namespace {

constexpr int RS_TT = 256;   // threads per CTA = outputs (down) / inputs (up) per CTA

template <typename T> __device__ __forceinline__ float ld_s(const T* p);
template <> __device__ __forceinline__ float ld_s<float>(const float* p) { return __ldg(p); }
template <> __device__ __forceinline__ float ld_s<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void st_s(T* p, float v);
template <> __device__ __forceinline__ void st_s<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void st_s<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }

// window [i0, i0 + W) of stream b into smem: history for i < 0, zero for i >= n
template <typename T>
__device__ __forceinline__ void stage_window(float* xs, const T* xb, const float* hb, int H, int n, int i0, int W) {
  for (int u = threadIdx.x; u < W; u += RS_TT) {
    const int i = i0 + u;
    float v = 0.f;
    if (i >= 0) { if (i < n) v = ld_s<T>(xb + i); }
    else if (hb && i >= -H) v = hb[H + i];
    xs[u] = v;
  }
}

template <typename T>
__global__ void __launch_bounds__(RS_TT)
resample_down_kernel(const T* __restrict__ x, T* __restrict__ y, const float* __restrict__ w, const float* __restrict__ hist,
                     int n, int n_out, int k, int ratio, int off, int H) {
  extern __shared__ float smem[];
  float* ws = smem;                       // [k]
  float* xs = smem + ((k + 3) & ~3);      // [(RS_TT - 1) * ratio + k]
  const int b = blockIdx.y, t0 = blockIdx.x * RS_TT;
  for (int j = threadIdx.x; j < k; j += RS_TT) ws[j] = w[j];
  stage_window<T>(xs, x + (size_t)b * n, hist ? hist + (size_t)b * H : nullptr, H, n, t0 * ratio - off,
                  (RS_TT - 1) * ratio + k);
  __syncthreads();
  const int t = t0 + threadIdx.x;
  if (t >= n_out) return;
  const float* xr = xs + threadIdx.x * ratio;
  float acc = 0.f;
  for (int j = 0; j < k; ++j) acc = fmaf(ws[j], xr[j], acc);
  st_s<T>(y + (size_t)b * n_out + t, acc);
}

template <typename T>
__global__ void __launch_bounds__(RS_TT)
resample_up_kernel(const T* __restrict__ x, T* __restrict__ y, const float* __restrict__ w /*[ratio][k]*/,
                   const float* __restrict__ hist, int n, int k, int ratio, int off, int H) {
  extern __shared__ float smem[];
  float* ws = smem;                              // [ratio * k]
  float* xs = smem + ((ratio * k + 3) & ~3);     // [RS_TT - 1 + k]
  float* ys = xs + ((RS_TT + k + 3) & ~3);       // [RS_TT * ratio]: staged so the stores are coalesced
  const int b = blockIdx.y, t0 = blockIdx.x * RS_TT;
  for (int j = threadIdx.x; j < ratio * k; j += RS_TT) ws[j] = w[j];
  stage_window<T>(xs, x + (size_t)b * n, hist ? hist + (size_t)b * H : nullptr, H, n, t0 - off, RS_TT - 1 + k);
  __syncthreads();
  const float* xr = xs + threadIdx.x;
  for (int ph = 0; ph < ratio; ++ph) {
    const float* wr = ws + ph * k;
    float acc = 0.f;
    for (int j = 0; j < k; ++j) acc = fmaf(wr[j], xr[j], acc);
    ys[threadIdx.x * ratio + ph] = acc;
  }
  __syncthreads();
  const long long o0 = (long long)t0 * ratio, n_out = (long long)n * ratio;
  T* yb = y + (size_t)b * n_out;
  for (int u = threadIdx.x; u < RS_TT * ratio; u += RS_TT)
    if (o0 + u < n_out) st_s<T>(yb + o0 + u, ys[u]);
}

// ---- register-blocked forms for the filter shapes the reference's design produces (ratio 2: k = 39 /
// 2 x 21, ratio 3: k = 57 / 3 x 19).  The generic kernels above read two shared-memory words per FMA
// and are bound by the shared-memory port (measured 1.75 ms for 512 x 2^19 samples, 7 % of HBM
// speed).  Here each thread owns R consecutive outputs (down) / inputs (up), keeps the whole filter
// in registers and reads each window sample ONCE for up to R (x ratio) FMAs.  The window is stored
// transposed in shared memory -- element u at [u % S][u / S], S = samples a thread advances -- so
// that lane t reading its c-th window sample touches word t + c / S of row c % S: conflict-free.
// Accumulation order per output is j ascending, exactly as in the generic kernels.
constexpr int RS_R = 8;
// the filter travels as a kernel parameter: it sits in the constant bank and every tap is an
// immediate constant operand of its FMA (no load, no register)
struct RsFilter { float w[64]; };

template <typename T, int S>
__device__ __forceinline__ void stage_window_t(float* xs, int pitch, const T* xb, const float* hb, int H, int n, int i0,
                                               int W) {
  for (int u = threadIdx.x; u < W; u += RS_TT) {
    const int i = i0 + u;
    float v = 0.f;
    if (i >= 0) { if (i < n) v = ld_s<T>(xb + i); }
    else if (hb && i >= -H) v = hb[H + i];
    xs[(u % S) * pitch + u / S] = v;
  }
}

template <typename T, int N>
__device__ __forceinline__ void store_run(T* y, long long o, long long n_out, const float* v) {
  if (o + N <= n_out) {
    if constexpr (sizeof(T) == 4) {
#pragma unroll
      for (int q = 0; q < N / 4; ++q)
        reinterpret_cast<float4*>(y + o)[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
    } else {
#pragma unroll
      for (int q = 0; q < N / 8; ++q) {
        __align__(16) __nv_bfloat16 e[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) e[i] = __float2bfloat16(v[8 * q + i]);
        reinterpret_cast<uint4*>(y + o)[q] = *reinterpret_cast<const uint4*>(e);
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < N; ++i)
      if (o + i < n_out) st_s<T>(y + o + i, v[i]);
  }
}

template <typename T, int RATIO, int K>
__global__ void __launch_bounds__(RS_TT)
resample_down_fast_kernel(const T* __restrict__ x, T* __restrict__ y, const __grid_constant__ RsFilter f,
                          const float* __restrict__ hist, int n, int n_out, int off, int H) {
  constexpr int S = RS_R * RATIO, WIN = (RS_R - 1) * RATIO + K;
  constexpr int PITCH = (RS_TT + (WIN + S - 1) / S) | 1;
  __shared__ float xs[S * PITCH];
  const int b = blockIdx.y;
  const long long o0 = (long long)blockIdx.x * RS_TT * RS_R;
  stage_window_t<T, S>(xs, PITCH, x + (size_t)b * n, hist ? hist + (size_t)b * H : nullptr, H, n,
                       (int)(o0 * RATIO) - off, (RS_TT - 1) * S + WIN);
  static_assert(K <= 64, "filter parameter block");
  __syncthreads();
  const long long o = o0 + (long long)threadIdx.x * RS_R;
  if (o >= n_out) return;
  float acc[RS_R];
#pragma unroll
  for (int r = 0; r < RS_R; ++r) acc[r] = 0.f;
#pragma unroll
  for (int c = 0; c < WIN; ++c) {
    const float xv = xs[(c % S) * PITCH + threadIdx.x + c / S];
#pragma unroll
    for (int r = 0; r < RS_R; ++r) {
      const int j = c - r * RATIO;
      if (j >= 0 && j < K) acc[r] = fmaf(f.w[j], xv, acc[r]);
    }
  }
  store_run<T, RS_R>(y + (size_t)b * n_out, o, n_out, acc);
}

template <typename T, int RATIO, int K>
__global__ void __launch_bounds__(RS_TT)
resample_up_fast_kernel(const T* __restrict__ x, T* __restrict__ y, const __grid_constant__ RsFilter f /*[RATIO][K]*/,
                        const float* __restrict__ hist, int n, int off, int H) {
  constexpr int S = RS_R, WIN = RS_R - 1 + K;
  constexpr int PITCH = (RS_TT + (WIN + S - 1) / S) | 1;
  __shared__ float xs[S * PITCH];
  const int b = blockIdx.y;
  const long long t0 = (long long)blockIdx.x * RS_TT * RS_R;
  stage_window_t<T, S>(xs, PITCH, x + (size_t)b * n, hist ? hist + (size_t)b * H : nullptr, H, n, (int)t0 - off,
                       (RS_TT - 1) * S + WIN);
  static_assert(RATIO * K <= 64, "filter parameter block");
  __syncthreads();
  const long long t = t0 + (long long)threadIdx.x * RS_R;
  if (t >= n) return;
  float acc[RS_R * RATIO];
#pragma unroll
  for (int r = 0; r < RS_R * RATIO; ++r) acc[r] = 0.f;
#pragma unroll
  for (int c = 0; c < WIN; ++c) {
    const float xv = xs[(c % S) * PITCH + threadIdx.x + c / S];
#pragma unroll
    for (int r = 0; r < RS_R; ++r) {
      const int j = c - r;
      if (j >= 0 && j < K) {
#pragma unroll
        for (int ph = 0; ph < RATIO; ++ph) acc[r * RATIO + ph] = fmaf(f.w[ph * K + j], xv, acc[r * RATIO + ph]);
      }
    }
  }
  store_run<T, RS_R * RATIO>(y + (size_t)b * n * RATIO, t * RATIO, (long long)n * RATIO, acc);
}

// new history = last H samples of (old history ++ x[0, n)); old and new are different buffers
template <typename T>
__global__ void resample_hist_kernel(const T* __restrict__ x, const float* __restrict__ old_h, float* __restrict__ new_h,
                                     int n, int H, int B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * H) return;
  const int b = i / H, u = i - b * H;
  const int src = n - H + u;                  // index into x, negative: old history
  new_h[i] = src >= 0 ? ld_s<T>(x + (size_t)b * n + src) : old_h[(size_t)b * H + H + src];
}

}  // namespace
}  // namespace rb

using namespace rb;

struct rb_resampler {
  int device = 0, ratio = 0, k_down = 0, k_up = 0, streaming = 0, max_streams = 0;
  int off_down = 0, off_up = 0, H_down = 0, H_up = 0;
  float *w_down = nullptr, *w_up = nullptr;
  RsFilter f_down, f_up;                       // host copies for the register-blocked kernels (first 64 taps)
  float* hist_down[2] = {nullptr, nullptr};   // ping-pong [max_streams][H]
  float* hist_up[2] = {nullptr, nullptr};
  int cur_down = 0, cur_up = 0;
  long long launches = 0;
};

extern "C" {

int rb_resampler_create(int device, int ratio, const float* down_w, int k_down, const float* up_w, int k_up,
                        int streaming, int max_streams, rb_resampler** out) {
  RB_CHECK(out && down_w && up_w, RB_ERR_INVALID, "rb_resampler_create: null argument");
  RB_CHECK(ratio >= 2 && ratio <= 16 && k_down >= 1 && k_down <= 4096 && k_up >= 1 && k_up <= 1024, RB_ERR_INVALID,
           "resampler: ratio %d, filter lengths %d / %d out of range", ratio, k_down, k_up);
  RB_CHECK(!streaming || max_streams > 0, RB_ERR_INVALID, "streaming resampler needs max_streams > 0");
  // resampler.py:20-23: the reference refuses odd ratios when cached_conv is on
  RB_CHECK(!streaming || ratio % 2 == 0, RB_ERR_INVALID,
           "When using streaming mode, resampling ratio must be a power of 2, got %d", ratio);
  int n_dev = 0;
  RB_CUDA(cudaGetDeviceCount(&n_dev));
  RB_CHECK(device >= 0 && device < n_dev, RB_ERR_INVALID, "device %d of %d", device, n_dev);
  RB_CUDA(cudaSetDevice(device));
  rb_resampler* r = new rb_resampler();
  r->device = device; r->ratio = ratio; r->k_down = k_down; r->k_up = k_up;
  r->streaming = streaming ? 1 : 0; r->max_streams = max_streams;
  // cached_conv.get_padding (centred): left (k-1)//2, right k//2; CachedConv1d: history = total padding
  // + stride delay (s - right % s) % s, and the delayed stream is what the filter sees
  const int pl_d = (k_down - 1) / 2, pr_d = k_down / 2, pl_u = (k_up - 1) / 2, pr_u = k_up / 2;
  if (streaming) {
    const int delay = (ratio - (pr_d % ratio)) % ratio;
    r->off_down = pl_d + pr_d + delay; r->H_down = r->off_down;
    r->off_up = pl_u + pr_u; r->H_up = r->off_up;
  } else {
    r->off_down = pl_d; r->off_up = pl_u;
  }
  auto fail = [&](int rc) { rb_resampler_destroy(r); return rc; };
  if (cudaMalloc((void**)&r->w_down, k_down * sizeof(float)) != cudaSuccess ||
      cudaMalloc((void**)&r->w_up, (size_t)ratio * k_up * sizeof(float)) != cudaSuccess) {
    set_error("resampler: cudaMalloc failed");
    return fail(RB_ERR_CUDA);
  }
  memset(&r->f_down, 0, sizeof(RsFilter));
  memset(&r->f_up, 0, sizeof(RsFilter));
  for (int j = 0; j < k_down && j < 64; ++j) r->f_down.w[j] = down_w[j];
  for (int j = 0; j < ratio * k_up && j < 64; ++j) r->f_up.w[j] = up_w[j];
  cudaMemcpy(r->w_down, down_w, k_down * sizeof(float), cudaMemcpyHostToDevice);
  cudaMemcpy(r->w_up, up_w, (size_t)ratio * k_up * sizeof(float), cudaMemcpyHostToDevice);
  if (streaming) {
    for (int i = 0; i < 2; ++i) {
      const size_t bd = (size_t)max_streams * std::max(r->H_down, 1) * sizeof(float);
      const size_t bu = (size_t)max_streams * std::max(r->H_up, 1) * sizeof(float);
      if (cudaMalloc((void**)&r->hist_down[i], bd) != cudaSuccess || cudaMalloc((void**)&r->hist_up[i], bu) != cudaSuccess) {
        set_error("resampler: cudaMalloc failed");
        return fail(RB_ERR_CUDA);
      }
      cudaMemset(r->hist_down[i], 0, bd);
      cudaMemset(r->hist_up[i], 0, bu);
    }
  }
  RB_CUDA(cudaDeviceSynchronize());
  *out = r;
  return RB_OK;
}

int rb_resampler_destroy(rb_resampler* r) {
  if (!r) return RB_OK;
  cudaSetDevice(r->device);
  cudaFree(r->w_down); cudaFree(r->w_up);
  for (int i = 0; i < 2; ++i) { cudaFree(r->hist_down[i]); cudaFree(r->hist_up[i]); }
  delete r;
  return RB_OK;
}

int rb_resampler_reset(rb_resampler* r, void* stream) {
  RB_CHECK(r, RB_ERR_INVALID, "rb_resampler_reset: null handle");
  if (!r->streaming) return RB_OK;
  RB_CUDA(cudaSetDevice(r->device));
  for (int i = 0; i < 2; ++i) {
    RB_CUDA(cudaMemsetAsync(r->hist_down[i], 0, (size_t)r->max_streams * std::max(r->H_down, 1) * sizeof(float), (cudaStream_t)stream));
    RB_CUDA(cudaMemsetAsync(r->hist_up[i], 0, (size_t)r->max_streams * std::max(r->H_up, 1) * sizeof(float), (cudaStream_t)stream));
  }
  return RB_OK;
}

int rb_resampler_launches(rb_resampler* r, int64_t* count) {
  RB_CHECK(r && count, RB_ERR_INVALID, "rb_resampler_launches: null argument");
  *count = r->launches;
  return RB_OK;
}

}  // extern "C"

template <typename T>
static int resample_run(rb_resampler* r, bool down, const T* x, int B, int n, T* y, cudaStream_t st) {
  const int k = down ? r->k_down : r->k_up, H = down ? r->H_down : r->H_up, off = down ? r->off_down : r->off_up;
  float** hist = down ? r->hist_down : r->hist_up;
  int& cur = down ? r->cur_down : r->cur_up;
  const float* h = (r->streaming && H > 0) ? hist[cur] : nullptr;
  if (down) {
    const int n_out = n / r->ratio;
    const size_t smem = (((k + 3) & ~3) + (size_t)(RS_TT - 1) * r->ratio + k) * sizeof(float);
    static bool attr = false;
    if (!attr) {
      RB_CUDA(cudaFuncSetAttribute(resample_down_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      attr = true;
    }
    RB_CHECK(smem <= 96 * 1024, RB_ERR_UNSUPPORTED, "resampler window needs %zu B of shared memory", smem);
    const dim3 gf((n_out + RS_TT * RS_R - 1) / (RS_TT * RS_R), B);
    // vector stores: every stream's output row must start 16-byte aligned
    const bool fast = !getenv("RB_RESAMPLE_GENERIC") && n_out % (16 / (int)sizeof(T)) == 0 && ((uintptr_t)y & 15) == 0;
    if (fast && r->ratio == 2 && k == 39)
      resample_down_fast_kernel<T, 2, 39><<<gf, RS_TT, 0, st>>>(x, y, r->f_down, h, n, n_out, off, H);
    else if (fast && r->ratio == 3 && k == 57)
      resample_down_fast_kernel<T, 3, 57><<<gf, RS_TT, 0, st>>>(x, y, r->f_down, h, n, n_out, off, H);
    else
    resample_down_kernel<T><<<dim3((n_out + RS_TT - 1) / RS_TT, B), RS_TT, smem, st>>>(x, y, r->w_down, h, n, n_out, k,
                                                                                   r->ratio, off, H);
  } else {
    const size_t smem = (((r->ratio * k + 3) & ~3) + ((RS_TT + k + 3) & ~3) + (size_t)RS_TT * r->ratio) * sizeof(float);
    static bool attr = false;
    if (!attr) {
      RB_CUDA(cudaFuncSetAttribute(resample_up_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      attr = true;
    }
    RB_CHECK(smem <= 96 * 1024, RB_ERR_UNSUPPORTED, "resampler window needs %zu B of shared memory", smem);
    const dim3 gf((n + RS_TT * RS_R - 1) / (RS_TT * RS_R), B);
    const bool fast = !getenv("RB_RESAMPLE_GENERIC") && ((long long)n * r->ratio) % (16 / (int)sizeof(T)) == 0 &&
                      ((uintptr_t)y & 15) == 0;
    if (fast && r->ratio == 2 && k == 21)
      resample_up_fast_kernel<T, 2, 21><<<gf, RS_TT, 0, st>>>(x, y, r->f_up, h, n, off, H);
    else if (fast && r->ratio == 3 && k == 19)
      resample_up_fast_kernel<T, 3, 19><<<gf, RS_TT, 0, st>>>(x, y, r->f_up, h, n, off, H);
    else
    resample_up_kernel<T><<<dim3((n + RS_TT - 1) / RS_TT, B), RS_TT, smem, st>>>(x, y, r->w_up, h, n, k, r->ratio, off, H);
  }
  ++r->launches;
  if (h) {
    resample_hist_kernel<T><<<(B * H + 255) / 256, 256, 0, st>>>(x, hist[cur], hist[cur ^ 1], n, H, B);
    // rows >= B keep their state in the other buffer: copy them across so the swap loses nothing
    if (B < r->max_streams)
      RB_CUDA(cudaMemcpyAsync(hist[cur ^ 1] + (size_t)B * H, hist[cur] + (size_t)B * H,
                              (size_t)(r->max_streams - B) * H * sizeof(float), cudaMemcpyDeviceToDevice, st));
    cur ^= 1;
    ++r->launches;
  }
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

static int resample_entry(rb_resampler* r, bool down, const void* x, int B, int n, int dtype, void* y, void* stream) {
  RB_CHECK(r && x && y, RB_ERR_INVALID, "rb_resampler: null argument");
  RB_CHECK(B > 0 && n > 0, RB_ERR_INVALID, "rb_resampler: B = %d, n = %d", B, n);
  RB_CHECK(!down || n % r->ratio == 0, RB_ERR_INVALID, "rb_resampler_to_model: n = %d is not a multiple of the ratio %d",
           n, r->ratio);
  RB_CHECK(!r->streaming || B <= r->max_streams, RB_ERR_INVALID, "rb_resampler: %d streams, allocated for %d", B,
           r->max_streams);
  RB_CHECK(dtype == RB_F32 || dtype == RB_BF16, RB_ERR_INVALID, "rb_resampler: dtype %d", dtype);
  RB_CUDA(cudaSetDevice(r->device));
  if (dtype == RB_F32) return resample_run<float>(r, down, (const float*)x, B, n, (float*)y, (cudaStream_t)stream);
  return resample_run<__nv_bfloat16>(r, down, (const __nv_bfloat16*)x, B, n, (__nv_bfloat16*)y, (cudaStream_t)stream);
}

extern "C" int rb_resampler_to_model(rb_resampler* r, const void* x, int B, int n, int dtype, void* y, void* stream) {
  return resample_entry(r, true, x, B, n, dtype, y, stream);
}
extern "C" int rb_resampler_from_model(rb_resampler* r, const void* x, int B, int n, int dtype, void* y, void* stream) {
  return resample_entry(r, false, x, B, n, dtype, y, stream);
}
