// pitch.cu -- the pitch front-end that produces the f0 conditioning channels (rave/pitch_utils.py):
//
//   rb_yin_estimate   estimate() :15-42 = _frame :45-49, _diff :52-75, _search :77-91
//   rb_f0_one_hot     extract_utterance_log_f0 / quantize_f0_norm ('abs') / get_f0_norm :116-135
//
// YIN: one CTA per (stream, frame).  The reference gets the autocorrelation from an FFT; here it is
// the direct sum  corr[tau] = sum_j x[j] x[j + tau]  over the zero-extended frame, register-blocked
// (each thread 4 consecutive lags x 4 consecutive j per step: three 16-byte shared-memory loads per
// 16 FMAs, conflict-free) -- same quantity, different rounding, so periods agree with the reference
// except where its own fp32 noise decides.  Everything after it (energy terms from the prefix sums of
// x^2, cumulative-mean normalisation, threshold + first local minimum search) follows the reference
// line by line.
#include <math.h>

#include <algorithm>

#include "common.cuh"

namespace rb {
namespace {

constexpr int YIN_TT = 256;

__device__ __forceinline__ float warp_scan_incl(float v) {
  const int l = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float u = __shfl_up_sync(0xffffffffu, v, o);
    if (l >= o) v += u;
  }
  return v;
}

// in-place inclusive prefix sum of a[0, n) in shared memory; `part` holds YIN_TT / 32 floats
__device__ void block_scan(float* a, int n, float* part) {
  const int chunk = (n + YIN_TT - 1) / YIN_TT;
  const int i0 = threadIdx.x * chunk, i1 = min(i0 + chunk, n);
  float s = 0.f;
  for (int i = i0; i < i1; ++i) s += a[i];
  const float incl = warp_scan_incl(s);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 31) part[w] = incl;
  __syncthreads();
  float base = 0.f;
  for (int q = 0; q < w; ++q) base += part[q];
  float run = base + incl - s;
  for (int i = i0; i < i1; ++i) { run += a[i]; a[i] = run; }
  __syncthreads();
}

__global__ void __launch_bounds__(YIN_TT)
yin_kernel(const float* __restrict__ x, float* __restrict__ f0, int T, int n_frames, int L, int stride, int tau_min,
           int tau_max, float threshold, float sample_rate, int LP) {
  extern __shared__ __align__(16) float sm[];
  float* xs = sm;                  // [LP] frame, zero-extended (LP multiple of 4, >= L + 4 * YIN_TT + 8)
  float* sq = xs + LP;             // [L + 1] prefix sums of x^2 (sq[0] = 0)
  float* d = sq + ((L + 4) & ~3);  // [tau_max] difference function, then cmdf
  __shared__ float part[YIN_TT / 32];
  __shared__ int s_first, s_pick;
  const int b = blockIdx.y, fr = blockIdx.x;
  const float* xb = x + (size_t)b * T + (size_t)fr * stride;
  const int avail = T - fr * stride;           // samples of the signal inside the frame (rest: zero padding)
  for (int i = threadIdx.x; i < LP; i += YIN_TT) xs[i] = (i < L && i < avail) ? __ldg(xb + i) : 0.f;
  if (threadIdx.x == 0) { sq[0] = 0.f; s_first = 0x7fffffff; s_pick = 0x7fffffff; }
  __syncthreads();
  for (int i = threadIdx.x; i < L; i += YIN_TT) sq[i + 1] = xs[i] * xs[i];
  __syncthreads();
  block_scan(sq + 1, L, part);
  // autocorrelation, lags [g0 + 4 t, g0 + 4 t + 4)
  for (int g0 = 0; g0 < tau_max; g0 += 4 * YIN_TT) {
    const int tau0 = g0 + 4 * threadIdx.x;
    if (tau0 < tau_max) {
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
      const float4* xa = reinterpret_cast<const float4*>(xs);
      const float4* xt = reinterpret_cast<const float4*>(xs + tau0);
      for (int j4 = 0; j4 < (L + 3) / 4; ++j4) {
        const float4 u = xa[j4];
        const float4 p = xt[j4], q = xt[j4 + 1];
        a0 = fmaf(u.x, p.x, a0); a1 = fmaf(u.x, p.y, a1); a2 = fmaf(u.x, p.z, a2); a3 = fmaf(u.x, p.w, a3);
        a0 = fmaf(u.y, p.y, a0); a1 = fmaf(u.y, p.z, a1); a2 = fmaf(u.y, p.w, a2); a3 = fmaf(u.y, q.x, a3);
        a0 = fmaf(u.z, p.z, a0); a1 = fmaf(u.z, p.w, a1); a2 = fmaf(u.z, q.x, a2); a3 = fmaf(u.z, q.y, a3);
        a0 = fmaf(u.w, p.w, a0); a1 = fmaf(u.w, q.x, a1); a2 = fmaf(u.w, q.y, a2); a3 = fmaf(u.w, q.z, a3);
      }
      const float acc[4] = {a0, a1, a2, a3};
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int tau = tau0 + r;
        // eq. 6 as upstream writes it: corr_0 + (sqrcs[L - tau] - sqrcs[tau]) - 2 corr[tau]
        if (tau < tau_max) d[tau] = sq[L] + (sq[L - tau] - sq[tau]) - 2.f * acc[r];
      }
    }
  }
  __syncthreads();
  // eq. 8: cmdf[i] = d[i] * i / max(cumsum(d[1..i]), 1e-5), i = 1 .. tau_max - 1; kept in d[i]
  const int n1 = tau_max - 1;
  float mine[8];                                // d[i] of this thread's chunk survives the scan in registers
  const int chunk = (n1 + YIN_TT - 1) / YIN_TT; // <= 8 for tau_max <= 2048; larger handled by the slow path below
  const bool small = chunk <= 8;
  const int i0 = 1 + threadIdx.x * chunk, i1 = min(i0 + chunk, tau_max);
  if (small)
    for (int i = i0; i < i1; ++i) mine[i - i0] = d[i];
  float* cum = small ? d : sq;                  // large tau_max: scan a copy (sq is free now, L + 1 >= tau_max)
  if (!small) {
    for (int i = 1 + threadIdx.x; i < tau_max; i += YIN_TT) sq[i] = d[i];
    __syncthreads();
  }
  block_scan(cum + 1, n1, part);
  if (small) {
    for (int i = i0; i < i1; ++i) d[i] = mine[i - i0] * (float)i / fmaxf(d[i], 1e-5f);
  } else {
    for (int i = 1 + threadIdx.x; i < tau_max; i += YIN_TT) d[i] = d[i] * (float)i / fmaxf(sq[i], 1e-5f);
  }
  __syncthreads();
  // search over c[k] = cmdf index k <-> d[k + 1 + tau_min], k in [0, n_c)
  const int n_c = tau_max - 1 - tau_min;
  const float* c = d + 1 + tau_min;
  int fb = 0x7fffffff;
  for (int k = threadIdx.x; k < n_c; k += YIN_TT)
    if (c[k] < threshold) { fb = k; break; }
  if (fb != 0x7fffffff) atomicMin(&s_first, fb);
  __syncthreads();
  int first = s_first;
  if (first == 0x7fffffff || first == 0) first = tau_max;      // argmax == 0 reads as "none below"
  int pick = 0x7fffffff;
  for (int k = first + threadIdx.x; k < n_c; k += YIN_TT) {
    const bool rising = (k == n_c - 1) || (c[k + 1] - c[k] >= 0.f);
    if (rising) { pick = k; break; }
  }
  if (pick != 0x7fffffff) atomicMin(&s_pick, pick);
  __syncthreads();
  if (threadIdx.x == 0) {
    const int tau = s_pick == 0x7fffffff ? 0 : s_pick;
    // `sample_rate / tensor` is torch's __rtruediv__: reciprocal(tensor) * sample_rate, two fp32 roundings
    f0[(size_t)b * n_frames + fr] = tau > 0 ? __fmul_rn(__frcp_rn((float)(tau + tau_min + 1)), sample_rate) : 0.f;
  }
}

// get_f0_norm: f0 [B][F] -> one-hot over bins + 1 classes and log_f0_norm (NaN where unvoiced)
__global__ void f0_one_hot_kernel(const float* __restrict__ f0, float* __restrict__ oh, float* __restrict__ lognorm,
                                  int F, int bins, int channels_first, long long total, float log40, float den) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int C = bins + 1;
  long long bf; int cls;
  if (channels_first) { const int t = (int)(i % F); const long long r = i / F; cls = (int)(r % C); bf = (r / C) * F + t; }
  else { cls = (int)(i % C); bf = i / C; }
  const float f = f0[bf];
  int idx = bins;                                     // NaN (unvoiced) sorts past the last boundary
  float ln = nanf("");
  if (f != 0.f && f == f) {
    ln = (logf(f) - log40) / den + 0.5f;
    // torch.bucketize(v, linspace(0, 1, bins + 1), right=True) - 1: number of boundaries <= v, minus one
    int cnt = 0;
    if (ln == ln) {
      int k = (int)floorf(ln * (float)bins);
      k = max(-1, min(k, bins));
      // boundaries are j / bins as linspace makes them: fix the guess against the exact comparisons
      while (k + 1 <= bins && (float)(k + 1) / (float)bins <= ln) ++k;
      while (k >= 0 && (float)k / (float)bins > ln) --k;
      cnt = k + 1;
    } else cnt = bins + 1;
    idx = cnt - 1;
  }
  if (idx < 0) idx += C;                              // python negative index of the eye() lookup (:110-113)
  oh[i] = cls == idx ? 1.f : 0.f;
  if (lognorm && cls == 0) lognorm[bf] = ln;
}

}  // namespace
}  // namespace rb

using namespace rb;

extern "C" int rb_yin_frames(int T, int sample_rate, double pitch_min, double frame_stride_s, int* n_frames) {
  RB_CHECK(n_frames && T > 0 && sample_rate > 0 && pitch_min > 0, RB_ERR_INVALID, "rb_yin_frames: bad argument");
  const int tau_max = (int)(sample_rate / pitch_min), L = 2 * tau_max, stride = (int)(frame_stride_s * sample_rate);
  RB_CHECK(stride > 0, RB_ERR_INVALID, "frame stride of %g s is less than one sample", frame_stride_s);
  *n_frames = (std::max(T, L) - L) / stride + 1;
  return RB_OK;
}

extern "C" int rb_yin_estimate(const float* x, int B, int T, int sample_rate, double pitch_min, double pitch_max,
                               double frame_stride_s, double threshold, float* f0, int n_frames, void* stream) {
  RB_CHECK(x && f0 && B > 0 && B <= 65535 && T > 0, RB_ERR_INVALID, "rb_yin_estimate: bad argument (1 <= B <= 65535)");
  RB_CHECK(sample_rate > 0 && pitch_min > 0 && pitch_max > pitch_min, RB_ERR_INVALID, "rb_yin_estimate: pitch range");
  const int tau_min = (int)(sample_rate / pitch_max), tau_max = (int)(sample_rate / pitch_min);
  const int L = 2 * tau_max, stride = (int)(frame_stride_s * sample_rate);
  RB_CHECK(stride > 0, RB_ERR_INVALID, "frame stride of %g s is less than one sample", frame_stride_s);
  RB_CHECK(tau_max - 1 - tau_min >= 2, RB_ERR_INVALID, "pitch range leaves no lags to search");
  const int want = (std::max(T, L) - L) / stride + 1;
  RB_CHECK(n_frames == want, RB_ERR_INVALID, "rb_yin_estimate: %d frames given, the signal has %d (rb_yin_frames)",
           n_frames, want);
  const int groups = (tau_max + 4 * YIN_TT - 1) / (4 * YIN_TT);
  const int LP = (std::max(L, (groups - 1) * 4 * YIN_TT + 4 * (YIN_TT - 1) + 4) + ((L + 3) & ~3) + 8 + 3) & ~3;
  const size_t smem = ((size_t)LP + ((L + 4) & ~3) + ((tau_max + 3) & ~3)) * sizeof(float);
  RB_CHECK(smem <= 200 * 1024, RB_ERR_UNSUPPORTED, "YIN frame of %d samples needs %zu B of shared memory (pitch_min too low)",
           L, smem);
  static bool attr = false;
  if (!attr) {
    RB_CUDA(cudaFuncSetAttribute(yin_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr = true;
  }
  yin_kernel<<<dim3(n_frames, B), YIN_TT, smem, (cudaStream_t)stream>>>(x, f0, T, n_frames, L, stride, tau_min, tau_max,
                                                                     (float)threshold, (float)sample_rate, LP);
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

extern "C" int rb_f0_one_hot(const float* f0, int B, int F, int num_bins, int channels_first, float* one_hot,
                             float* log_f0_norm, void* stream) {
  RB_CHECK(f0 && one_hot && B > 0 && F > 0 && num_bins > 0, RB_ERR_INVALID, "rb_f0_one_hot: bad argument");
  const long long total = (long long)B * F * (num_bins + 1);
  // quantize_f0_norm 'abs' (:121-122): (log f0 - log 40) / log(400 - log 40), in fp32 like the tensors upstream
  const float log40 = logf(40.f), den = logf(400.f - log40);
  f0_one_hot_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(f0, one_hot, log_f0_norm, F, num_bins,
                                                                                    channels_first, total, log40, den);
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}
