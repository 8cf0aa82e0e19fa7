// latent.cu -- export-time latent post/pre-processing (scripts/export.py:482-539), the steps either side of
// encode / decode in the scripted wrapper (in this fork they are defined but not on the live `myforward`
// path):
//
//   rb_latent_post / rb_latent_pre      VariationalScriptedRAVE.post_process_latent / pre_process_latent
//                                       :484-499: z = pca[:n] (z - mean)   /   z = pca^T cat(z, noise) + mean
//   rb_unit_norm_to_angles / _inverse   rave/blocks.py:926-956 (SphericalScriptedRAVE :533-539)
//
// (DiscreteScriptedRAVE :502-517 is rb_rvq_encode / rb_rvq_decode plus a clamp; the Wasserstein variant is a
// noise concat.)  Latents are tiny ([B, <=128, frames]); one thread per output element / per (stream, frame).
#include <math.h>

#include "common.cuh"

namespace rb {
namespace {

// out[b][o][t] = sum_c W[o][c] (z[b][c][t] - sub[c]) + add[o];   W row stride ldw, transposed: W[c][o]
__global__ void latent_matmul_kernel(const float* __restrict__ z, const float* __restrict__ z2, int C1, int C,
                                     const float* __restrict__ W, int transposed, const float* __restrict__ sub,
                                     const float* __restrict__ add, float* __restrict__ out, int O, int F,
                                     long long total) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int t = (int)(i % F);
  const long long r = i / F;
  const int o = (int)(r % O);
  const long long b = r / O;
  float acc = 0.f;
  for (int c = 0; c < C; ++c) {
    // input channel c: first C1 channels from z, the rest from z2 (noise; zeros when absent)
    float v = c < C1 ? z[(b * C1 + c) * F + t] : (z2 ? z2[(b * (C - C1) + (c - C1)) * F + t] : 0.f);
    if (sub) v -= sub[c];
    acc = fmaf(transposed ? W[(size_t)c * C + o] : W[(size_t)o * C + c], v, acc);
  }
  out[i] = acc + (add ? add[o] : 0.f);
}

// blocks.py:926-939
__global__ void to_angles_kernel(const float* __restrict__ x, float* __restrict__ y, int C, int F, long long total) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int t = (int)(i % F);
  const long long b = i / F;
  const float* xb = x + b * C * F + t;
  float* yb = y + b * (C - 1) * F + t;
  const float pi = 3.14159265358979323846f;
  float s = xb[(size_t)(C - 1) * F];
  s = s * s;
  for (int j = C - 2; j >= 0; --j) {
    const float xj = xb[(size_t)j * F];
    s += xj * xj;                                  // norms[j]^2 = sum_{m >= j} x[m]^2, accumulated from the tail
    float a = acosf(xj / sqrtf(s));
    if (j == C - 2) {
      if (!(xb[(size_t)(C - 1) * F] >= 0.f)) a = 2.f * pi - a;
      a = a / (2.f * pi);
    } else {
      a = a / pi;
    }
    yb[(size_t)j * F] = 2.f * (a - .5f);
  }
}

// blocks.py:942-956
__global__ void from_angles_kernel(const float* __restrict__ a, float* __restrict__ y, int C, int F, long long total) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int t = (int)(i % F);
  const long long b = i / F;
  const float* ab = a + b * (C - 1) * F + t;
  float* yb = y + b * C * F + t;
  const float pi = 3.14159265358979323846f;
  float prod = 1.f;
  for (int j = 0; j < C - 1; ++j) {
    float v = ab[(size_t)j * F] / 2.f + .5f;
    v = v - floorf(v);                             // python % 1
    v *= (j == C - 2) ? 2.f * pi : pi;
    yb[(size_t)j * F] = cosf(v) * prod;
    prod *= sinf(v);
  }
  yb[(size_t)(C - 1) * F] = prod;
}

}  // namespace
}  // namespace rb

using namespace rb;

struct rb_latent {
  int device = 0, C = 0;
  float *pca = nullptr, *mean = nullptr;           // [C][C], [C]
};

extern "C" {

int rb_latent_create(int device, int C, const float* pca, const float* mean, rb_latent** out) {
  RB_CHECK(out && pca && mean && C > 0 && C <= 4096, RB_ERR_INVALID, "rb_latent_create: bad argument");
  RB_CUDA(cudaSetDevice(device));
  rb_latent* h = new rb_latent();
  h->device = device; h->C = C;
  if (cudaMalloc((void**)&h->pca, (size_t)C * C * sizeof(float)) != cudaSuccess ||
      cudaMalloc((void**)&h->mean, C * sizeof(float)) != cudaSuccess) {
    cudaFree(h->pca);
    delete h;
    set_error("rb_latent_create: cudaMalloc failed");
    return RB_ERR_CUDA;
  }
  cudaMemcpy(h->pca, pca, (size_t)C * C * sizeof(float), cudaMemcpyHostToDevice);
  cudaMemcpy(h->mean, mean, C * sizeof(float), cudaMemcpyHostToDevice);
  *out = h;
  return RB_OK;
}

int rb_latent_destroy(rb_latent* h) {
  if (!h) return RB_OK;
  cudaSetDevice(h->device);
  cudaFree(h->pca); cudaFree(h->mean);
  delete h;
  return RB_OK;
}

// post_process_latent after reparametrize (export.py:486-489): out [B,n_keep,F] = pca[:n_keep] (z - mean)
int rb_latent_post(rb_latent* h, const float* z, int B, int F, int n_keep, float* out, void* stream) {
  RB_CHECK(h && z && out && B > 0 && F > 0 && n_keep > 0 && n_keep <= h->C, RB_ERR_INVALID, "rb_latent_post: bad argument");
  RB_CUDA(cudaSetDevice(h->device));
  const long long total = (long long)B * n_keep * F;
  latent_matmul_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      z, nullptr, h->C, h->C, h->pca, 0, h->mean, nullptr, out, n_keep, F, total);
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

// pre_process_latent (export.py:491-499): out [B,C,F] = pca^T cat(z [B,n_keep,F], noise [B,C-n_keep,F]) + mean;
// the reference draws the noise with randn -- the caller passes it (NULL: zeros)
int rb_latent_pre(rb_latent* h, const float* z, const float* noise, int B, int F, int n_keep, float* out, void* stream) {
  RB_CHECK(h && z && out && B > 0 && F > 0 && n_keep > 0 && n_keep <= h->C, RB_ERR_INVALID, "rb_latent_pre: bad argument");
  RB_CUDA(cudaSetDevice(h->device));
  const long long total = (long long)B * h->C * F;
  latent_matmul_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      z, noise, n_keep, h->C, h->pca, 1, nullptr, h->mean, out, h->C, F, total);
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

int rb_unit_norm_to_angles(const float* x, int B, int C, int F, float* angles, void* stream) {
  RB_CHECK(x && angles && B > 0 && C >= 2 && F > 0, RB_ERR_INVALID, "rb_unit_norm_to_angles: bad argument");
  const long long total = (long long)B * F;
  to_angles_kernel<<<(unsigned)((total + 127) / 128), 128, 0, (cudaStream_t)stream>>>(x, angles, C, F, total);
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

int rb_angles_to_unit_norm(const float* angles, int B, int C, int F, float* x, void* stream) {
  RB_CHECK(x && angles && B > 0 && C >= 2 && F > 0, RB_ERR_INVALID, "rb_angles_to_unit_norm: bad argument");
  const long long total = (long long)B * F;
  from_angles_kernel<<<(unsigned)((total + 127) / 128), 128, 0, (cudaStream_t)stream>>>(angles, x, C, F, total);
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  return RB_OK;
}

}  // extern "C"
