// speaker.cu -- the speaker encoder `SpeakerRAVE` (rave/CombinedRave.py:200-328), the step just
// before the path: model.py:246-247 turns the PQMF bands of a reference utterance into the 256-d
// vector that RAVE.encode appends to every latent frame (model.py:618-620).  It runs once per
// utterance, offline (centred zero padding), in fp32 on the CUDA-core conv kernel of the fp32
// engine plus a few small kernels for the pooling tail.  Architecture constants are the
// reference's (they are literals in CombinedRave.py, not configuration).
#include <math.h>

#include <algorithm>

#include "common.cuh"

namespace rb {

namespace {

struct SpkLayerDef { int c_in, c_out, k, stride, dil, pre_act; };
// forward order; RB_SPK_* in include/rave_b200.h index this table
const SpkLayerDef kLayers[RB_SPK_N_LAYERS] = {
    {16, 128, 7, 1, 1, RB_ACT_NONE},      // in_layer                         :208-214
    {128, 128, 3, 1, 1, RB_ACT_LEAKY},    // layer2 DilatedUnit k3 d=1        :221-225
    {128, 128, 1, 1, 1, RB_ACT_LEAKY},    //        DilatedUnit 1x1 (+ residual)
    {128, 256, 8, 4, 1, RB_ACT_LEAKY},    //        strided conv r=4          :227-231
    {256, 256, 3, 1, 3, RB_ACT_LEAKY},    // layer3 d=3                       :238-242
    {256, 256, 1, 1, 1, RB_ACT_LEAKY},
    {256, 256, 8, 4, 1, RB_ACT_LEAKY},    //        r=4                       :244-248
    {256, 256, 3, 1, 5, RB_ACT_LEAKY},    // layer4 d=5                       :255-259
    {256, 256, 1, 1, 1, RB_ACT_LEAKY},
    {256, 256, 4, 2, 1, RB_ACT_LEAKY},    //        r=2                       :261-265
    {256, 256, 1, 1, 1, RB_ACT_NONE},     // cat_layer                        :267-270
    {768, 768, 3, 1, 1, RB_ACT_NONE},     // out_layer                        :272-275
    {2304, 128, 1, 1, 1, RB_ACT_NONE},    // attention.0                      :284
    {128, 768, 1, 1, 1, RB_ACT_NONE},     // attention.3                      :287
    {1536, 256, 1, 1, 1, RB_ACT_NONE},    // fc6                              :293
};
constexpr int SPK_C = 768;                // attention_projection

// x2 [B,256,2n] max-pooled by 2, x3, x4 [B,256,n] -> y [B,768,n]               (:305, mp2 :299)
__global__ void spk_cat_kernel(const float* __restrict__ x2, const float* __restrict__ x3,
                               const float* __restrict__ x4, float* __restrict__ y, int n, long long total) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int t = (int)(i % n);
  const long long r = i / n;
  const int c = (int)(r % SPK_C);
  const long long b = r / SPK_C;
  float v;
  if (c < 256) {
    const float* p = x2 + ((b * 256 + c) * 2 * (long long)n) + 2 * t;
    v = fmaxf(p[0], p[1]);
  } else if (c < 512) {
    v = x3[(b * 256 + (c - 256)) * (long long)n + t];
  } else {
    v = x4[(b * 256 + (c - 512)) * (long long)n + t];
  }
  y[i] = v;
}

__device__ __forceinline__ float block_sum(float v, float* red) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  v = l < nw ? red[l] : 0.f;
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = blockDim.x >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  v = l < nw ? red[l] : -INFINITY;
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// One CTA per (b, c) row of u = out_layer(...) [B,768,n]: x = LeakyReLU(u) (:307), then the three
// channel blocks of global_x [B,2304,n] = x | mean over time | sqrt(clamp(unbiased var))  (:311-315)
__global__ void __launch_bounds__(128)
spk_global_kernel(const float* __restrict__ u, float* __restrict__ g, int n) {
  __shared__ float red[4];
  const int c = blockIdx.x, b = blockIdx.y;
  const float* ur = u + ((size_t)b * SPK_C + c) * n;
  float* g0 = g + ((size_t)b * 3 * SPK_C + c) * n;
  float* g1 = g0 + (size_t)SPK_C * n;
  float* g2 = g1 + (size_t)SPK_C * n;
  float s = 0.f;
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    float v = ur[t];
    v = v > 0.f ? v : 0.2f * v;
    g0[t] = v;
    s += v;
  }
  const float mean = block_sum(s, red) / (float)n;
  float q = 0.f;
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    float v = ur[t];
    v = (v > 0.f ? v : 0.2f * v) - mean;
    q = fmaf(v, v, q);
  }
  const float var = block_sum(q, red) / (float)(n - 1);
  const float sd = sqrtf(fminf(fmaxf(var, 1e-4f), 1e4f));
  for (int t = threadIdx.x; t < n; t += blockDim.x) { g1[t] = mean; g2[t] = sd; }
}

// a [B,C,n] <- BatchNorm_eval(ReLU(a)), scale = gamma / sqrt(var + eps), shift = beta - mean * scale  (:285-286)
__global__ void spk_relu_bn_kernel(float* __restrict__ a, const float* __restrict__ scale,
                                   const float* __restrict__ shift, int C, int n, long long total) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c = (int)((i / n) % C);
  a[i] = fmaf(fmaxf(a[i], 0.f), scale[c], shift[c]);
}

// One CTA per (b, c): w = softmax over time of the attention logits (:288), mu = sum x w,
// sg = sqrt(clamp(sum x^2 w - mu^2)) (:319-320), then bn5 (eval) on cat(mu, sg) (:322-323)
__global__ void __launch_bounds__(128)
spk_pool_kernel(const float* __restrict__ logits, const float* __restrict__ g, const float* __restrict__ scale,
                const float* __restrict__ shift, float* __restrict__ v, int n) {
  __shared__ float red[4];
  const int c = blockIdx.x, b = blockIdx.y;
  const float* lr = logits + ((size_t)b * SPK_C + c) * n;
  const float* xr = g + ((size_t)b * 3 * SPK_C + c) * n;
  float m = -INFINITY;
  for (int t = threadIdx.x; t < n; t += blockDim.x) m = fmaxf(m, lr[t]);
  m = block_max(m, red);
  float se = 0.f, s1 = 0.f, s2 = 0.f;
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    const float e = expf(lr[t] - m);
    const float x = xr[t];
    se += e;
    s1 = fmaf(x, e, s1);
    s2 = fmaf(x * x, e, s2);
  }
  se = block_sum(se, red);
  s1 = block_sum(s1, red);
  s2 = block_sum(s2, red);
  if (threadIdx.x == 0) {
    const float mu = s1 / se;
    const float sg = sqrtf(fminf(fmaxf(s2 / se - mu * mu, 1e-4f), 1e4f));
    float* vb = v + (size_t)b * 2 * SPK_C;
    vb[c] = fmaf(mu, scale[c], shift[c]);
    vb[SPK_C + c] = fmaf(sg, scale[SPK_C + c], shift[SPK_C + c]);
  }
}

// pqmf.py:13-17 reverse_half: negate odd bands at even times
__global__ void spk_reverse_half_kernel(float* __restrict__ x, int M, int n, long long total) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int t = (int)(i % n), c = (int)((i / n) % M);
  if ((c & 1) && !(t & 1)) x[i] = -x[i];
}

// emb[b][o] = fc6(v[b])  (:324): one warp per output
__global__ void spk_linear_kernel(const float* __restrict__ v, const float* __restrict__ w /*[out][in]*/,
                                  const float* __restrict__ bias, float* __restrict__ y, int n_in, int n_out,
                                  int B) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, l = threadIdx.x & 31;
  if (gw >= B * n_out) return;
  const int b = gw / n_out, o = gw - b * n_out;
  const float* wr = w + (size_t)o * n_in;
  const float* vr = v + (size_t)b * n_in;
  float s = 0.f;
  for (int i = l; i < n_in; i += 32) s = fmaf(wr[i], vr[i], s);
  for (int d = 16; d; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
  if (l == 0) y[gw] = s + bias[o];
}

}  // namespace
}  // namespace rb

using namespace rb;

struct rb_speaker {
  int device = 0;
  float* w[RB_SPK_N_LAYERS] = {};      // convs: packed [c_in][k][c_out]; fc6: [out][in] as given
  float* bias[RB_SPK_N_LAYERS] = {};
  float* bn_scale[2] = {};             // 0: attention.2 (128), 1: bn5 (1536)
  float* bn_shift[2] = {};
  float* pqmf_w = nullptr;             // analysis bank packed [1][k][M]
  int pqmf_k = 0;
  float* bands = nullptr;              // [B,16,T/16] scratch of rb_speaker_embed_audio
  size_t bands_bytes = 0;
  float* ws = nullptr;
  size_t ws_bytes = 0;
  long long launches = 0;
};

static int spk_upload(const float* host, size_t n, float** dev) {
  cudaFree(*dev);
  *dev = nullptr;
  RB_CUDA(cudaMalloc((void**)dev, n * sizeof(float)));
  RB_CUDA(cudaMemcpy(*dev, host, n * sizeof(float), cudaMemcpyHostToDevice));
  return RB_OK;
}

extern "C" {

int rb_speaker_create(int device, rb_speaker** out) {
  RB_CHECK(out, RB_ERR_INVALID, "rb_speaker_create: null out");
  int n_dev = 0;
  RB_CUDA(cudaGetDeviceCount(&n_dev));
  RB_CHECK(device >= 0 && device < n_dev, RB_ERR_INVALID, "device %d of %d", device, n_dev);
  rb_speaker* s = new rb_speaker();
  s->device = device;
  *out = s;
  return RB_OK;
}

int rb_speaker_destroy(rb_speaker* s) {
  if (!s) return RB_OK;
  cudaSetDevice(s->device);
  for (int i = 0; i < RB_SPK_N_LAYERS; ++i) { cudaFree(s->w[i]); cudaFree(s->bias[i]); }
  for (int i = 0; i < 2; ++i) { cudaFree(s->bn_scale[i]); cudaFree(s->bn_shift[i]); }
  cudaFree(s->ws);
  cudaFree(s->pqmf_w);
  cudaFree(s->bands);
  delete s;
  return RB_OK;
}

int rb_speaker_load(rb_speaker* s, int layer, const float* weight, int64_t w_numel, const float* bias,
                    int64_t b_numel) {
  RB_CHECK(s && weight && bias, RB_ERR_INVALID, "rb_speaker_load: null argument");
  RB_CHECK(layer >= 0 && layer < RB_SPK_N_LAYERS, RB_ERR_INVALID, "speaker layer %d out of range", layer);
  const SpkLayerDef& L = kLayers[layer];
  RB_CHECK(w_numel == (int64_t)L.c_out * L.c_in * L.k && b_numel == L.c_out, RB_ERR_INVALID,
           "speaker layer %d: weight %lld / bias %lld elements, expected %d*%d*%d / %d", layer,
           (long long)w_numel, (long long)b_numel, L.c_out, L.c_in, L.k, L.c_out);
  RB_CUDA(cudaSetDevice(s->device));
  int rc;
  if (layer == RB_SPK_FC6) {
    rc = spk_upload(weight, (size_t)w_numel, &s->w[layer]);
  } else {
    std::vector<float> p((size_t)w_numel);          // [c_out][c_in][k] -> [c_in][k][c_out]
    for (int co = 0; co < L.c_out; ++co)
      for (int ci = 0; ci < L.c_in; ++ci)
        for (int j = 0; j < L.k; ++j)
          p[((size_t)ci * L.k + j) * L.c_out + co] = weight[((size_t)co * L.c_in + ci) * L.k + j];
    rc = spk_upload(p.data(), p.size(), &s->w[layer]);
  }
  if (rc) return rc;
  return spk_upload(bias, (size_t)b_numel, &s->bias[layer]);
}

int rb_speaker_load_bn(rb_speaker* s, int which, const float* gamma, const float* beta, const float* mean,
                       const float* var, int n, float eps) {
  RB_CHECK(s && gamma && beta && mean && var, RB_ERR_INVALID, "rb_speaker_load_bn: null argument");
  RB_CHECK((which == 0 && n == 128) || (which == 1 && n == 2 * SPK_C), RB_ERR_INVALID,
           "batch norm %d with %d channels", which, n);
  std::vector<float> sc(n), sh(n);
  for (int i = 0; i < n; ++i) {
    sc[i] = gamma[i] / sqrtf(var[i] + eps);
    sh[i] = beta[i] - mean[i] * sc[i];
  }
  RB_CUDA(cudaSetDevice(s->device));
  int rc = spk_upload(sc.data(), n, &s->bn_scale[which]);
  if (rc) return rc;
  return spk_upload(sh.data(), n, &s->bn_shift[which]);
}

int rb_speaker_load_pqmf(rb_speaker* s, const float* w, int n_band, int k) {
  RB_CHECK(s && w, RB_ERR_INVALID, "rb_speaker_load_pqmf: null argument");
  RB_CHECK(n_band == kLayers[RB_SPK_IN].c_in && k > 0 && (k & 1), RB_ERR_INVALID,
           "speaker PQMF: %d bands (need %d), k = %d (need odd)", n_band, kLayers[RB_SPK_IN].c_in, k);
  std::vector<float> p((size_t)n_band * k);         // [M][1][k] -> [1][k][M]
  for (int m = 0; m < n_band; ++m)
    for (int j = 0; j < k; ++j) p[(size_t)j * n_band + m] = w[(size_t)m * k + j];
  RB_CUDA(cudaSetDevice(s->device));
  s->pqmf_k = k;
  return spk_upload(p.data(), p.size(), &s->pqmf_w);
}

int rb_speaker_launches(rb_speaker* s, int64_t* count) {
  RB_CHECK(s && count, RB_ERR_INVALID, "rb_speaker_launches: null argument");
  *count = s->launches;
  return RB_OK;
}

}  // extern "C"

static int spk_conv(rb_speaker* s, int layer, const float* x, int n_in, float* y, const float* res, int B,
                    cudaStream_t st) {
  const SpkLayerDef& L = kLayers[layer];
  ConvArgs a;
  a.x.p = const_cast<float*>(x); a.x.C = L.c_in; a.x.ld = n_in; a.x.hist = 0; a.x.n = n_in;
  const int n_out = n_in / L.stride;
  a.y.p = y; a.y.C = L.c_out; a.y.ld = n_out; a.y.hist = 0; a.y.n = n_out;
  if (res) { a.res.p = const_cast<float*>(res); a.res.C = L.c_out; a.res.ld = n_out; a.res.hist = 0; a.res.n = n_out; }
  a.w = s->w[layer]; a.bias = s->bias[layer]; a.alpha = nullptr;
  a.B = B; a.c_in = L.c_in; a.c_out = L.c_out; a.k = L.k; a.stride = L.stride; a.dil = L.dil;
  a.off = ((L.k - 1) * L.dil) / 2;      // cached_conv.get_padding centred: left = (p - 1) // 2, p = (k-1) d + 1
  a.ileave = 1; a.pre_act = L.pre_act; a.res_delay = 0; a.lo = 0;
  return launch_conv_f32(a, st);
}

// workspace of one rb_speaker_embed call in floats: three ping-pong buffers sized for the widest stage +
// the tail tensors (every one of them is written by its producer before it is read: nothing needs zeroing)
static size_t spk_ws_floats(int B, int T) {
  const size_t n0 = T, n1 = T / 4, n2 = T / 16, n3 = T / 32;
  const size_t stage = std::max((size_t)128 * n0, (size_t)256 * n1);
  return (size_t)B * (3 * stage + 256 * n2 + 2 * 256 * n3 + 2 * SPK_C * n3 + 3 * SPK_C * n3 +
                      128 * n3 + SPK_C * n3 + 2 * SPK_C);
}

static int spk_check(rb_speaker* s, const float* x, int B, int T, float* emb) {
  RB_CHECK(s && x && emb, RB_ERR_INVALID, "rb_speaker_embed: null argument");
  RB_CHECK(B > 0 && T >= 64 && T % 32 == 0, RB_ERR_INVALID,
           "rb_speaker_embed: T = %d band samples must be a multiple of 32 (the strides) and >= 64", T);
  for (int i = 0; i < RB_SPK_N_LAYERS; ++i)
    RB_CHECK(s->w[i] && s->bias[i], RB_ERR_STATE, "speaker layer %d has no weights", i);
  RB_CHECK(s->bn_scale[0] && s->bn_scale[1], RB_ERR_STATE, "speaker batch norms not loaded");
  return RB_OK;
}

static int spk_embed_ws(rb_speaker* s, const float* x, int B, int T, float* emb, float* ws, cudaStream_t st);

extern "C" int rb_speaker_embed(rb_speaker* s, const float* x, int B, int T, float* emb, void* stream) {
  int rc0 = spk_check(s, x, B, T, emb);
  if (rc0) return rc0;
  RB_CUDA(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  const size_t total = spk_ws_floats(B, T);
  if (total * sizeof(float) > s->ws_bytes) {
    RB_CUDA(cudaStreamSynchronize(st));
    cudaFree(s->ws);
    s->ws = nullptr; s->ws_bytes = 0;
    RB_CUDA(cudaMalloc((void**)&s->ws, total * sizeof(float)));
    s->ws_bytes = total * sizeof(float);
  }
  return spk_embed_ws(s, x, B, T, emb, s->ws, st);
}

// ---- allocation-free variant.  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops).  The
// workspace is ONE file-scope device array, held by every GPU that loads the library until exit and shared by
// every call on that GPU: calls that overlap on the device race.  Each buffer inside it is written by its
// producer before it is read, so nothing left over from the previous call is ever consumed.
static __device__ __align__(256) float g_spk_scratch[RB_SPEAKER_NOALLOC_MAX_FLOATS];

extern "C" int rb_speaker_embed_noalloc(rb_speaker* s, const float* x, int B, int T, float* emb, void* stream) {
  int rc0 = spk_check(s, x, B, T, emb);
  if (rc0) return rc0;
  const size_t total = spk_ws_floats(B, T);
  RB_CHECK(total <= (size_t)RB_SPEAKER_NOALLOC_MAX_FLOATS, RB_ERR_INVALID,
           "rb_speaker_embed_noalloc: B = %d x T = %d band samples needs %zu floats of workspace, the static "
           "scratch holds %zu (about B * T <= 4096); use rb_speaker_embed", B, T, total,
           (size_t)RB_SPEAKER_NOALLOC_MAX_FLOATS);
  RB_CHECK(s->device >= 0 && s->device < 64, RB_ERR_INVALID, "device index %d out of range", s->device);
  RB_CUDA(cudaSetDevice(s->device));
  // In host code `g_spk_scratch` is a host-side stand-in, not a device address: resolve it once per device,
  // with that device current
  static float* scratch_of[64] = {};
  if (!scratch_of[s->device]) {
    float* ptr = nullptr;
    RB_CUDA(cudaGetSymbolAddress((void**)&ptr, g_spk_scratch));
    RB_CHECK(ptr, RB_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch pointer");
    scratch_of[s->device] = ptr;
  }
  return spk_embed_ws(s, x, B, T, emb, scratch_of[s->device], (cudaStream_t)stream);
}

static int spk_embed_ws(rb_speaker* s, const float* x, int B, int T, float* emb, float* ws, cudaStream_t st) {
  const size_t n0 = T, n1 = T / 4, n2 = T / 16, n3 = T / 32;
  const size_t stage = std::max((size_t)128 * n0, (size_t)256 * n1);
  float* p = ws;
  auto take = [&](size_t n) { float* r = p; p += (size_t)B * n; return r; };
  float *a = take(stage), *h = take(stage), *c = take(stage);
  float* x2 = take(256 * n2);
  float *x3 = take(256 * n3), *x4 = take(256 * n3);
  float *cat = take(SPK_C * n3), *u = take(SPK_C * n3), *g = take(3 * SPK_C * n3);
  float *att = take(128 * n3), *logits = take(SPK_C * n3), *v = take(2 * SPK_C);
  const long long l0 = g_launches;
  int rc;
#define SPK(call) do { if ((rc = (call))) return rc; } while (0)
  SPK(spk_conv(s, RB_SPK_IN, x, (int)n0, a, nullptr, B, st));
  // layer2: Residual(DilatedUnit) then the strided conv; a -> h -> c (= a + DU(a)) -> a
  SPK(spk_conv(s, RB_SPK_L2_DIL, a, (int)n0, h, nullptr, B, st));
  SPK(spk_conv(s, RB_SPK_L2_PW, h, (int)n0, c, a, B, st));
  SPK(spk_conv(s, RB_SPK_L2_DOWN, c, (int)n0, a, nullptr, B, st));          // x1 [B,256,n1]
  SPK(spk_conv(s, RB_SPK_L3_DIL, a, (int)n1, h, nullptr, B, st));
  SPK(spk_conv(s, RB_SPK_L3_PW, h, (int)n1, c, a, B, st));
  SPK(spk_conv(s, RB_SPK_L3_DOWN, c, (int)n1, x2, nullptr, B, st));         // x2 [B,256,n2]
  SPK(spk_conv(s, RB_SPK_L4_DIL, x2, (int)n2, h, nullptr, B, st));
  SPK(spk_conv(s, RB_SPK_L4_PW, h, (int)n2, c, x2, B, st));
  SPK(spk_conv(s, RB_SPK_L4_DOWN, c, (int)n2, x3, nullptr, B, st));         // x3 [B,256,n3]
  SPK(spk_conv(s, RB_SPK_CAT, x3, (int)n3, x4, nullptr, B, st));
  {
    const long long tot = (long long)B * SPK_C * n3;
    spk_cat_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(x2, x3, x4, cat, (int)n3, tot);
    ++g_launches;
  }
  SPK(spk_conv(s, RB_SPK_OUT, cat, (int)n3, u, nullptr, B, st));
  spk_global_kernel<<<dim3(SPK_C, B), 128, 0, st>>>(u, g, (int)n3);
  ++g_launches;
  SPK(spk_conv(s, RB_SPK_ATT1, g, (int)n3, att, nullptr, B, st));
  {
    const long long tot = (long long)B * 128 * n3;
    spk_relu_bn_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(att, s->bn_scale[0], s->bn_shift[0], 128,
                                                                      (int)n3, tot);
    ++g_launches;
  }
  SPK(spk_conv(s, RB_SPK_ATT2, att, (int)n3, logits, nullptr, B, st));
  spk_pool_kernel<<<dim3(SPK_C, B), 128, 0, st>>>(logits, g, s->bn_scale[1], s->bn_shift[1], v, (int)n3);
  ++g_launches;
  spk_linear_kernel<<<(B * 256 * 32 + 255) / 256, 256, 0, st>>>(v, s->w[RB_SPK_FC6], s->bias[RB_SPK_FC6], emb,
                                                               2 * SPK_C, 256, B);
  ++g_launches;
#undef SPK
  RB_CUDA(cudaGetLastError());
  s->launches += g_launches - l0;
  return RB_OK;
}

// model.py:246-247: speaker = speaker_encoder(pqmf(utterance)); CachedPQMF.forward = strided conv with the
// analysis bank, centred padding, then reverse_half (pqmf.py:269-273)
extern "C" int rb_speaker_embed_audio(rb_speaker* s, const float* audio, int B, int T, float* emb, void* stream) {
  RB_CHECK(s && audio && emb, RB_ERR_INVALID, "rb_speaker_embed_audio: null argument");
  RB_CHECK(s->pqmf_w, RB_ERR_STATE, "speaker PQMF bank not loaded");
  const int M = kLayers[RB_SPK_IN].c_in;
  RB_CHECK(B > 0 && T >= 2 * M * 32 && T % (M * 32) == 0, RB_ERR_INVALID,
           "rb_speaker_embed_audio: T = %d samples must be a multiple of %d and at least %d", T, M * 32, 2 * M * 32);
  RB_CUDA(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int n = T / M;
  const size_t need = (size_t)B * M * n * sizeof(float);
  if (need > s->bands_bytes) {
    RB_CUDA(cudaStreamSynchronize(st));
    cudaFree(s->bands);
    s->bands = nullptr; s->bands_bytes = 0;
    RB_CUDA(cudaMalloc((void**)&s->bands, need));
    s->bands_bytes = need;
  }
  const long long l0 = g_launches;
  ConvArgs a;
  a.x.p = const_cast<float*>(audio); a.x.C = 1; a.x.ld = T; a.x.hist = 0; a.x.n = T;
  a.y.p = s->bands; a.y.C = M; a.y.ld = n; a.y.hist = 0; a.y.n = n;
  a.w = s->pqmf_w; a.bias = nullptr; a.alpha = nullptr;
  a.B = B; a.c_in = 1; a.c_out = M; a.k = s->pqmf_k; a.stride = M; a.dil = 1; a.off = (s->pqmf_k - 1) / 2;
  a.ileave = 1; a.pre_act = RB_ACT_NONE; a.res_delay = 0; a.lo = 0;
  int rc = launch_conv_f32(a, st);
  if (rc) return rc;
  const long long tot = (long long)B * M * n;
  spk_reverse_half_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(s->bands, M, n, tot);
  ++g_launches;
  RB_CUDA(cudaGetLastError());
  s->launches += g_launches - l0;
  return rb_speaker_embed(s, s->bands, B, n, emb, stream);
}
