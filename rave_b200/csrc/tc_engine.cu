// tc_engine.cu -- bf16 tensor-core engine: layer planning / weight packing for the tcgen05
// implicit-GEMM kernel (tc_conv.cu), the small group-major elementwise kernels, and the
// op interpreter for dtype == RB_BF16.
//
// Activation layout: group-major  [stream][Cp/8][row][8 channels]  bf16, Cp = channels rounded
// up to 16 (audio slots: Cp = 1, the caller's dense [stream][T]).  A stride-s conv reads its
// input through the view [rows/s][s*Cp], a transposed conv writes its output through
// [rows][r*Cp'] -- same bytes, no copies -- so every conv of the path becomes a stride-1
// row-shift GEMM (tc_conv.cuh).  LeakyReLU nets store tensors already activated (Plan::sact).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "engine.h"
#include "tc_conv.cuh"

namespace rb {
namespace tc {

static inline int round16(int v) { return (v + 15) & ~15; }
static inline int floordiv(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

// One convolution-like op, planned: geometry in VIEW units + packed device weights.
struct Layer {
  bool valid = false;
  int fin = 1, fout = 1;          // view factors of the input / output slot
  int cin_v = 0;                  // view input channels (fin * Cp_in)
  int n_taps = 1, dil = 1, off = 0;
  int N_total = 0, N_TILE = 0, n_ntiles = 0;
  int KC = 0, n_kchunks = 0;
  int out_mask = 0;
  std::vector<int> nt_base;       // [n_ntiles + 1]
  int total_blocks = 0;
  uint32_t blk_bytes = 0;
  __nv_bfloat16* w = nullptr;     // device, packed blocks
  __nv_bfloat16* w2 = nullptr;    // the same blocks split by output rows for cta_group::2 pairs (N_TILE = 256:
                                  // [half][KC/8][128][8] -- each CTA of a pair loads and holds one half), or nullptr
  uint64_t* masks = nullptr;      // device [n_ntiles][n_kchunks]
  float* bias = nullptr;          // device [N_total] (nullptr: none)
  float* alpha = nullptr;         // device [cin_v]   (Snake pre-activation)
  // fused DilatedUnit: this layer also carries the 1x1 conv that follows it
  bool fused = false;
  bool pairdu = false;            // C = 256 unit run by tc_du_pair_kernel (cluster pair, both convs in one launch)
  int fused_with = -1;            // op index of the 1x1 conv
  // op whose dst this launch writes: the fused 1x1 conv, or a HEAD_V2 op folded into the epilogue
  int tail = -1;
  int head = 0;                   // epilogue computes tanh(wave * sigmoid(amp)) (blocks.py amplitude_modulation)
  float* bias2 = nullptr;
  float* alpha2 = nullptr;
  int act2 = 0;
  // stream-as-row twin (Plan::sr_layers): geometry of the conv it was unrolled from
  int sr_taps = 0, sr_dil = 1, sr_off = 0;
};

struct Plan {
  std::vector<Layer> layers;      // per op
  // Streaming, deep layers (<= 8 rows per stream and chunk): the same conv with its taps unrolled into
  // the K dimension (im2col over taps, n_taps = 1), run with the STREAM index as the row axis of the
  // views (and the rows of the chunk as its stream axis) -- 128 streams fill a 128-row tile instead of running
  // one nearly empty tile (and one pass over the whole weight set) per stream.  See run_layer.
  std::vector<Layer> sr_layers;
  std::vector<int> skip;          // op is the tail of a fused pair
  std::vector<int> cp;            // per slot channel pitch
  // per slot: tensor is stored LeakyReLU-activated (every consumer wants act(x) as conv input, or
  // x as a residual, which min(z, 5z) recovers) -- consumers then load it straight into the MMA
  std::vector<int> sact;
  // Folded conditioning (offline): the conv right behind CONCAT_COND, k = 3 with centred zero padding, reads only
  // the latent channels; the time-constant speaker channels enter as a per-stream bias with the exact edge rows
  // (the first row of a stream misses tap 0, the last row tap 2).  fold_layers[i]: the conv packed over the latent
  // channels only; fold_wc[i]: its speaker-channel weights, bf16-rounded, fp32 [3][c_spk][N_total] on the device.
  std::vector<Layer> fold_layers;
  std::vector<float*> fold_wc;
  float* fold_sbias = nullptr;    // [rows][3][N_total], recomputed when the speaker table or the batch changes
  size_t fold_sbias_cap = 0;
  const float* fold_key_spk = nullptr;
  long long fold_key_stride = -1;
  int fold_key_b0 = -1, fold_key_B = -1, fold_key_op = -1;
  unsigned long long fold_key_version = ~0ull;
  cudaStream_t fold_key_stream = nullptr;
  int fold_now = -1;              // op whose folded form runs next (set by its CONCAT_COND, which then writes nothing)
  int run_oe = 0;                 // end of the op range of the running program (tc_run_program)
};
static thread_local const float* g_sbias = nullptr;   // per-stream bias of the NEXT run_layer call (folded conditioning)
static thread_local int g_sbias_n = 0;

static int slot_cp(const rb_handle* h, int sl) {
  if (sl == h->spec.enc_in_slot || sl == h->spec.dec_out_slot) return 1;   // audio [B][T]
  return round16(h->slots[sl].channels);
}

template <typename T>
static int dev_upload(const std::vector<T>& v, T** dev) {
  RB_CUDA(cudaMalloc((void**)dev, std::max<size_t>(v.size(), 1) * sizeof(T)));
  RB_CUDA(cudaMemcpy(*dev, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  return RB_OK;
}

static inline __nv_bfloat16 f2bf(float f) { return __float2bfloat16_rn(f); }

// Snake alpha of n channels, followed by 1 / (alpha + 1e-9) (blocks.py:845-853) so that the kernels multiply
static int upload_alpha(const std::vector<float>& a, float** dev) {
  std::vector<float> t(a);
  for (float v : a) t.push_back(1.0f / (v + 1e-9f));
  return dev_upload(t, dev);
}

// View-space weight of one op: Wv[n][c][j], n < N_total, c < cin_v, j < n_taps (zero padded).
struct ViewWeight {
  int N_total, cin_v, n_taps, dil, off, fin, fout, out_mask;
  std::vector<float> w;
  std::vector<float> bias;    // [N_total] (empty: none)
  std::vector<float> alpha;   // [cin_v]   (empty: none)
  float& at(int n, int c, int j) { return w[((size_t)n * cin_v + c) * n_taps + j]; }
};

// conv geometry -> view geometry.  x index of tap j at output t: t*stride + j*dil - off.
static void view_geometry(int k, int stride, int dil, int off, ViewWeight& v) {
  if (stride == 1) {
    v.n_taps = k; v.dil = dil; v.off = off;
  } else {
    const int jmin = floordiv(-off, stride), jmax = floordiv((k - 1) * dil - off, stride);
    v.n_taps = jmax - jmin + 1; v.dil = 1; v.off = -jmin;
  }
}

static int build_view_weight(const rb_handle* h, int oi, const Plan& pl, ViewWeight& v) {
  const rb_op& op = h->ops[oi];
  const OpWeights& ow = h->w[oi];
  const int cp_in = pl.cp[op.src];
  RB_CHECK(op.src_ch_off == 0, RB_ERR_UNSUPPORTED, "op %d: src_ch_off is not supported by the bf16 engine", oi);
  if (op.kind == RB_OP_CONV) {
    const int r = std::max(op.ileave, 1);
    const int c_out_ph = op.c_out / r;                 // channels per phase
    const int cpo = round16(c_out_ph);
    RB_CHECK(pl.cp[op.dst] == cpo, RB_ERR_INVALID, "op %d: output pitch mismatch", oi);
    v.fin = op.stride; v.fout = r; v.out_mask = 0;
    v.cin_v = op.stride * cp_in;
    v.N_total = r * cpo;
    view_geometry(op.k, op.stride, op.dilation, op.off, v);
    v.w.assign((size_t)v.N_total * v.cin_v * v.n_taps, 0.f);
    RB_CHECK(ow.w_host.size() == (size_t)op.c_out * op.c_in * op.k, RB_ERR_STATE, "op %d: no host weight", oi);
    const int jmin = op.stride == 1 ? 0 : floordiv(-op.off, op.stride);
    for (int co = 0; co < op.c_out; ++co) {
      const int ph = co / c_out_ph, cc = co - ph * c_out_ph;
      const int n = ph * cpo + cc;
      for (int ci = 0; ci < op.c_in; ++ci)
        for (int j = 0; j < op.k; ++j) {
          const float wv = ow.w_host[((size_t)co * op.c_in + ci) * op.k + j];
          if (op.stride == 1) {
            v.at(n, ci, j) = wv;
          } else {
            const int q = j * op.dilation - op.off;
            const int row = floordiv(q, op.stride), sub = q - row * op.stride;
            v.at(n, sub * cp_in + ci, row - jmin) = wv;
          }
        }
    }
    if (op.has_bias) {
      RB_CHECK((int)ow.bias_host.size() == c_out_ph, RB_ERR_STATE, "op %d: no host bias", oi);
      v.bias.assign(v.N_total, 0.f);
      for (int ph = 0; ph < r; ++ph)
        for (int cc = 0; cc < c_out_ph; ++cc) v.bias[ph * cpo + cc] = ow.bias_host[cc];
    }
    if (op.pre_act == RB_ACT_SNAKE) {
      RB_CHECK((int)ow.alpha_host.size() == op.c_in, RB_ERR_STATE, "op %d: Snake alpha not loaded", oi);
      v.alpha.assign(v.cin_v, 1.f);
      for (int sub = 0; sub < op.stride; ++sub)
        for (int ci = 0; ci < op.c_in; ++ci) v.alpha[sub * cp_in + ci] = ow.alpha_host[ci];
    }
    return RB_OK;
  }
  if (op.kind == RB_OP_PQMF_FWD) {
    // Conv1d(1 -> M, k, stride M) on the audio viewed as [T/M][M]; keep c_out bands; the
    // reverse_half sign flip (pqmf.py:13-17) is the epilogue's out_mask.
    const int M = op.n_band;
    RB_CHECK(ow.w_host.size() == (size_t)M * op.k, RB_ERR_STATE, "pqmf analysis weight missing");
    v.fin = M; v.fout = 1; v.out_mask = 1;
    v.cin_v = M * cp_in;                       // cp_in == 1
    v.N_total = round16(op.c_out);
    view_geometry(op.k, M, 1, op.off, v);
    v.w.assign((size_t)v.N_total * v.cin_v * v.n_taps, 0.f);
    const int jmin = floordiv(-op.off, M);
    for (int m = 0; m < op.c_out; ++m)
      for (int j = 0; j < op.k; ++j) {
        const int q = j - op.off, row = floordiv(q, M), sub = q - row * M;
        v.at(m, sub, row - jmin) = ow.w_host[(size_t)m * op.k + j];
      }
    return RB_OK;
  }
  if (op.kind == RB_OP_PQMF_INV) {
    // out[t*M + (M-1-m)] = M * sum_c sum_j w[m][c][j] xr[c][t + j - off]  (pqmf.py:275-284);
    // xr = reverse_half(x) is written by the head kernels (the only producers of this slot).
    const int M = op.n_band;
    RB_CHECK(ow.w_host.size() == (size_t)M * M * op.k, RB_ERR_STATE, "pqmf synthesis weight missing");
    RB_CHECK(cp_in == M, RB_ERR_UNSUPPORTED, "pqmf synthesis expects %d-channel input", M);
    v.fin = 1; v.fout = M; v.out_mask = 0;
    v.cin_v = cp_in;
    v.N_total = M;
    view_geometry(op.k, 1, 1, op.off, v);
    v.w.assign((size_t)v.N_total * v.cin_v * v.n_taps, 0.f);
    for (int m = 0; m < M; ++m)
      for (int c = 0; c < M; ++c)
        for (int j = 0; j < op.k; ++j)
          v.at(M - 1 - m, c, j) = (float)M * ow.w_host[((size_t)m * M + c) * op.k + j];
    return RB_OK;
  }
  set_error("op %d is not a convolution", oi);
  return RB_ERR_INVALID;
}

// Row batching: F consecutive rows become ONE row of F x the channels on both sides,
//   D'[t', s*N + n] = sum_j sum_c A[F t' + s + j*dil - off, c] W[n, c, j],   s < F,
// i.e. a conv over rows of F*cin channels with ceil((F + halo) / F) taps.  Same bytes in memory
// (view factors fin*F / fout*F); narrow layers (PQMF banks, first / last convs) get an MMA N
// that is F times wider, F times fewer tiles and output chunks for every epilogue warp.  The
// zero (tap, sub-row) blocks this creates are skipped by the packer's tap masks.
static void batch_rows(ViewWeight& v, int F) {
  ViewWeight o;
  o.fin = v.fin * F; o.fout = v.fout * F; o.cin_v = v.cin_v * F; o.N_total = v.N_total * F;
  const int amin = floordiv(-v.off, F), amax = floordiv(F - 1 + (v.n_taps - 1) * v.dil - v.off, F);
  o.n_taps = amax - amin + 1; o.dil = 1; o.off = -amin; o.out_mask = 0;
  o.w.assign((size_t)o.N_total * o.cin_v * o.n_taps, 0.f);
  for (int s = 0; s < F; ++s)
    for (int n = 0; n < v.N_total; ++n) {
      // reverse_half (out_mask): odd channels of even rows change sign; F is even, so row parity = s parity
      const float sign = (v.out_mask && !(s & 1) && (n & 1)) ? -1.f : 1.f;
      for (int j = 0; j < v.n_taps; ++j) {
        const int m = s + j * v.dil - v.off, a = floordiv(m, F), b = m - a * F;
        for (int c = 0; c < v.cin_v; ++c)
          o.at(s * v.N_total + n, b * v.cin_v + c, a - amin) = sign * v.at(n, c, j);
      }
    }
  if (!v.bias.empty()) {
    o.bias.resize(o.N_total);
    for (int s = 0; s < F; ++s)
      for (int n = 0; n < v.N_total; ++n)
        o.bias[s * v.N_total + n] = ((v.out_mask && !(s & 1) && (n & 1)) ? -1.f : 1.f) * v.bias[n];
  }
  if (!v.alpha.empty()) {
    o.alpha.resize(o.cin_v);
    for (int b = 0; b < F; ++b)
      for (int c = 0; c < v.cin_v; ++c) o.alpha[b * v.cin_v + c] = v.alpha[c];
  }
  v = std::move(o);
}

// Batching factor for a narrow layer: widest N <= 128 whose weights stay resident in smem.
static int choose_row_batch(const ViewWeight& v, int rows_per_frame) {
  if (v.N_total > 64) return 1;
  for (int F = 4; F >= 2; F >>= 1) {
    if (rows_per_frame % F) continue;
    if (v.N_total * F > 128) continue;
    const int taps = floordiv(F - 1 + (v.n_taps - 1) * v.dil - v.off, F) - floordiv(-v.off, F) + 1;
    // non-zero (tap, sub-row) blocks: one per distinct row offset s + j*dil - off
    const size_t bytes = (size_t)(v.N_total * F) * v.cin_v * (size_t)(F + (v.n_taps - 1) * v.dil) * 2;
    if (taps > 64 || bytes > (size_t)112 * 1024) continue;
    return F;
  }
  return 1;
}

static int largest_div16(int v, int cap) {
  int best = 16;
  for (int d = 16; d <= cap && d <= v; d += 16)
    if (v % d == 0) best = d;
  return best;
}

static void free_layer(Layer& L) {
  cudaFree(L.w); cudaFree(L.w2); cudaFree(L.masks); cudaFree(L.bias); cudaFree(L.alpha); cudaFree(L.bias2); cudaFree(L.alpha2);
  L = Layer();
}

// Pack Wv into [N_TILE][KC] blocks (UMMA K-major no-swizzle core-matrix order) in the order
// the kernel consumes them: n-tile, K chunk, active tap.  `extra` (fused 1x1 conv) appended.
static int pack_layer(const ViewWeight& v, int kc_cap, const ViewWeight* extra, Layer& L, int ncap_arg = 0,
                      bool twin = false) {
  L.fin = v.fin; L.fout = v.fout; L.cin_v = v.cin_v; L.n_taps = v.n_taps; L.dil = v.dil; L.off = v.off;
  L.N_total = v.N_total; L.out_mask = v.out_mask;
  RB_CHECK(v.n_taps <= 64, RB_ERR_UNSUPPORTED, "conv with %d row taps (max 64)", v.n_taps);
  int ncap = 256;
  if (const char* e = getenv("RB_TC_NTILE_CAP")) ncap = std::max(16, atoi(e));   // experiments only
  if (const char* e = getenv("RB_TC_NTILE_CAP_IL")) {                             // experiments only
    const int fo = atoi(e) / 1000, cap = atoi(e) % 1000;     // "4128": layers with fout >= 4 get cap 128
    if (v.fout >= fo) ncap = cap;
  }
  if (ncap_arg > 0) ncap = ncap_arg;
  L.N_TILE = largest_div16(v.N_total, ncap);
  L.n_ntiles = v.N_total / L.N_TILE;
  RB_CHECK(L.n_ntiles <= TC_MAX_NT, RB_ERR_UNSUPPORTED, "too many output-channel tiles");
  L.KC = largest_div16(v.cin_v, kc_cap);
  L.n_kchunks = v.cin_v / L.KC;
  L.blk_bytes = (uint32_t)L.N_TILE * L.KC * 2;
  std::vector<uint64_t> masks((size_t)L.n_ntiles * L.n_kchunks, 0);
  for (int nt = 0; nt < L.n_ntiles; ++nt) {
    bool any = false;
    for (int kc = 0; kc < L.n_kchunks; ++kc) {
      uint64_t m = 0;
      for (int j = 0; j < v.n_taps; ++j) {
        bool nz = false;
        for (int n = nt * L.N_TILE; n < (nt + 1) * L.N_TILE && !nz; ++n)
          for (int c = kc * L.KC; c < (kc + 1) * L.KC; ++c)
            if (v.w[((size_t)n * v.cin_v + c) * v.n_taps + j] != 0.f) { nz = true; break; }
        if (nz) m |= (uint64_t)1 << j;
      }
      if (extra && m == 0) m = 1;                      // fused unit: every chunk is loaded (E2 tracks the A ring)
      masks[(size_t)nt * L.n_kchunks + kc] = m;
      any |= m != 0;
    }
    if (!any) masks[(size_t)nt * L.n_kchunks] = 1;   // keep the accumulator defined
  }
  L.nt_base.assign(L.n_ntiles + 1, 0);
  std::vector<__nv_bfloat16> blocks;
  int nblk = 0;
  for (int nt = 0; nt < L.n_ntiles; ++nt) {
    L.nt_base[nt] = nblk;
    for (int kc = 0; kc < L.n_kchunks; ++kc)
      for (int j = 0; j < v.n_taps; ++j) {
        if (!((masks[(size_t)nt * L.n_kchunks + kc] >> j) & 1)) continue;
        const size_t base = blocks.size();
        blocks.resize(base + (size_t)L.N_TILE * L.KC);
        for (int n = 0; n < L.N_TILE; ++n)
          for (int c = 0; c < L.KC; ++c)
            blocks[base + (size_t)(c >> 3) * (L.N_TILE * 8) + (size_t)n * 8 + (c & 7)] =
                f2bf(v.w[((size_t)(nt * L.N_TILE + n) * v.cin_v + kc * L.KC + c) * v.n_taps + j]);
        ++nblk;
      }
  }
  L.nt_base[L.n_ntiles] = nblk;
  L.total_blocks = nblk;
  if (extra) {
    // 1x1 conv of the fused unit: K = N_TILE channels in chunks of KC, same block shape
    for (int kc = 0; kc < L.N_TILE / L.KC; ++kc) {
      const size_t base = blocks.size();
      blocks.resize(base + (size_t)L.N_TILE * L.KC);
      for (int n = 0; n < L.N_TILE; ++n)
        for (int c = 0; c < L.KC; ++c)
          blocks[base + (size_t)(c >> 3) * (L.N_TILE * 8) + (size_t)n * 8 + (c & 7)] =
              f2bf(extra->w[((size_t)n * extra->cin_v + kc * L.KC + c) * extra->n_taps]);
    }
  }
  int rc = dev_upload(blocks, &L.w);
  if (rc) return rc;
  if (L.N_TILE == 256 || twin) {
    // N-split twin for cta_group::2: block [n][c] -> half n / (N / 2), then the same core-matrix order over the
    // N / 2 rows of a half (a pair DilatedUnit's 1x1 conv blocks included)
    std::vector<__nv_bfloat16> b2(blocks.size());
    const size_t be = (size_t)L.N_TILE * L.KC, he = be / 2;
    const int nb_all = (int)(blocks.size() / be), H = L.N_TILE / 2;
    for (int b = 0; b < nb_all; ++b)
      for (int n = 0; n < L.N_TILE; ++n)
        for (int c = 0; c < L.KC; ++c)
          b2[b * be + (size_t)(n / H) * he + (size_t)(c >> 3) * (H * 8) + (size_t)(n % H) * 8 + (c & 7)] =
              blocks[b * be + (size_t)(c >> 3) * (L.N_TILE * 8) + (size_t)n * 8 + (c & 7)];
    if ((rc = dev_upload(b2, &L.w2))) return rc;
  }
  if ((rc = dev_upload(masks, &L.masks))) return rc;
  if (!v.bias.empty() && (rc = dev_upload(v.bias, &L.bias))) return rc;
  if (!v.alpha.empty() && (rc = upload_alpha(v.alpha, &L.alpha))) return rc;
  L.valid = true;
  return RB_OK;
}

// ------------------------------------------------------------------ buffers and launches
// One activation tensor of the bf16 engine, in un-reshaped row units.  Layout: group-major
// [stream][cp/8][ld rows][8 channels] (cp = channels rounded up to 16); audio (cp == 1) is the
// caller's dense [stream][ld samples].
struct Buf {
  __nv_bfloat16* p = nullptr;
  int cp = 0, ld = 0, hist = 0, n = 0, lo = 0;
};

static int make_view(const Buf& b, int f, TcView* v) {
  RB_CHECK(b.ld % f == 0 && b.hist % f == 0 && b.n % f == 0, RB_ERR_INVALID,
           "tensor rows (%d,%d,%d) not divisible by the view factor %d", b.ld, b.hist, b.n, f);
  v->p = b.p; v->hist = b.hist / f; v->n = b.n / f; v->lo = floordiv(b.lo, f);
  if (b.cp == 1) {                              // audio: rows of f samples, f/8 groups per row
    RB_CHECK(f % 8 == 0 && b.ld % 8 == 0, RB_ERR_INVALID, "audio view needs a factor that is a multiple of 8");
    v->C = f; v->G1 = 1; v->sA = 0; v->sB = 1; v->sT = f / 8; v->sb = b.ld / 8;   // group g of a row = unit g
  } else {
    const int G = b.cp / 8;
    v->C = b.cp * f; v->G1 = G; v->sA = b.ld; v->sB = 1; v->sT = f; v->sb = (long long)G * b.ld;
  }
  v->g1_shift = -1;
  for (int sh = 0; sh < 20; ++sh)
    if ((1 << sh) == v->G1) v->g1_shift = sh;
  return RB_OK;
}

static uint32_t align_up(uint32_t v, uint32_t a) { return (v + a - 1) / a * a; }

static unsigned long long* g_trace = nullptr;   // debug hook only: device buffer [64][16] of clock stamps
// Offline programs alternate the direction in which consecutive launches walk their row tiles (TcParams::reverse):
// a layer then starts on the rows its producer wrote last, which are still in L2.  -1: off (streaming, debug hook).
static thread_local int g_reverse = -1;
static inline int next_direction() {
  if (g_reverse < 0) return 0;
  const int r = g_reverse;
  g_reverse ^= 1;
  return r;
}

// x_act / res_act / out_act: the tensor is (to be) stored LeakyReLU-activated (Plan::sact).
// sr_rows > 0: L is a stream-as-row twin (Plan::sr_layers); computes the sr_rows rows of the chunk for all
// streams in one launch (rows of the views = streams, "streams" of the views = rows of the chunk).
static int run_layer(const Layer& L, const rb_op& op, const rb_op* op2, const Buf& src, const Buf& dst,
                     const Buf* res, int B, cudaStream_t s, int x_act = 0, int res_act = 0, int out_act = 0,
                     int sr_rows = 0) {
  TcParams p;
  memset(&p, 0, sizeof(p));
  int rc = make_view(src, L.fin, &p.x);
  if (rc) return rc;
  if ((rc = make_view(dst, L.fout, &p.y))) return rc;
  p.sbias = g_sbias; p.sbias_n = g_sbias_n;
  g_sbias = nullptr;
  const int n_streams = B;
  if (sr_rows > 0) {
    // rows = streams: row stride = the per-stream stride; tap j (and sub-row of a strided view) becomes
    // the phase g / G1 of the view, `sB` units apart; the first tap's row is folded into the pointer;
    // the chunk's rows take the place of the stream index (stride = one view row)
    TcView& x = p.x;
    RB_CHECK(src.cp != 1 && dst.cp != 1 && L.sr_taps > 0, RB_ERR_INVALID, "stream-as-row layer on an audio slot");
    RB_CHECK(x.hist - L.sr_off >= 0, RB_ERR_INVALID, "stream-as-row: history shorter than the filter");
    x.p = src.p + (long long)(x.hist - L.sr_off) * x.sT * 8;
    x.sB = L.fin == 1 ? L.sr_dil : 1;
    { const int row_units = x.sT; x.sT = (int)x.sb; x.sb = row_units; }
    x.hist = 0; x.lo = 0; x.n = n_streams; x.C = L.cin_v;
    TcView& y = p.y;
    y.p = dst.p + (long long)y.hist * y.sT * 8;
    { const int row_units = y.sT; y.sT = (int)y.sb; y.sb = row_units; }
    y.hist = 0; y.lo = 0; y.n = n_streams;
    B = sr_rows;
  }
  RB_CHECK(p.x.C == L.cin_v && p.y.C == (L.head ? L.N_total / 2 : L.N_total), RB_ERR_INVALID,
           "layer/view channel mismatch (%d vs %d, %d vs %d)", p.x.C, L.cin_v, p.y.C, L.N_total);
  p.head = L.head;
  const rb_op& rop = op2 ? *op2 : op;
  if (res) {
    if ((rc = make_view(*res, 1, &p.res))) return rc;
    RB_CHECK(L.fout == 1 && p.res.C == L.N_total, RB_ERR_INVALID, "residual shape mismatch");
    p.res_delay = rop.res_delay;
    if (sr_rows > 0) {
      TcView& r = p.res;
      RB_CHECK(res->cp != 1 && r.hist - rop.res_delay >= 0, RB_ERR_INVALID, "stream-as-row: residual history");
      r.p = res->p + (long long)(r.hist - rop.res_delay) * 8;
      { const int row_units = r.sT; r.sT = (int)r.sb; r.sb = row_units; }
      r.hist = 0; r.lo = 0; r.n = n_streams;
      p.res_delay = 0;
    }
  }
  p.w = L.w; p.masks = L.masks; p.bias = L.bias; p.alpha = L.alpha; p.alpha_n = L.cin_v;
  p.pre_act = x_act ? RB_ACT_NONE : op.pre_act;
  p.out_act = out_act ? RB_ACT_LEAKY : RB_ACT_NONE;
  p.res_act = (res && res_act) ? 1 : 0;
  // fused unit whose residual is its own input: E2 reads it back from the A tile in smem
  p.res_smem = (L.fused && res && res->p == src.p && rop.res_delay == 0 && L.n_kchunks == 1 &&
                p.pre_act == RB_ACT_NONE && !getenv("RB_TC_NORESSMEM")) ? 1 : 0;   // (an activation pass would overwrite the tile)
  p.B = B;
  p.n_taps = L.n_taps; p.dil = L.dil; p.off = L.off;
  p.KC = L.KC; p.n_kchunks = L.n_kchunks;
  p.N_TILE = L.N_TILE; p.n_ntiles = L.n_ntiles;
  for (int i = 0; i <= L.n_ntiles; ++i) p.nt_base[i] = L.nt_base[i];
  p.out_mask = L.out_mask;
  p.trace = g_trace;
  p.fused = L.fused ? 1 : 0;
  p.act2 = L.act2; p.bias2 = L.bias2; p.alpha2 = L.alpha2; p.n_blk1 = L.total_blocks;
  const int halo = (L.n_taps - 1) * L.dil;
  const uint32_t budget = 224u * 1024u;
  const uint32_t w_total = (uint32_t)(L.total_blocks + (L.fused ? L.N_TILE / L.KC : 0)) * L.blk_bytes;
  p.w_blk_bytes = L.blk_bytes;
  p.off_bias = 640 + TC_MAX_MASKS * 8;    // after the mbarriers, the TMEM address slot and the tap masks
  RB_CHECK(L.n_ntiles * L.n_kchunks <= TC_MAX_MASKS, RB_ERR_UNSUPPORTED, "too many (n-tile, K-chunk) pairs (%d x %d)",
           L.n_ntiles, L.n_kchunks);
  const uint32_t fixed = align_up(p.off_bias + (uint32_t)(L.N_total + 3 * L.N_TILE) * 4, 128);   // bias, bias2, alpha2, 1/alpha2
  // accumulator stages: as many as TMEM holds (<= 4); the fused unit keeps one h tile per stage
  const int acc_cols1 = L.fused ? 2 * L.N_TILE : L.N_TILE;
  const int sd_max = std::max(1, std::min(4, 512 / acc_cols1));
  p.SH = L.fused ? std::min(sd_max, p.res_smem ? 3 : 4) : 0;
  // fused unit: h = act(conv3 + b) goes back into the accumulator's TMEM columns (packed bf16) and the 1x1 conv
  // takes its A operand from there: no h tile in shared memory (its stores and the MMA's reads of it were a
  // sixth of the L1 data-pipe wavefronts of a C = 128 unit) and no single h stage to queue behind.  Measured:
  // C = 128 units 545 -> 531 kcycles; C = 64 units 790 -> 815 (issue-bound: the TMEM store + wait cost more
  // instructions than the two shared-memory stores they replace), so only units wider than 64 channels use it.
  // (one kernel variant carries it: LeakyReLU units on activated storage with the residual in global memory)
  p.h_tmem = (L.fused && L.N_TILE > 64 && (L.N_TILE & 31) == 0 && p.pre_act == RB_ACT_NONE && L.act2 == RB_ACT_LEAKY &&
              res && !p.res_smem && !getenv("RB_TC_NOHTMEM")) ? 1 : 0;
  p.h_stage_bytes = (L.fused && !p.h_tmem) ? (uint32_t)(L.N_TILE / 8) * 2048u : 0u;
  // resident weights when the whole layer fits beside >= 2 A stages, else a streamed ring
  for (int attempt = 0; attempt < 2; ++attempt) {
    const bool resident = attempt == 0 && (L.n_ntiles == 1) && w_total <= 160u * 1024u;
    if (attempt == 0 && !resident) continue;
    RB_CHECK(resident || !L.fused, RB_ERR_UNSUPPORTED, "fused unit does not fit in shared memory");
    p.w_resident = resident ? 1 : 0;
    {
      // Two warps issue GEMM1 of alternate work items where one tile's MMAs are short: the issuing warp shares its
      // scheduler with four busy output warps and needs ~140 cycles per tcgen05.mma (event trace of the C = 128
      // unit: 3300 cycles to issue 24 MMAs that execute in 1536), so a single issuer sets the tile rate.
      // N_TILE <= 64 always; N_TILE = 128 when a tile is at most 2048 tensor-pipe cycles (first conv 527 -> 425
      // kcycles, C = 128 units 455 -> 446; the 64->128 strided conv, 3072 cycles of MMAs per tile, is left alone)
      const int mma_clk = L.total_blocks * (L.KC / 16) * (L.N_TILE / 2);
      int dual_max = 128;
      if (const char* e = getenv("RB_TC_DUAL_MAXN")) dual_max = atoi(e);   // experiments only
      p.dual_g1 = (resident && L.N_TILE <= dual_max && (L.N_TILE <= 64 || mma_clk <= 2048) && !getenv("RB_TC_NODUAL")) ? 2 : 0;
      // (a third issuer -- the second-GEMM warp, idle in unfused kernels -- measured: PQMF banks 249 -> 247, last
      // conv 587 -> 582 kcycles: two are enough.  RB_TC_TRI=1 turns it on for experiments.)
      if (p.dual_g1 && !L.fused && L.N_TILE <= 64 && getenv("RB_TC_TRI")) p.dual_g1 = 3;
    }
    p.MT = (!resident && p.y.n >= 256 && sr_rows == 0) ? 2 : 1;   // stream-as-row: few row tiles, keep them all busy
    // streamed transposed convs run as cluster pairs: the two CTAs split every weight block (half
    // the L2 -> SM weight traffic each), which pays for one sub-tile per accumulator stage and so
    // for two stages at N_TILE = 256 -- the MMAs of the next work item overlap the epilogue, and
    // these layers (r phases of output per input row) are epilogue-heavy.  Measured neutral or
    // worse on the plain k = 1 / k = 3 layers (the 4-deep weight ring is then consumed in ~2000
    // cycles, about its refill latency), so those keep two sub-tiles per weight block.
    p.pair = (!resident && !L.fused && p.pre_act == RB_ACT_NONE && p.y.n >= 256 && L.fout >= 2 && sr_rows == 0 &&
              !getenv("RB_TC_NOPAIR")) ? 1 : 0;
    if (const char* e = getenv("RB_TC_PAIRALL"))           // experiments only
      p.pair = (atoi(e) && !resident && !L.fused && p.pre_act == RB_ACT_NONE && p.y.n >= 256) ? 1 : 0;
    // cta_group::2 (tc_conv.cu, CG2): every streamed plain layer with 256-wide output tiles whose pair gets two
    // real row tiles.  One 128-row tile per CTA and accumulator stage (two stages of 256 columns: the MMAs of the
    // next item run under the epilogue), each CTA loads half of every weight block.
    p.cg2 = (!resident && !L.fused && L.w2 && L.N_TILE == 256 && p.pre_act == RB_ACT_NONE && p.y.n >= 256 &&
             sr_rows == 0 && !g_trace && !getenv("RB_TC_NOCG2")) ? 1 : 0;
    if (p.cg2) {
      p.cg2_spin = 0;
      if (const char* e = getenv("RB_TC_CG2_SPIN")) p.cg2_spin = atoi(e);   // experiments only
      p.pair = 1;
      // two sub-tiles per weight block (one accumulator stage of 512 columns) where a tile is long enough
      // (>= 8 K chunks x taps) that the un-overlapped epilogue is the smaller cost than streaming every weight
      // block twice as often; else one sub-tile and two accumulator stages (measured per layer,
      // profiles/r02_cg2_cycles.txt)
      // (single-tap layers too since the accumulator is split: conv1 + res C = 512 321 -> 306 kcycles)
      p.MT = (p.y.n >= 512 && L.n_kchunks >= 8) ? 2 : 1;
      if (const char* e = getenv("RB_TC_CG2_MT")) p.MT = (atoi(e) == 2 && p.y.n >= 512) ? 2 : 1;   // experiments only
      // short tensors (the deepest layers: 256 rows per stream) too: the units of a cluster are numbered across
      // streams (flat_units), so each CTA still runs two sub-tiles per weight block -- these layers stream their
      // whole weight set from L2 once per 128 rows otherwise, and that traffic is what bounds them
      if (p.MT == 1 && p.y.n >= 256 && p.y.n < 512 && B >= 2 && L.n_taps >= 2 && L.n_kchunks >= 8 &&
          !getenv("RB_TC_NOPAIRSTREAMS"))
        p.MT = 2;
    }
    if (p.pair && !p.cg2) p.MT = (2 * L.N_TILE <= 256 && p.y.n >= 512) ? 2 : 1;
    p.R = p.MT * 128 + halo;
    {
      // TMA boxes of the A tile: whole chunk in one box when the rows fit a box (<= 256), else
      // one box per channel group and row half
      const int c8n = L.KC / 8;
      const int groups = p.x.C / 8;
      const int g1 = groups <= p.x.G1 ? groups : p.x.G1;
      const int r8 = (p.R + 7) & ~7;
      if (r8 <= 256) {
        p.R_ALLOC = r8; p.a_rsplit = 1; p.a_box_rows = r8;
        int gb = c8n;
        while (g1 % gb) --gb;
        while (c8n % gb) --gb;
        p.a_gb = gb;
      } else {
        p.R_ALLOC = (p.R + 15) & ~15; p.a_rsplit = 2; p.a_box_rows = p.R_ALLOC / 2; p.a_gb = 1;
      }
      RB_CHECK(p.a_box_rows <= 256, RB_ERR_UNSUPPORTED, "A tile of %d rows exceeds two TMA boxes", p.R);
      p.a_tx_bytes = (uint32_t)c8n * p.R_ALLOC * 16u;
      {
        int flat_min = 0;
        if (const char* e = getenv("RB_TC_FLATMIN")) flat_min = atoi(e);   // experiments only
        p.a_flat = (p.x.sT == 1 && p.x.C / 8 <= p.x.G1 && p.R_ALLOC >= flat_min) ? 1 : 0;
      }
    }
    p.a_stage_bytes = (uint32_t)(L.KC / 8) * p.R_ALLOC * 16u;
    // W ring depth: a single-tap layer consumes one A chunk per W block, so it needs A stages in
    // flight (HBM latency) more than W stages (L2 latency); multi-tap layers the other way round
    p.SW = resident ? 0 : (L.n_taps == 1 ? 2 : 4);
    const uint32_t blk_bytes = p.cg2 ? L.blk_bytes / 2 : L.blk_bytes;      // one ring stage
    if (p.cg2) { p.SW = 4; p.w = L.w2; p.w_blk_bytes = blk_bytes; }
    if (!resident && sr_rows == 0)
      if (const char* e = getenv("RB_TC_SW")) p.SW = std::min(8, std::max(2, atoi(e)));   // experiments only
    if (!resident && sr_rows > 0) {
      // stream-as-row: a CTA's work is one short pass over its slice of the weights, bound by the L2
      // round trip of each block -- keep as many blocks in flight as fit beside three A stages
      int sw = 6;                                                               // 4 -> 6: C3 1.086 -> 1.051 ms / step
      if (const char* e = getenv("RB_TC_SR_SW")) sw = std::min(8, std::max(2, atoi(e)));   // experiments only (<= MAX_SW)
      p.SW = sw;
      while (p.SW > 2 && fixed + p.SW * L.blk_bytes + 3 * p.a_stage_bytes > budget) --p.SW;
    }
    uint32_t w_bytes = resident ? align_up(w_total, 128) : p.SW * blk_bytes;
    if (!resident && fixed + w_bytes + 2 * p.a_stage_bytes > budget) { p.SW = 2; w_bytes = p.SW * blk_bytes; }
    if (L.fused) {
      // fewer h stages when smem is tight (wide units): keep at least 3 A stages if possible
      while (p.SH > 1 && fixed + w_bytes + p.SH * p.h_stage_bytes + 3 * p.a_stage_bytes > budget) --p.SH;
    }
    const uint32_t h_total = p.SH * p.h_stage_bytes;
    if (fixed + w_bytes + h_total + p.a_stage_bytes > budget) {
      RB_CHECK(resident, RB_ERR_UNSUPPORTED, "conv tile does not fit in shared memory");
      continue;
    }
    const uint32_t room = budget - fixed - w_bytes - h_total;
    p.SA = (int)std::min<uint32_t>(12, room / p.a_stage_bytes);
    if (resident && p.SA < 2) continue;
    p.off_a = fixed;
    p.off_w = p.off_a + p.SA * p.a_stage_bytes;
    p.off_h = p.off_w + w_bytes;
    p.smem_bytes = p.off_h + h_total;
    break;
  }
  RB_CHECK(p.SA >= 1, RB_ERR_UNSUPPORTED, "conv tile does not fit in shared memory");
  // Several GEMM1 issuers run up to one work item apart.  An issuer may wait for fill k + 1 of an A stage only
  // when fill k is certainly complete (an mbarrier parity wait passes on a barrier that is TWO phases behind):
  // true when the stage's previous fill belongs to the issuer's own previous item, i.e. the ring holds the
  // chunks of as many items as there are issuers.  (The stream-as-row twin of a C = 128 unit has 6 chunks per
  // item and 5 stages: with two issuers its second launch read stages that had not been filled.)
  if (p.dual_g1 > 1 && p.SA < p.dual_g1 * L.n_kchunks) p.dual_g1 = 0;
  const int acc_cols = L.fused ? 2 * L.N_TILE : p.MT * L.N_TILE;
  p.SD = std::max(1, std::min(4, 512 / acc_cols));
  if (L.fused) p.SD = std::min(sd_max, 4);
  if (p.dual_g1 > p.SD) p.dual_g1 = p.SD >= 2 ? p.SD : 0;   // same rule for the accumulator stages
  if (p.h_tmem) p.SH = p.SD;
  if (p.res_smem && p.SA < 5) p.res_smem = 0;           // too few A stages to hold tiles until E2
  RB_CHECK(!L.fused || p.SD >= 2, RB_ERR_UNSUPPORTED, "fused unit needs two accumulator stages");
  uint32_t cols = 32;
  while (cols < (uint32_t)(p.SD * acc_cols)) cols <<= 1;
  p.tmem_cols = cols;
  // sector-complete stores (tc_conv.cu drain_pair): rows that interleave in memory leave in 32-byte pieces
  p.pair_mode = 0;
  p.pair_ph_tile = 0;
  if (!L.fused && !res && sr_rows == 0 && p.pre_act == RB_ACT_NONE && !L.out_mask &&
      (p.y.sT & 1) == 0 && (p.y.sb & 1) == 0 && ((p.y.hist * p.y.sT) & 1) == 0 &&
      (reinterpret_cast<uintptr_t>(p.y.p) & 31) == 0 && !getenv("RB_TC_NOPAIRST")) {
    if (L.head) {
      if (L.n_ntiles == 1 && L.head == 2 && L.N_total == 64 && p.MT == 1 && (p.y.sA & 1) == 0) p.pair_mode = 3;
    } else if (dst.cp == 1) {
      if (L.n_ntiles == 1 && (L.N_total & 31) == 0) p.pair_mode = 2;
    } else if (L.fout >= 2 && (L.fout & 1) == 0 && p.y.sB == 1 && (p.y.sA & 1) == 0) {
      // several n-tiles: each must hold whole pairs of phases (the 256->128 x4 transposed conv: 4 phases of 128
      // channels in two tiles of 256 columns)
      const int dcol = L.N_total / L.fout;
      if (L.n_ntiles == 1 || (dcol % 16 == 0 && L.N_TILE % (2 * dcol) == 0 && L.N_TILE * L.n_ntiles == L.N_total)) {
        p.pair_mode = 1;
        p.pair_dcol = dcol;
        p.pair_units = (L.N_TILE / (2 * dcol)) * (dcol >> 4);
        p.pair_ph_tile = L.n_ntiles > 1 ? L.N_TILE / dcol : 0;
      }
    }
  }
  // MT = 2 leaves one accumulator stage: publish / take back its sub-tiles separately (tc_conv.cu gemm1_cg2)
  p.split = (p.cg2 && p.MT == 2 && p.SD == 1 && !p.pair_mode && L.n_taps <= p.SW && !g_trace &&
             !getenv("RB_TC_NOSPLIT")) ? 1 : 0;
  p.tiles_per_stream = (p.y.n + p.MT * 128 - 1) / (p.MT * 128);
  if (p.cg2) {
    p.flat_units = p.tiles_per_stream;
    const long long units = (long long)B * p.flat_units, per = 1ll << p.pair;
    p.n_work = L.n_ntiles * (int)((units + per - 1) / per);
  } else {
    if (p.pair) p.tiles_per_stream = (p.tiles_per_stream + 1) / 2;   // work slots: two row tiles each
    p.n_work = L.n_ntiles * B * p.tiles_per_stream;
  }
  if (getenv("RB_TC_PLAN_DEBUG"))
    fprintf(stderr, "tc layer: cin_v %d N %d x %d taps %d dil %d fin %d fout %d KC %d | resident %d fused %d pair %d cg2 %d MT %d "
            "SA %d SW %d SD %d pair_mode %d split %d rows %d work %d smem %u\n", L.cin_v, L.N_TILE, L.n_ntiles, L.n_taps, L.dil, L.fin,
            L.fout, L.KC, p.w_resident, p.fused, p.pair, p.cg2, p.MT, p.SA, p.SW, p.SD, p.pair_mode, p.split, p.y.n, p.n_work,
            p.smem_bytes);
  p.reverse = sr_rows == 0 ? next_direction() : 0;
  return launch_tc_conv(p, s);
}

// Residual(DilatedUnit) at C = 256 through tc_du_pair_kernel: x (activated storage) -> y, residual = x.
static int run_pair_du(const Layer& L, const rb_op& op, const Buf& src, const Buf& dst, int B, cudaStream_t s,
                       int res_act, int out_act) {
  TcParams p;
  memset(&p, 0, sizeof(p));
  int rc = make_view(src, 1, &p.x);
  if (rc) return rc;
  if ((rc = make_view(dst, 1, &p.y))) return rc;
  if ((rc = make_view(src, 1, &p.res))) return rc;
  RB_CHECK(p.x.C == L.N_TILE && p.y.C == L.N_TILE, RB_ERR_INVALID, "pair DilatedUnit: channel mismatch");
  p.w = L.w2; p.masks = L.masks; p.bias = L.bias; p.bias2 = L.bias2;
  p.pre_act = RB_ACT_NONE; p.act2 = RB_ACT_LEAKY;
  p.out_act = out_act ? RB_ACT_LEAKY : RB_ACT_NONE;
  p.res_act = res_act ? 1 : 0;
  p.B = B;
  p.n_taps = L.n_taps; p.dil = L.dil; p.off = L.off;
  p.KC = L.KC; p.n_kchunks = L.n_kchunks;
  p.N_TILE = L.N_TILE; p.n_ntiles = 1;
  p.nt_base[0] = 0; p.nt_base[1] = L.total_blocks;
  p.MT = 1; p.pair = 1; p.cg2 = 1; p.fused = 1; p.h_tmem = 1; p.n_blk1 = L.total_blocks;
  p.w_blk_bytes = L.blk_bytes / 2;
  const int halo = (L.n_taps - 1) * L.dil;
  p.R = 128 + halo;
  {
    const int c8n = L.KC / 8;
    const int r8 = (p.R + 7) & ~7;
    RB_CHECK(r8 <= 256, RB_ERR_UNSUPPORTED, "pair DilatedUnit: dilation too large");
    p.R_ALLOC = r8; p.a_rsplit = 1; p.a_box_rows = r8;
    int gb = c8n;
    const int g1 = p.x.G1;
    while (g1 % gb) --gb;
    while (c8n % gb) --gb;
    p.a_gb = gb;
    p.a_tx_bytes = (uint32_t)c8n * p.R_ALLOC * 16u;
    p.a_flat = (p.x.sT == 1) ? 1 : 0;
  }
  p.a_stage_bytes = (uint32_t)(L.KC / 8) * p.R_ALLOC * 16u;
  const uint32_t budget = 224u * 1024u;
  p.off_bias = 640 + TC_MAX_MASKS * 8;
  const uint32_t fixed = align_up(p.off_bias + 512u * 4u, 128);
  // shared memory: the 1x1 conv's half blocks resident; conv3's too when they fit (C <= 128: 48 KB per CTA at
  // C = 128), else a ring of them deep enough to hold the first half of the next tile's weights while E1 / GEMM2
  // of this tile run; and >= 3 A stages
  const uint32_t w1_bytes = (uint32_t)L.n_kchunks * p.w_blk_bytes;
  p.w_resident = (uint32_t)L.total_blocks * p.w_blk_bytes <= 64u * 1024u ? 1 : 0;
  p.SW = 6;
  if (const char* e = getenv("RB_TC_PD_SW")) p.SW = std::min(8, std::max(2, atoi(e)));   // experiments only
  while (!p.w_resident && p.SW > 2 && fixed + w1_bytes + p.SW * p.w_blk_bytes + 3 * p.a_stage_bytes > budget) --p.SW;
  const uint32_t w_bytes = p.w_resident ? (uint32_t)L.total_blocks * p.w_blk_bytes : p.SW * p.w_blk_bytes;
  p.SA = (int)std::min<uint32_t>(12, (budget - fixed - w1_bytes - w_bytes) / p.a_stage_bytes);
  RB_CHECK(p.SA >= 2, RB_ERR_UNSUPPORTED, "pair DilatedUnit does not fit in shared memory");
  p.off_a = fixed;
  p.off_w = p.off_a + p.SA * p.a_stage_bytes;
  p.off_h = p.off_w + w_bytes;                             // (resident 1x1-conv weights)
  p.smem_bytes = p.off_h + w1_bytes;
  p.SD = 1; p.tmem_cols = 512;
  if (const char* e = getenv("RB_TC_PD_DBG")) p.dbg = atoi(e);                           // experiments only
  p.tiles_per_stream = ((p.y.n + 127) / 128 + 1) / 2;     // work slots: one 128-row tile per CTA of a pair
  p.n_work = B * p.tiles_per_stream;
  if (getenv("RB_TC_PLAN_DEBUG"))
    fprintf(stderr, "tc layer: pair DilatedUnit C %d dil %d | resident %d SA %d SW %d rows %d work %d smem %u\n", L.N_TILE,
            L.dil, p.w_resident, p.SA, p.SW, p.y.n, p.n_work, p.smem_bytes);
  p.reverse = next_direction();
  return launch_tc_du_pair(p, s);
}

// ------------------------------------------------------------------- elementwise kernels
// Bandwidth-bound single passes over group-major tensors; the flat index runs in the memory
// order of the output (stream, channel group, row, channel-in-group).
__device__ __forceinline__ size_t eidx(const Buf& b, int bi, int t, int c) {
  if (b.cp == 1) return (size_t)bi * b.ld + b.hist + t;
  return (((size_t)bi * (b.cp >> 3) + (c >> 3)) * b.ld + b.hist + t) * 8 + (c & 7);
}
__device__ __forceinline__ void decode_gm(size_t i, const Buf& y, int& b, int& t, int& c) {
  const int e = (int)(i & 7);
  size_t r = i >> 3;
  t = (int)(r % y.n);
  r /= y.n;
  const int G = y.cp >> 3;
  c = (int)(r % G) * 8 + e;
  b = (int)(r / G);
}

// [B][C][n] (caller layout, dense) -> engine layout
__global__ void cf_to_gm_kernel(const __nv_bfloat16* __restrict__ src, Buf dst, int C, int B) {
  const size_t total = (size_t)B * dst.n * dst.cp;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int b, t, c;
    decode_gm(i, dst, b, t, c);
    dst.p[eidx(dst, b, t, c)] = c < C ? src[((size_t)b * C + c) * dst.n + t] : __float2bfloat16_rn(0.f);
  }
}
// engine layout -> [B][C][n]
__global__ void gm_to_cf_kernel(Buf src, __nv_bfloat16* __restrict__ dst, int C, int B) {
  const size_t total = (size_t)B * C * src.n;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(i % src.n);
    const size_t r = i / src.n;
    const int c = (int)(r % C), b = (int)(r / C);
    dst[i] = src.p[eidx(src, b, t, c)];
  }
}
// model.py:618-620: z = cat(z, speaker.repeat(B, 1, F)) (+ zero pad channels)
// One thread per 16-byte unit (8 channels of one row): z groups are copied, conditioning groups are
// the same 8 values for every row of every stream, pad groups are zero.
__global__ void concat_gm_kernel(Buf z, Buf y, CondArgs ca, int c_z, int B) {
  const int c_cond = ca.c_spk;
  const int G = y.cp >> 3, Gz = z.cp >> 3;
  const long long units = (long long)B * G * y.n;
  for (long long u = blockIdx.x * (long long)blockDim.x + threadIdx.x; u < units; u += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(u % y.n);
    const int r = (int)(u / y.n);
    const int g = r % G, b = r / G;
    const int c0 = g * 8;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (c0 + 8 <= c_z) {
      v = reinterpret_cast<const uint4*>(z.p)[((size_t)b * Gz + g) * z.ld + z.hist + t];
    } else {
      __align__(16) __nv_bfloat16 e[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int c = c0 + k;
        float v32 = 0.f;
        if (c < c_z) v32 = __bfloat162float(z.p[eidx(z, b, t, c)]);
        else if (c < c_z + c_cond) v32 = ca.spk[(size_t)(ca.b0 + b) * ca.spk_stride + (c - c_z)];
        else if (ca.f0 && c < c_z + c_cond + ca.f0_C)
          v32 = ca.f0[((size_t)(ca.b0 + b) * ca.f0_C + (c - c_z - c_cond)) * ca.f0_F + t];
        e[k] = __float2bfloat16_rn(v32);
      }
      v = *reinterpret_cast<const uint4*>(e);
    }
    reinterpret_cast<uint4*>(y.p)[((size_t)b * G + g) * y.ld + y.hist + t] = v;
  }
}
// AdaIN transfer (blocks.py:885-891) as x = a[b][c]*x + b[b][c], in place.  One thread per 16-byte unit (8
// channels of one row); consecutive threads take consecutive rows of one channel group: 512 contiguous bytes
// per warp each way (the scalar version moved 2 bytes per thread and ran at 1 TB/s).
__global__ void affine_gm_kernel(Buf x, const float* __restrict__ a, const float* __restrict__ bb, int C, int B) {
  const int G = x.cp >> 3;
  const long long units = (long long)B * G * x.n;
  for (long long u = blockIdx.x * (long long)blockDim.x + threadIdx.x; u < units; u += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(u % x.n);
    const int r = (int)(u / x.n);
    const int g = r % G, b = r / G;
    uint4* q = reinterpret_cast<uint4*>(x.p) + (((size_t)b * G + g) * x.ld + x.hist + t);
    uint4 v = *q;
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
    const float* ap = a + (size_t)b * C + g * 8;
    const float* bp = bb + (size_t)b * C + g * 8;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int c = g * 8 + 2 * k;
      float lo = __uint_as_float(w[k] << 16), hi = __uint_as_float(w[k] & 0xffff0000u);
      if (c < C) lo = fmaf(lo, __ldg(ap + 2 * k), __ldg(bp + 2 * k));
      if (c + 1 < C) hi = fmaf(hi, __ldg(ap + 2 * k + 1), __ldg(bp + 2 * k + 1));
      __nv_bfloat162 h2 = __floats2bfloat162_rn(lo, hi);
      w[k] = *reinterpret_cast<uint32_t*>(&h2);
    }
    *q = make_uint4(w[0], w[1], w[2], w[3]);
  }
}
// AdaIN running statistics on a group-major tensor: one CTA per (stream, channel), two passes.
__global__ void __launch_bounds__(256)
adain_stats_gm_kernel(Buf x, int C, float* __restrict__ mean_t, float* __restrict__ std_t, int count) {
  __shared__ float sh[8];
  const int b = blockIdx.x / C, c = blockIdx.x - b * C;
  const __nv_bfloat16* p = x.p + eidx(x, b, 0, c);        // element t sits 8 * t further on
  auto bsum = [&](float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sh[w];
    return t;
  };
  float s1 = 0.f;
  for (int t = threadIdx.x; t < x.n; t += 256) s1 += __bfloat162float(p[(size_t)t * 8]);
  const float mean = bsum(s1) / (float)x.n;
  float s2 = 0.f;
  for (int t = threadIdx.x; t < x.n; t += 256) { const float d = __bfloat162float(p[(size_t)t * 8]) - mean; s2 = fmaf(d, d, s2); }
  const float var = bsum(s2) / (float)(x.n - 1);
  if (threadIdx.x == 0) {
    const float inv = 1.f / (float)(count + 1);
    const int row = blockIdx.x;
    mean_t[row] += (mean - mean_t[row]) * inv;
    std_t[row] += (sqrtf(var) - std_t[row]) * inv;
  }
}
// GeneratorV2.forward tail (blocks.py:699-707) + reverse_half of the synthesis bank input
__global__ void head_v2_gm_kernel(Buf w, Buf nz, Buf y, int data, int am, int B) {
  const size_t total = (size_t)B * y.n * y.cp;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int b, t, c;
    decode_gm(i, y, b, t, c);
    float v = 0.f;
    if (c < data) {
      v = __bfloat162float(w.p[eidx(w, b, t, c)]);
      if (am) v *= 1.f / (1.f + __expf(-__bfloat162float(w.p[eidx(w, b, t, data + c)])));
      if (nz.p) v += __bfloat162float(nz.p[eidx(nz, b, t, c)]);
      v = tanhf(v);
      if ((c & 1) && !(t & 1)) v = -v;
    }
    y.p[eidx(y, b, t, c)] = __float2bfloat16_rn(v);
  }
}
// Generator.forward tail (blocks.py:412): tanh(waveform) * mod_sigmoid(loudness)
__global__ void head_v1_gm_kernel(Buf wv, Buf ld, Buf y, int data, int B) {
  const size_t total = (size_t)B * y.n * y.cp;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int b, t, c;
    decode_gm(i, y, b, t, c);
    float v = 0.f;
    if (c < data) {
      const float sg = 1.f / (1.f + expf(-__bfloat162float(ld.p[eidx(ld, b, t, 0)])));
      v = tanhf(__bfloat162float(wv.p[eidx(wv, b, t, c)])) * (2.f * powf(sg, 2.3f) + 1e-7f);
      if ((c & 1) && !(t & 1)) v = -v;
    }
    y.p[eidx(y, b, t, c)] = __float2bfloat16_rn(v);
  }
}
// NoiseGeneratorV2.forward tail (blocks.py:281-291, core.py:66,95-129) in closed form; one
// thread per (stream, block q, band d) -- see kernels_f32.cu for the derivation.
__global__ void noise_v2_gm_kernel(Buf amp, const float* __restrict__ noise, long long nbs, Buf y, int data,
                                   int nblk, unsigned long long seed, int B) {
  const size_t total = (size_t)B * nblk * data;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int q = (int)(i % nblk);
    const size_t r = i / nblk;
    const int d = (int)(r % data), b = (int)(r / data);
    float a[5];
#pragma unroll
    for (int n = 0; n < 5; ++n) {
      const float hh = __bfloat162float(amp.p[eidx(amp, b, q, d * 5 + n)]) - 5.f;
      const float sg = 1.f / (1.f + expf(-hh));
      a[n] = 2.f * powf(sg, 2.3f) + 1e-7f;
    }
    float ir[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const float c1 = cospif(t * 0.25f), c2 = cospif(t * 0.5f), c3 = cospif(t * 0.75f);
      const float c4 = (t & 1) ? -1.f : 1.f;
      const float v = (a[0] + 2.f * (a[1] * c1 + a[2] * c2 + a[3] * c3) + a[4] * c4) * 0.125f;
      const int u = (t + 4) & 7;
      ir[t] = v * 0.5f * (1.f - cospif(u * 0.25f));
    }
    float nzv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (noise) {
        nzv[j] = noise[(size_t)b * nbs + ((size_t)q * data + d) * 8 + j];
      } else {
        unsigned long long z = seed + (((unsigned long long)b * nblk + q) * data * 8 + d * 8 + j) * 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        nzv[j] = (float)(z >> 40) * (2.0f / 16777216.0f) - 1.0f;
      }
    }
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      float sacc = 0.f;
#pragma unroll
      for (int j = 0; j <= t; ++j) sacc = fmaf(nzv[j], ir[t - j], sacc);
      y.p[eidx(y, b, q * 8 + t, d)] = __float2bfloat16_rn(sacc);
    }
  }
}
__global__ void zero_pad_channels_kernel(Buf y, int C, int B) {
  const size_t total = (size_t)B * y.n * y.cp;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int b, t, c;
    decode_gm(i, y, b, t, c);
    if (c >= C) y.p[eidx(y, b, t, c)] = __float2bfloat16_rn(0.f);
  }
}

static inline int ew_grid(size_t total) {
  return (int)std::min<size_t>((total + 255) / 256, 148 * 16);
}
#define RB_LAUNCH_EW(kernel, total, s, ...)                          \
  do {                                                                \
    if ((total) > 0) {                                                \
      kernel<<<ew_grid(total), 256, 0, s>>>(__VA_ARGS__);             \
      ++g_launches;                                                   \
      RB_CUDA(cudaGetLastError());                                    \
    }                                                                 \
  } while (0)

}  // namespace tc

using namespace tc;

static void free_fold(Plan* pl) {
  for (auto& L : pl->fold_layers) free_layer(L);
  for (float* p : pl->fold_wc) cudaFree(p);
  cudaFree(pl->fold_sbias);
}

void tc_free(rb_handle* h) {
  Plan* pl = (Plan*)h->tc;
  if (!pl) return;
  for (auto& L : pl->layers) free_layer(L);
  for (auto& L : pl->sr_layers) free_layer(L);
  free_fold(pl);
  delete pl;
  h->tc = nullptr;
}

// Plans every conv-like op (called when all weights are present; re-run on every reload).
int tc_pack_weights(rb_handle* h) {
  tc_free(h);
  Plan* pl = new Plan();
  h->tc = pl;
  const int nops = (int)h->ops.size();
  pl->layers.resize(nops);
  pl->sr_layers.resize(nops);
  pl->fold_layers.resize(nops);
  pl->fold_wc.assign(nops, nullptr);
  pl->skip.assign(nops, 0);
  pl->cp.resize(h->slots.size());
  for (size_t s = 0; s < h->slots.size(); ++s) pl->cp[s] = slot_cp(h, (int)s);
  // Activated storage: a slot written by a conv and read only by convs that pre-activate it
  // with LeakyReLU (or add it as a residual) is stored as leaky(x) by its producer.
  {
    const int ns = (int)h->slots.size();
    std::vector<int> ok(ns, 1), conv_reads(ns, 0), conv_writes(ns, 0);
    auto bad = [&](int sl) { if (sl >= 0 && sl < ns) ok[sl] = 0; };
    for (int i = 0; i < nops; ++i) {
      const rb_op& op = h->ops[i];
      if (op.kind == RB_OP_CONV) {
        if (op.pre_act == RB_ACT_LEAKY && op.src_ch_off == 0) ++conv_reads[op.src]; else bad(op.src);
        ++conv_writes[op.dst];
      } else {
        bad(op.src); bad(op.dst); bad(op.res);
      }
      bad(op.src2);
    }
    bad(h->spec.enc_in_slot); bad(h->spec.enc_out_slot); bad(h->spec.dec_in_slot); bad(h->spec.dec_out_slot);
    bad(h->spec.noise_slot);
    pl->sact.assign(ns, 0);
    for (int sl = 0; sl < ns; ++sl) pl->sact[sl] = ok[sl] && conv_reads[sl] > 0 && conv_writes[sl] > 0;
    if ((int)h->tc_sact_hint.size() == ns) pl->sact = h->tc_sact_hint;     // rb_debug_conv_bf16
  }
  for (int i = 0; i < nops; ++i) {
    const rb_op& op = h->ops[i];
    if (pl->skip[i]) continue;
    if (op.kind != RB_OP_CONV && op.kind != RB_OP_PQMF_FWD && op.kind != RB_OP_PQMF_INV) continue;
    ViewWeight v;
    int rc = build_view_weight(h, i, *pl, v);
    if (rc) return rc;
    // DilatedUnit fusion: conv(k, dil) -> act -> conv(1x1) + residual of the unit's input
    bool fuse = false, pairdu = false;
    ViewWeight v2;
    if (op.kind == RB_OP_CONV && i + 1 < nops && i + 1 != h->spec.n_enc_ops) {
      const rb_op& o2 = h->ops[i + 1];
      fuse = o2.kind == RB_OP_CONV && o2.k == 1 && o2.stride == 1 && o2.ileave <= 1 && op.stride == 1 &&
             op.ileave <= 1 && o2.src == op.dst && o2.res == op.src && op.res < 0 && o2.pre_act == op.pre_act &&
             v.N_total == v.cin_v && v.N_total <= 128 && pl->cp[o2.dst] == v.N_total &&
             (size_t)v.N_total * v.cin_v * (v.n_taps + 1) * 2 <= (size_t)128 * 1024;
      // offline LeakyReLU nets on activated storage, C = 256: both convs in one cluster-pair launch
      // (tc_du_pair_kernel).  The kernel also runs C = 128 / 64 (two tiles in tensor memory, weights resident) but
      // measured SLOWER there than the one-CTA fused kernel -- 517 vs 464 and 1296 vs 785 kcycles per launch: the
      // narrow units are bound by the instructions of their output warps, not by the chain between them -- so
      // those widths are behind RB_TC_PD128 / RB_TC_PD64 (experiments, and the tests of the two-stage path)
      const bool pd_width = v.N_total == 256 || (v.N_total == 128 && getenv("RB_TC_PD128")) ||
                            (v.N_total == 64 && getenv("RB_TC_PD64"));
      pairdu = o2.kind == RB_OP_CONV && o2.k == 1 && o2.stride == 1 && o2.ileave <= 1 && op.stride == 1 &&
               op.ileave <= 1 && o2.src == op.dst && o2.res == op.src && op.res < 0 && o2.pre_act == op.pre_act &&
               op.pre_act == RB_ACT_LEAKY && pd_width && v.cin_v == v.N_total && v.n_taps == 3 &&
               pl->cp[o2.dst] == v.N_total && !h->spec.streaming && pl->sact[op.src] && op.has_bias && o2.has_bias &&
               !getenv("RB_TC_NOPAIRDU");
      if (pairdu)
        for (int j = 0; j < nops; ++j)                   // h must have no other reader
          if (j != i + 1 && (h->ops[j].src == op.dst || h->ops[j].res == op.dst || h->ops[j].src2 == op.dst)) pairdu = false;
      if (pairdu) fuse = false;
      if (fuse || pairdu) {
        rc = build_view_weight(h, i + 1, *pl, v2);
        if (rc) return rc;
      }
    }
    if (!fuse && !pairdu && op.res < 0 && !h->spec.streaming) {   // (streaming keeps plain row views: history alignment)
      const int F = choose_row_batch(v, h->slots[op.dst].per_frame / v.fout);
      if (F > 1) batch_rows(v, F);
    }
    Layer& L = pl->layers[i];
    rc = pack_layer(v, 64, (fuse || pairdu) ? &v2 : nullptr, L, 0, pairdu);
    if (rc) return rc;
    if (getenv("RB_TC_PLAN_DEBUG") && op.kind == RB_OP_CONV && i > 0 && h->ops[i - 1].kind == RB_OP_CONCAT_COND)
      fprintf(stderr, "fold candidate op %d: src match %d fuse %d pairdu %d streaming %d k %d stride %d dil %d ileave %d res %d "
              "pre_act %d taps %d off %d fin %d fout %d cin_v %d N %d\n", i, (int)(h->ops[i - 1].dst == op.src), (int)fuse,
              (int)pairdu, (int)h->spec.streaming, op.k, op.stride, op.dilation, op.ileave, op.res, op.pre_act, v.n_taps,
              v.off, v.fin, v.fout, v.cin_v, v.N_total);
    // folded conditioning (Plan::fold_layers): CONCAT_COND -> this conv, offline
    // (in the reference's graph the concatenation is the LAST encoder op -- encode() returns the conditioned latent,
    // model.py:615-620 -- and this conv the first decoder op: the fold applies to calls that span both programs)
    if (op.kind == RB_OP_CONV && i > 0 && h->ops[i - 1].kind == RB_OP_CONCAT_COND &&
        (h->ops[i - 1].dst == op.src || (i == h->spec.n_enc_ops && h->ops[i - 1].dst == h->spec.enc_out_slot &&
                                         op.src == h->spec.dec_in_slot)) &&
        !fuse && !pairdu && !h->spec.streaming && op.k == 3 && op.stride == 1 && op.dilation == 1 &&
        std::max(op.ileave, 1) == 1 && op.res < 0 && op.pre_act == RB_ACT_NONE && v.n_taps == 3 && v.off == 1 &&
        v.fin == 1 && v.fout == 1 && !getenv("RB_TC_NOFOLD")) {
      const rb_op& cc = h->ops[i - 1];
      const int c_z = cc.c_in, c_spk = h->spec.cond_channels;
      bool only_reader = true;
      for (int j = 0; j < nops; ++j)
        if (j != i && (h->ops[j].src == op.src || h->ops[j].res == op.src || h->ops[j].src2 == op.src)) only_reader = false;
      if (getenv("RB_TC_PLAN_DEBUG"))
        fprintf(stderr, "  fold: only_reader %d c_z %d c_spk %d cp_src %d cp_dst %d\n", (int)only_reader, c_z, c_spk,
                pl->cp[cc.src], pl->cp[op.dst]);
      if (only_reader && c_spk > 0 && c_z % 16 == 0 && pl->cp[cc.src] == c_z && c_z + c_spk <= v.cin_v &&
          v.N_total % 32 == 0 && (pl->cp[op.dst] / 8) % 4 == 0) {
        ViewWeight vz;
        vz.N_total = v.N_total; vz.cin_v = c_z; vz.n_taps = 3; vz.dil = 1; vz.off = 1; vz.fin = vz.fout = 1; vz.out_mask = 0;
        vz.w.assign((size_t)vz.N_total * c_z * 3, 0.f);
        for (int n = 0; n < v.N_total; ++n)
          for (int c = 0; c < c_z; ++c)
            for (int j = 0; j < 3; ++j) vz.at(n, c, j) = v.at(n, c, j);
        vz.bias = v.bias;
        if ((rc = pack_layer(vz, 64, nullptr, pl->fold_layers[i]))) return rc;
        if (pl->fold_layers[i].N_TILE % 32 != 0) free_layer(pl->fold_layers[i]);   // needs the drain_fast variant
        // the speaker channels' weights as the MMA would see them (bf16), [tap][channel][n]
        std::vector<float> wc((size_t)3 * c_spk * v.N_total);
        for (int j = 0; j < 3; ++j)
          for (int c = 0; c < c_spk; ++c)
            for (int n = 0; n < v.N_total; ++n)
              wc[((size_t)j * c_spk + c) * v.N_total + n] = __bfloat162float(f2bf(v.at(n, c_z + c, j)));
        if (pl->fold_layers[i].valid && (rc = dev_upload(wc, &pl->fold_wc[i]))) return rc;
      }
    }
    if (pairdu) {
      RB_CHECK(L.w2 && L.N_TILE == v.N_total && L.KC == 64 && L.n_kchunks * 64 == L.N_TILE &&
                   L.total_blocks == 3 * L.n_kchunks, RB_ERR_STATE, "pair DilatedUnit: unexpected packing");
      L.pairdu = true;
      L.fused_with = i + 1;
      L.tail = i + 1;
      L.act2 = h->ops[i + 1].pre_act;
      if ((rc = dev_upload(v2.bias, &L.bias2))) return rc;
      pl->skip[i + 1] = 1;
    }
    if (h->spec.streaming && op.kind == RB_OP_CONV && v.out_mask == 0 &&
        pl->cp[op.src] != 1 && pl->cp[op.dst] != 1 &&
        h->slots[op.dst].per_frame / v.fout <= (fuse ? 32 : 8) &&      // fused units: also the 32-row C = 128 stage
        (size_t)v.cin_v * v.n_taps <= 16384 && !getenv("RB_TC_NOSR")) {
      ViewWeight u;                                    // taps unrolled into channels: c' = j * cin_v + c
      u.N_total = v.N_total; u.cin_v = v.cin_v * v.n_taps; u.n_taps = 1; u.dil = 1; u.off = 0;
      u.fin = v.fin; u.fout = v.fout; u.out_mask = 0;
      u.w.assign((size_t)u.N_total * u.cin_v, 0.f);
      for (int n = 0; n < v.N_total; ++n)
        for (int c = 0; c < v.cin_v; ++c)
          for (int j = 0; j < v.n_taps; ++j)
            u.w[(size_t)n * u.cin_v + (size_t)j * v.cin_v + c] = v.w[((size_t)n * v.cin_v + c) * v.n_taps + j];
      u.bias = v.bias;
      if (!v.alpha.empty())
        for (int j = 0; j < v.n_taps; ++j) u.alpha.insert(u.alpha.end(), v.alpha.begin(), v.alpha.end());
      // 1024 streams are only 8 row tiles: narrow output tiles spread one launch over many SMs (each CTA
      // streams its slice of the weights serially).  A K chunk must stay inside one tap / sub-row phase
      // and the kernel's mask table holds 128 (n-tile, chunk) pairs: first shape that fits.
      Layer& S = pl->sr_layers[i];
      const int cps = pl->cp[op.src];
      const int n_small = std::max(64, round16((u.N_total + TC_MAX_NT - 1) / TC_MAX_NT));
      const int shapes[4][2] = {{n_small, 64}, {n_small, 128}, {256, 64}, {256, 128}};
      if (fuse) {                                       // a fused unit is one n-tile; 64-channel K chunks
        const int kc = largest_div16(u.cin_v, largest_div16(cps, 64));
        if (cps % kc == 0 && u.cin_v / kc <= TC_MAX_MASKS) {
          rc = pack_layer(u, largest_div16(cps, 64), &v2, S, 256);
          if (rc) return rc;
        }
      }
      for (int a = 0; a < 4 && !S.valid && !fuse; ++a) {
        const int nt = largest_div16(u.N_total, shapes[a][0]), kc = largest_div16(u.cin_v, largest_div16(cps, shapes[a][1]));
        if (cps % kc != 0 || (u.N_total / nt) > TC_MAX_NT || (u.N_total / nt) * (u.cin_v / kc) > TC_MAX_MASKS) continue;
        rc = pack_layer(u, largest_div16(cps, shapes[a][1]), nullptr, S, shapes[a][0]);
        if (rc) return rc;
      }
      S.sr_taps = v.n_taps; S.sr_dil = v.dil; S.sr_off = v.off;
      if (fuse && S.valid) {
        S.fused = true; S.fused_with = i + 1; S.tail = i + 1; S.act2 = h->ops[i + 1].pre_act;
        if (!v2.bias.empty() && (rc = dev_upload(v2.bias, &S.bias2))) return rc;
        if (!v2.alpha.empty() && (rc = upload_alpha(v2.alpha, &S.alpha2))) return rc;
      }
    }
    // conv -> HEAD_V2 (no noise branch, amplitude modulation, 16 bands): the head runs in the epilogue
    if (!fuse && op.kind == RB_OP_CONV && op.res < 0 && i + 1 < nops && i + 1 != h->spec.n_enc_ops) {
      const rb_op& o2 = h->ops[i + 1];
      bool only_reader = true;
      for (int j = 0; j < nops; ++j)
        if (j != i + 1 && (h->ops[j].src == op.dst || h->ops[j].src2 == op.dst || h->ops[j].res == op.dst)) only_reader = false;
      if (o2.kind == RB_OP_HEAD_V2 && o2.src == op.dst && o2.src2 < 0 && o2.k == 1 && o2.c_out == 16 &&
          op.c_out == 32 && std::max(op.ileave, 1) == 1 && only_reader && pl->cp[o2.dst] == 16 &&
          op.dst != h->spec.dec_out_slot && L.N_total % 32 == 0) {
        L.head = L.N_total / 32;
        L.tail = i + 1;
        pl->skip[i + 1] = 1;
      }
    }
    if (fuse) {
      L.fused = true;
      L.fused_with = i + 1;
      L.tail = i + 1;
      L.act2 = h->ops[i + 1].pre_act;
      if (!v2.bias.empty() && (rc = dev_upload(v2.bias, &L.bias2))) return rc;
      if (!v2.alpha.empty() && (rc = upload_alpha(v2.alpha, &L.alpha2))) return rc;
      pl->skip[i + 1] = 1;
    }
  }
  return RB_OK;
}

// Per-stream bias of a folded conditioning conv: V_j[n] = sum_c Wc[j][c][n] * bf16(spk[b][c]) for the three taps;
// row classes: interior V0 + V1 + V2, first row of a stream (tap 0 falls on the zero padding) V1 + V2, last row
// V0 + V1.  One block per stream, thread n-strided; the speaker row goes through shared memory.
__global__ void cond_fold_kernel(const float* __restrict__ wc, const float* __restrict__ spk, long long spk_stride,
                                 int b0, int c_spk, int N, float* __restrict__ out) {
  extern __shared__ float srow[];
  const int b = blockIdx.x;
  for (int c = threadIdx.x; c < c_spk; c += blockDim.x)
    srow[c] = __bfloat162float(__float2bfloat16_rn(spk[(size_t)(b0 + b) * spk_stride + c]));
  __syncthreads();
  for (int n = threadIdx.x; n < N; n += blockDim.x) {
    float v0 = 0.f, v1 = 0.f, v2 = 0.f;
    for (int c = 0; c < c_spk; ++c) {
      const float sv = srow[c];
      v0 = fmaf(wc[((size_t)0 * c_spk + c) * N + n], sv, v0);
      v1 = fmaf(wc[((size_t)1 * c_spk + c) * N + n], sv, v1);
      v2 = fmaf(wc[((size_t)2 * c_spk + c) * N + n], sv, v2);
    }
    float* o = out + (size_t)b * 3 * N + n;
    o[0] = v0 + v1 + v2;
    o[N] = v1 + v2;
    o[2 * N] = v0 + v1;
  }
}

// ------------------------------------------------------------------- offline op interpreter
static int tc_exec_op(rb_handle* h, Plan* pl, int i, std::vector<Buf>& bufs, int B, const float* noise,
                      int64_t noise_bstride, cudaStream_t s) {
  const rb_op& op = h->ops[i];
  const Layer& L = pl->layers[i];
  switch (op.kind) {
    case RB_OP_CONV:
    case RB_OP_PQMF_FWD:
    case RB_OP_PQMF_INV: {
      RB_CHECK(L.valid, RB_ERR_STATE, "op %d has no bf16 plan", i);
      const std::vector<int>& sa = pl->sact;
      if (pl->fold_now == i) {
        pl->fold_now = -1;
        const rb_op& cc = h->ops[i - 1];
        const Layer& FL = pl->fold_layers[i];
        CondArgs ca;
        int rc = cond_args(h, i - 1, B, bufs[op.dst].n, &ca);
        if (rc) return rc;
        const int N = FL.N_total;
        const size_t need = (size_t)B * 3 * N;
        if (pl->fold_sbias_cap < need) {
          RB_CUDA(cudaStreamSynchronize(s));
          cudaFree(pl->fold_sbias);
          pl->fold_sbias = nullptr; pl->fold_sbias_cap = 0;
          RB_CUDA(cudaMalloc((void**)&pl->fold_sbias, need * sizeof(float)));
          pl->fold_sbias_cap = need;
          pl->fold_key_op = -1;
        }
        if (pl->fold_key_op != i || pl->fold_key_spk != ca.spk || pl->fold_key_stride != ca.spk_stride ||
            pl->fold_key_b0 != ca.b0 || pl->fold_key_B != B || pl->fold_key_version != h->cond_version ||
            pl->fold_key_stream != s) {
          cond_fold_kernel<<<B, 256, (size_t)ca.c_spk * sizeof(float), s>>>(pl->fold_wc[i], ca.spk, ca.spk_stride, ca.b0,
                                                                            ca.c_spk, N, pl->fold_sbias);
          ++g_launches;
          RB_CUDA(cudaGetLastError());
          pl->fold_key_op = i; pl->fold_key_spk = ca.spk; pl->fold_key_stride = ca.spk_stride; pl->fold_key_b0 = ca.b0;
          pl->fold_key_B = B; pl->fold_key_version = h->cond_version; pl->fold_key_stream = s;
        }
        g_sbias = pl->fold_sbias; g_sbias_n = N;
        return run_layer(FL, op, nullptr, bufs[cc.src], bufs[op.dst], nullptr, B, s, sa[cc.src], 0, sa[op.dst]);
      }
      if (L.pairdu) {
        const rb_op& o2 = h->ops[L.fused_with];
        RB_CHECK(sa[op.src], RB_ERR_STATE, "pair DilatedUnit: input is not stored activated");
        return run_pair_du(L, op, bufs[op.src], bufs[o2.dst], B, s, sa[o2.res], sa[o2.dst]);
      }
      if (L.fused) {
        const rb_op& o2 = h->ops[L.fused_with];
        const Layer& S = pl->sr_layers[i];
        const int rows = bufs[o2.dst].n;
        const int sr_min_b = getenv("RB_TC_SR_MINB") ? atoi(getenv("RB_TC_SR_MINB")) : 128;
        if (S.valid && S.fused && h->spec.streaming && B >= sr_min_b && rows >= 1 && rows <= 32) {
          if (getenv("RB_TC_SR_DEBUG")) fprintf(stderr, "op %d fused unit C %d rows %d sr_valid 1\n", i, op.c_in, rows);
          return run_layer(S, op, &o2, bufs[op.src], bufs[o2.dst], &bufs[o2.res], B, s, sa[op.src], sa[o2.res],
                           sa[o2.dst], rows);
        }
        return run_layer(L, op, &o2, bufs[op.src], bufs[o2.dst], &bufs[o2.res], B, s, sa[op.src], sa[o2.res],
                         sa[o2.dst]);
      }
      if (L.head)
        return run_layer(L, op, nullptr, bufs[op.src], bufs[h->ops[L.tail].dst], nullptr, B, s, sa[op.src], 0, 0);
      {
        const Layer& S = pl->sr_layers[i];
        const Buf& src = bufs[op.src];
        const int rows = bufs[op.dst].n / L.fout;
        const int sr_min_b = getenv("RB_TC_SR_MINB") ? atoi(getenv("RB_TC_SR_MINB")) : 128;   // tests lower it
        if (getenv("RB_TC_SR_DEBUG"))
          fprintf(stderr, "op %d conv %d->%d k%d s%d il%d rows %d sr_valid %d lo %d N_TILE %d KC %d nk %d\n", i, op.c_in,
                  op.c_out, op.k, op.stride, op.ileave, rows, (int)S.valid, src.lo, S.valid ? S.N_TILE : L.N_TILE,
                  S.valid ? S.KC : L.KC, S.valid ? S.n_kchunks : L.n_kchunks);
        if (S.valid && h->spec.streaming && B >= sr_min_b && rows >= 1 && rows <= 8) {
          return run_layer(S, op, nullptr, src, bufs[op.dst], op.res >= 0 ? &bufs[op.res] : nullptr, B, s,
                           sa[op.src], op.res >= 0 ? sa[op.res] : 0, sa[op.dst], rows);
        }
      }
      return run_layer(L, op, nullptr, bufs[op.src], bufs[op.dst], op.res >= 0 ? &bufs[op.res] : nullptr, B, s,
                       sa[op.src], op.res >= 0 ? sa[op.res] : 0, sa[op.dst]);
    }
    case RB_OP_AFFINE: {
      const OpWeights& w = h->w[i];
      RB_CHECK(w.aff_a && w.aff_b, RB_ERR_STATE, "op %d: AdaIN statistics not loaded", i);
      const int b0 = h->cur_b0;
      RB_CHECK(b0 + B <= w.aff_rows, RB_ERR_INVALID,
               "batch %d exceeds the %d rows of AdaIN statistics (cc.MAX_BATCH_SIZE)", b0 + B, w.aff_rows);
      const Buf& x = bufs[op.src];
      RB_CHECK(x.cp != 1, RB_ERR_UNSUPPORTED, "AdaIN on an audio-layout tensor");
      const size_t ro = (size_t)b0 * op.c_in;
      bool apply = true;
      int rc = adain_step(h, h->w[i], op.c_in, s, [&](float* mean_t, float* std_t, int count) -> int {
        if (x.n > 0 && B > 0) {
          adain_stats_gm_kernel<<<B * op.c_in, 256, 0, s>>>(x, op.c_in, mean_t + ro, std_t + ro, count);
          ++g_launches;
          RB_CUDA(cudaGetLastError());
        }
        return (int)RB_OK;
      }, &apply);
      if (rc) return rc;
      if (apply) RB_LAUNCH_EW(affine_gm_kernel, (size_t)B * x.n * (x.cp / 8), s, x, w.aff_a + ro, w.aff_b + ro, op.c_in, B);
      return RB_OK;
    }
    case RB_OP_HEAD_V2: {
      const Buf& y = bufs[op.dst];
      RB_LAUNCH_EW(head_v2_gm_kernel, (size_t)B * y.n * y.cp, s, bufs[op.src],
                   op.src2 >= 0 ? bufs[op.src2] : Buf(), y, op.c_out, op.k, B);
      return RB_OK;
    }
    case RB_OP_HEAD_V1: {
      const Buf& y = bufs[op.dst];
      RB_LAUNCH_EW(head_v1_gm_kernel, (size_t)B * y.n * y.cp, s, bufs[op.src], bufs[op.src2], y, op.c_out, B);
      return RB_OK;
    }
    case RB_OP_NOISE_V2: {
      const Buf& amp = bufs[op.src];
      const Buf& y = bufs[op.dst];
      if (y.cp > op.c_out) RB_LAUNCH_EW(zero_pad_channels_kernel, (size_t)B * y.n * y.cp, s, y, op.c_out, B);
      RB_LAUNCH_EW(noise_v2_gm_kernel, (size_t)B * amp.n * op.c_out, s, amp, noise, (long long)noise_bstride, y,
                   op.c_out, amp.n, h->noise_seed++, B);
      return RB_OK;
    }
    case RB_OP_CONCAT_COND: {
      const Buf& y = bufs[op.dst];
      CondArgs ca;
      int rc = cond_args(h, i, B, y.n, &ca);
      if (rc) return rc;
      pl->fold_now = -1;
      if (i + 1 < pl->run_oe && pl->fold_layers[i + 1].valid && !ca.f0 && y.n >= 2 && !pl->skip[i + 1] &&
          !getenv("RB_TC_NOFOLD")) {
        pl->fold_now = i + 1;                 // the next conv reads the latent alone; nothing to materialise
        return RB_OK;
      }
      RB_LAUNCH_EW(concat_gm_kernel, (size_t)B * y.n * (y.cp / 8), s, bufs[op.src], y, ca, op.c_in, B);
      return RB_OK;
    }
  }
  set_error("op %d: unknown kind %d", i, op.kind);
  return RB_ERR_INVALID;
}

// Runs ops [ob, oe) on bf16 buffers.  `in` / `out` are in the CALLER's layout ([B,1,T] audio
// or [B,C,F] latents, dense); a range that spans both programs keeps the latent on the
// device in group-major form (rb_forward).
int tc_run_program(rb_handle* h, int ob, int oe, const void* in, void* out, int B, int F, cudaStream_t s) {
  Plan* pl = (Plan*)h->tc;
  RB_CHECK(pl, RB_ERR_STATE, "bf16 engine: weights are not packed");
  const int ns = (int)h->slots.size();
  const int n_enc = h->spec.n_enc_ops;
  const bool has_enc = ob < n_enc, has_dec = oe > n_enc;
  const int in_slot = has_enc ? h->spec.enc_in_slot : h->spec.dec_in_slot;
  const int out_slot = has_dec ? h->spec.dec_out_slot : h->spec.enc_out_slot;
  if (has_enc && has_dec)
    RB_CHECK(pl->cp[h->spec.enc_out_slot] == pl->cp[h->spec.dec_in_slot], RB_ERR_INVALID,
             "encoder output and decoder input widths differ");
  std::vector<int> last_use(ns, -1), buf_of(ns, -1);
  pl->run_oe = oe;
  pl->fold_now = -1;
  for (int i = ob; i < oe; ++i) {
    if (pl->skip[i]) continue;
    const rb_op& op = h->ops[i];
    const rb_op& ol = pl->layers[i].tail >= 0 ? h->ops[pl->layers[i].tail] : op;
    last_use[op.src] = i;
    if (ol.res >= 0 && !pl->layers[i].head) last_use[ol.res] = i;
    if (op.src2 >= 0) last_use[op.src2] = i;
    last_use[ol.dst] = std::max(last_use[ol.dst], i);
    // folded conditioning: the conv behind CONCAT_COND reads the concat's INPUT
    if (pl->fold_layers[i].valid && i > ob) last_use[h->ops[i - 1].src] = std::max(last_use[h->ops[i - 1].src], i);
  }
  if (has_enc && has_dec) last_use[h->spec.enc_out_slot] = std::max(last_use[h->spec.enc_out_slot], last_use[h->spec.dec_in_slot]);
  auto slot_rows = [&](int sl) { return h->slots[sl].per_frame * F; };
  auto slot_bytes = [&](int sl, int nb) { return (size_t)nb * slot_rows(sl) * pl->cp[sl] * sizeof(__nv_bfloat16); };
  int mb = h->microbatch;
  if (mb <= 0) {
    size_t per_stream = 0;
    for (int sl = 0; sl < ns; ++sl) per_stream = std::max(per_stream, slot_bytes(sl, 1));
    const size_t budget = (size_t)2 << 30;   // persistent kernels want >= several waves of tiles per launch
    mb = (int)std::max<size_t>(1, std::min<size_t>(B, budget / std::max<size_t>(per_stream, 1)));
  }
  const int in_ch = h->slots[in_slot].channels, out_ch = h->slots[out_slot].channels;
  for (int b0 = 0; b0 < B; b0 += mb) {
    const int nb = std::min(mb, B - b0);
    std::vector<Buf> bufs(ns);
    h->cur_b0 = b0;
    for (auto& pb : h->pool) pb.busy = false;
    std::fill(buf_of.begin(), buf_of.end(), -1);
    auto alloc_slot = [&](int sl) -> int {
      int bi;
      int rc = pool_get(h, std::max<size_t>(slot_bytes(sl, nb), 256), &bi);
      if (rc) return rc;
      buf_of[sl] = bi;
      Buf& b = bufs[sl];
      b.p = (__nv_bfloat16*)h->pool[bi].p;
      b.cp = pl->cp[sl]; b.n = slot_rows(sl); b.ld = b.n; b.hist = 0; b.lo = 0;
      return RB_OK;
    };
    // program input
    if (pl->cp[in_slot] == 1) {
      Buf& b = bufs[in_slot];
      b.p = (__nv_bfloat16*)in + (size_t)b0 * slot_rows(in_slot);
      b.cp = 1; b.n = slot_rows(in_slot); b.ld = b.n;
    } else {
      int rc = alloc_slot(in_slot);
      if (rc) return rc;
      const Buf& b = bufs[in_slot];
      RB_LAUNCH_EW(cf_to_gm_kernel, (size_t)nb * b.n * b.cp, s,
                   (const __nv_bfloat16*)in + (size_t)b0 * in_ch * b.n, b, in_ch, nb);
    }
    if (pl->cp[out_slot] == 1) {
      Buf& b = bufs[out_slot];
      b.p = (__nv_bfloat16*)out + (size_t)b0 * slot_rows(out_slot);
      b.cp = 1; b.n = slot_rows(out_slot); b.ld = b.n;
    }
    const float* noise = nullptr;
    int64_t noise_bstride = 0;
    if (h->spec.noise_slot >= 0 && h->noise) {
      const rb_slot& nsl = h->slots[h->spec.noise_slot];
      noise_bstride = (int64_t)nsl.channels * nsl.per_frame * F;
      RB_CHECK(h->noise_numel >= (int64_t)B * noise_bstride, RB_ERR_INVALID, "noise tensor has %lld elements, need %lld",
               (long long)h->noise_numel, (long long)B * noise_bstride);
      noise = h->noise + (size_t)b0 * noise_bstride;
    }
    tc::g_reverse = getenv("RB_TC_NOREVERSE") ? -1 : 0;
    for (int i = ob; i < oe; ++i) {
      if (i == n_enc && has_enc && has_dec) {      // latent stays on the device, group-major
        bufs[h->spec.dec_in_slot] = bufs[h->spec.enc_out_slot];
        buf_of[h->spec.dec_in_slot] = buf_of[h->spec.enc_out_slot];
        buf_of[h->spec.enc_out_slot] = -1;
      }
      if (pl->skip[i]) continue;
      const rb_op& op = h->ops[i];
      const rb_op& ol = pl->layers[i].tail >= 0 ? h->ops[pl->layers[i].tail] : op;
      if (!bufs[ol.dst].p) {
        int rc = alloc_slot(ol.dst);
        if (rc) return rc;
      }
      RB_CHECK(bufs[op.src].p, RB_ERR_STATE, "op %d reads slot %d before it is written", i, op.src);
      if (op.src2 >= 0 && op.kind != RB_OP_NOISE_V2)
        RB_CHECK(bufs[op.src2].p, RB_ERR_STATE, "op %d reads slot %d before it is written", i, op.src2);
      int rc;
      {
        OpTimer tm(h, i, s);
        rc = tc_exec_op(h, pl, i, bufs, nb, noise, noise_bstride, s);
      }
      if (rc) return rc;
      const int used[3] = {op.src, pl->layers[i].head ? -1 : ol.res, op.src2};
      for (int u : used)
        if (u >= 0 && last_use[u] == i && buf_of[u] >= 0) {
          h->pool[buf_of[u]].busy = false;
          buf_of[u] = -1;
        }
    }
    if (pl->cp[out_slot] != 1) {
      const Buf& b = bufs[out_slot];
      RB_CHECK(b.p, RB_ERR_STATE, "program output was never written");
      RB_LAUNCH_EW(gm_to_cf_kernel, (size_t)nb * out_ch * b.n, s, b,
                   (__nv_bfloat16*)out + (size_t)b0 * out_ch * b.n, out_ch, nb);
    }
  }
  h->cur_b0 = 0;
  tc::g_reverse = -1;
  adain_finalize(h, ob, oe);
  return RB_OK;
}

}  // namespace rb

// ------------------------------------------------------------------------- bf16 streaming
// cached_conv semantics (scripts/export.py:542-594) on the tensor-core engine.  Same scheme as
// the fp32 engine (stream.cu): every slot owns one persistent group-major buffer
// [max_streams][Cp/8][hist + max rows][8] whose first `hist` rows are the left context its
// consumers need (CachedPadding1d state of each conv, AlignBranches delay line, transposed-conv
// overlap).  A step writes the new rows behind the context, runs the same kernels as the offline
// path with lo = -hist (the TMA tensor map starts at the history rows), then rolls the tail of
// each row window to the front.  hist is rounded up to 16 rows so that every strided / phase /
// audio view of the slot stays aligned.
namespace rb {
namespace tc {

__global__ void roll_gm_kernel(Buf b, int rows_per_stream_group, int n) {
  // one CTA per (stream, channel group): rows [n, n + hist) -> [0, hist), 16-byte units
  uint4* base = reinterpret_cast<uint4*>(b.p) + (size_t)blockIdx.x * rows_per_stream_group;
  for (int r0 = 0; r0 < b.hist; r0 += blockDim.x) {
    const int r = r0 + threadIdx.x;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r < b.hist) v = base[r + n];
    __syncthreads();
    if (r < b.hist) base[r] = v;
    __syncthreads();
  }
}
// All group-major slots of a step in ONE launch (a step used to end with ~35 roll launches, 17 % of
// its time): CTA -> (slot, stream, channel group) through the prefix table of the descriptor block.
constexpr int ROLL_MAX = 48;
struct RollSet {
  uint4* base[ROLL_MAX];
  int ld[ROLL_MAX], hist[ROLL_MAX], n[ROLL_MAX];
  int first[ROLL_MAX + 1];    // first WARP of each slot
  int skip[ROLL_MAX];         // hist is rounded up to 16 rows for alignment; only the last hist - skip rows are state
  int nblk[ROLL_MAX];         // (stream, channel group) blocks of the slot
  int lpb[ROLL_MAX];          // lanes per block: a warp rolls 32 / lpb blocks at once (state is 2-18 rows per block)
  int count;
};
__global__ void __launch_bounds__(256) roll_many_kernel(const __grid_constant__ RollSet rs) {
  // A warp takes 32 / lpb consecutive (stream, channel group) blocks of one slot, lpb lanes each; rows are read
  // before they are written (the windows overlap when a chunk is shorter than the state).  The slot of a warp is
  // found by bisection of the prefix table: the first version walked it linearly (~30 steps for each of the ~10^6
  // one-block warps of a 1024-stream step, to then move 2-16 rows): 0.2 ms per step for 86 MB of state.
  const int gw = (int)((blockIdx.x * 256u + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (gw >= rs.first[rs.count]) return;
  int lo = 0, hi = rs.count;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (gw >= rs.first[mid]) lo = mid; else hi = mid;
  }
  const int k = lo, lpb = rs.lpb[k];
  const int blk = (gw - rs.first[k]) * (32 / lpb) + lane / lpb, l = lane % lpb;
  const bool live = blk < rs.nblk[k];
  uint4* base = rs.base[k] + (size_t)(live ? blk : 0) * rs.ld[k];
  const int hist = rs.hist[k], n = rs.n[k];
  for (int r0 = rs.skip[k]; r0 < hist; r0 += lpb) {
    const int r = r0 + l;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (live && r < hist) v = base[r + n];
    __syncwarp();
    if (live && r < hist) base[r] = v;
    __syncwarp();
  }
}
__global__ void roll_audio_kernel(__nv_bfloat16* p, int ld, int hist, int n) {
  __nv_bfloat16* base = p + (size_t)blockIdx.x * ld;
  for (int r0 = 0; r0 < hist; r0 += blockDim.x) {
    const int r = r0 + threadIdx.x;
    __nv_bfloat16 v = __float2bfloat16_rn(0.f);
    if (r < hist) v = base[r + n];
    __syncthreads();
    if (r < hist) base[r] = v;
    __syncthreads();
  }
}
__global__ void copy_gm_kernel(Buf src, Buf dst, int B) {
  const size_t total = (size_t)B * dst.n * dst.cp;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int b, t, c;
    decode_gm(i, dst, b, t, c);
    dst.p[eidx(dst, b, t, c)] = src.p[eidx(src, b, t, c)];
  }
}

static Buf stream_buf(rb_handle* h, Plan* pl, int sl, int F) {
  Buf b;
  b.p = (__nv_bfloat16*)h->sbuf[sl];
  b.cp = pl->cp[sl]; b.ld = h->s_ld[sl]; b.hist = h->s_hist[sl];
  b.n = h->slots[sl].per_frame * F; b.lo = -b.hist;
  return b;
}

}  // namespace tc

int tc_stream_alloc(rb_handle* h) {
  Plan* pl = (Plan*)h->tc;
  RB_CHECK(pl, RB_ERR_STATE, "bf16 engine: weights are not packed");
  const int ns = (int)h->slots.size();
  for (int sl = 0; sl < ns; ++sl) {
    const rb_slot& sd = h->slots[sl];
    if (sl == h->spec.noise_slot) continue;
    h->s_hist[sl] = (sd.hist + 15) & ~15;
    h->s_ld[sl] = h->s_hist[sl] + ((sd.per_frame * h->max_frames + 15) & ~15);
    const size_t bytes = (size_t)h->max_streams * pl->cp[sl] * h->s_ld[sl] * sizeof(__nv_bfloat16);
    RB_CUDA(cudaMalloc(&h->sbuf[sl], bytes));
    RB_CUDA(cudaMemset(h->sbuf[sl], 0, bytes));
  }
  return RB_OK;
}

int tc_stream_reset(rb_handle* h, const int32_t* ids, int n) {
  Plan* pl = (Plan*)h->tc;
  RB_CHECK(pl, RB_ERR_STATE, "bf16 engine: weights are not packed");
  for (size_t sl = 0; sl < h->slots.size(); ++sl) {
    if (!h->sbuf[sl]) continue;
    const size_t row = (size_t)pl->cp[sl] * h->s_ld[sl] * sizeof(__nv_bfloat16);
    if (!ids) {
      RB_CUDA(cudaMemset(h->sbuf[sl], 0, row * h->max_streams));
    } else {
      for (int i = 0; i < n; ++i) {
        RB_CHECK(ids[i] >= 0 && ids[i] < h->max_streams, RB_ERR_INVALID, "stream id %d out of range", ids[i]);
        RB_CUDA(cudaMemset((char*)h->sbuf[sl] + row * ids[i], 0, row));
      }
    }
  }
  return RB_OK;
}

// Runs ops [ob, oe) of one program on the persistent buffers.  `in` (caller layout) is copied
// behind the history of the program's input slot when given; `out` receives the program output.
int tc_stream_run(rb_handle* h, int ob, int oe, int B, int F, const void* in, void* out, cudaStream_t s) {
  Plan* pl = (Plan*)h->tc;
  RB_CHECK(pl, RB_ERR_STATE, "bf16 engine: weights are not packed");
  const int ns = (int)h->slots.size();
  const bool is_enc = ob < h->spec.n_enc_ops;
  const int in_slot = is_enc ? h->spec.enc_in_slot : h->spec.dec_in_slot;
  const int out_slot = is_enc ? h->spec.enc_out_slot : h->spec.dec_out_slot;
  std::vector<Buf> bufs(ns);
  for (int sl = 0; sl < ns; ++sl)
    if (h->sbuf[sl]) bufs[sl] = stream_buf(h, pl, sl, F);
  if (in) {
    const Buf& b = bufs[in_slot];
    if (b.cp == 1) {
      RB_CUDA(cudaMemcpy2DAsync(b.p + b.hist, (size_t)b.ld * 2, in, (size_t)b.n * 2, (size_t)b.n * 2, B,
                                cudaMemcpyDeviceToDevice, s));
    } else {
      const int in_ch = h->slots[in_slot].channels;
      RB_LAUNCH_EW(cf_to_gm_kernel, (size_t)B * b.n * b.cp, s, (const __nv_bfloat16*)in, b, in_ch, B);
    }
  }
  const bool direct_out = out && pl->cp[out_slot] == 1 && h->s_hist[out_slot] == 0;
  if (direct_out) {
    Buf& b = bufs[out_slot];
    b.p = (__nv_bfloat16*)out; b.ld = b.n; b.hist = 0; b.lo = 0;
  }
  const float* noise = nullptr;
  int64_t noise_bstride = 0;
  if (h->spec.noise_slot >= 0 && h->noise) {
    const rb_slot& nsl = h->slots[h->spec.noise_slot];
    noise_bstride = (int64_t)nsl.channels * nsl.per_frame * F;
    RB_CHECK(h->noise_numel >= (int64_t)B * noise_bstride, RB_ERR_INVALID, "noise tensor too small");
    noise = h->noise;
  }
  for (int i = ob; i < oe; ++i) {
    if (pl->skip[i]) continue;
    int rc = tc_exec_op(h, pl, i, bufs, B, noise, noise_bstride, s);
    if (rc) return rc;
  }
  if (out && !direct_out) {
    const Buf& b = bufs[out_slot];
    if (b.cp == 1) {
      RB_CUDA(cudaMemcpy2DAsync(out, (size_t)b.n * 2, b.p + b.hist, (size_t)b.ld * 2, (size_t)b.n * 2, B,
                                cudaMemcpyDeviceToDevice, s));
    } else {
      const int out_ch = h->slots[out_slot].channels;
      RB_LAUNCH_EW(gm_to_cf_kernel, (size_t)B * out_ch * b.n, s, b, (__nv_bfloat16*)out, out_ch, B);
    }
  }
  // state update: every slot this program reads with history
  std::vector<char> seen(ns, 0);
  RollSet rs;
  rs.count = 0;
  rs.first[0] = 0;
  auto flush = [&]() -> int {
    if (rs.count == 0) return RB_OK;
    roll_many_kernel<<<(rs.first[rs.count] + 7) / 8, 256, 0, s>>>(rs);
    RB_CUDA(cudaGetLastError());
    ++g_launches;
    rs.count = 0;
    return RB_OK;
  };
  for (int i = ob; i < oe; ++i) {
    const rb_op& op = h->ops[i];
    const int used[3] = {op.src, op.res, op.src2};
    for (int u : used) {
      if (u < 0 || seen[u] || !h->sbuf[u] || h->s_hist[u] == 0 || u == h->spec.noise_slot) continue;
      seen[u] = 1;
      const Buf& b = bufs[u];
      if (b.cp == 1) {
        roll_audio_kernel<<<B, 256, 0, s>>>(b.p, b.ld, b.hist, b.n);
        RB_CUDA(cudaGetLastError());
        ++g_launches;
      } else if (getenv("RB_TC_ROLL_EACH")) {
        // streams [0, B) x groups are contiguous blocks of ld rows
        roll_gm_kernel<<<B * (b.cp / 8), 128, 0, s>>>(b, b.ld, b.n);
        RB_CUDA(cudaGetLastError());
        ++g_launches;
      } else {
        const int k = rs.count++;
        rs.base[k] = reinterpret_cast<uint4*>(b.p);
        rs.ld[k] = b.ld; rs.hist[k] = b.hist; rs.n[k] = b.n;
        rs.skip[k] = std::max(0, b.hist - h->slots[u].hist);
        const int state = b.hist - rs.skip[k];
        int lpb = 2;
        while (lpb < 32 && lpb < state) lpb <<= 1;
        rs.lpb[k] = lpb;
        rs.nblk[k] = B * (b.cp / 8);
        rs.first[k + 1] = rs.first[k] + (rs.nblk[k] + 32 / lpb - 1) / (32 / lpb);
        if (rs.count == ROLL_MAX) { int rc = flush(); if (rc) return rc; }
      }
    }
  }
  { int rc = flush(); if (rc) return rc; }
  adain_finalize(h, ob, oe);
  return RB_OK;
}

// latent hand-over of rb_stream_step: encoder output slot -> decoder input slot (behind its history)
int tc_stream_handoff(rb_handle* h, int B, int F, cudaStream_t s) {
  Plan* pl = (Plan*)h->tc;
  const Buf src = stream_buf(h, pl, h->spec.enc_out_slot, F), dst = stream_buf(h, pl, h->spec.dec_in_slot, F);
  RB_CHECK(src.cp == dst.cp, RB_ERR_INVALID, "encoder output and decoder input widths differ");
  RB_LAUNCH_EW(copy_gm_kernel, (size_t)B * dst.n * dst.cp, s, src, dst, B);
  return RB_OK;
}

}  // namespace rb

// ------------------------------------------------------------------------- testing hook
// Runs one conv op (or a conv + 1x1 conv pair, fused when the planner fuses it) of the bf16
// engine on host fp32 data: x [B][c_in][n_in] -> y [B][c_out/ileave][n_out].  Used by
// tests/test_gpu_tc_conv.py to check the tcgen05 kernel layer by layer against numpy.
extern "C" int rb_debug_conv_bf16(const rb_debug_conv* d) {
  using namespace rb;
  RB_CHECK(d && d->n_ops >= 1 && d->n_ops <= 2 && d->x && d->y && d->w[0], RB_ERR_INVALID, "rb_debug_conv_bf16: bad arguments");
  rb_handle hh;
  rb_handle* h = &hh;
  h->spec.enc_in_slot = h->spec.dec_out_slot = -1;
  h->spec.n_enc_ops = 0;
  h->spec.noise_slot = -1;
  const rb_op& o0 = d->op[0];
  const int r0 = std::max(o0.ileave, 1);
  const int n_mid = d->n_in / o0.stride * r0;
  // slots: 0 in, 1 out of op0, 2 out of op1, 3 explicit residual
  h->slots.push_back({o0.c_in, d->n_in, 0, 0});
  h->slots.push_back({o0.c_out / r0, n_mid, 0, 0});
  if (d->n_ops == 2) h->slots.push_back({d->op[1].c_out, n_mid, 0, 0});
  else h->slots.push_back({1, 1, 0, 0});
  h->slots.push_back({o0.c_out / r0, n_mid, 0, 0});
  for (int i = 0; i < d->n_ops; ++i) {
    rb_op op = d->op[i];
    op.kind = RB_OP_CONV;
    op.src = i == 0 ? 0 : 1;
    op.dst = i == 0 ? 1 : 2;
    op.src2 = -1;
    op.res = i == 1 ? 0 : (d->res ? 3 : -1);
    op.src_ch_off = 0;
    h->ops.push_back(op);
  }
  h->w.resize(h->ops.size());
  for (int i = 0; i < d->n_ops; ++i) {
    const rb_op& op = h->ops[i];
    h->w[i].w_host.assign(d->w[i], d->w[i] + (size_t)op.c_out * op.c_in * op.k);
    if (op.has_bias) {
      RB_CHECK(d->bias[i], RB_ERR_INVALID, "bias missing");
      h->w[i].bias_host.assign(d->bias[i], d->bias[i] + op.c_out / std::max(op.ileave, 1));
    }
    if (op.pre_act == RB_ACT_SNAKE) {
      RB_CHECK(d->alpha[i], RB_ERR_INVALID, "alpha missing");
      h->w[i].alpha_host.assign(d->alpha[i], d->alpha[i] + op.c_in);
    }
  }
  {
    // activated-storage flags of the four slots, known to the planner (a C = 256 pair needs its input activated)
    const int osl = d->n_ops == 2 ? 2 : 1;
    h->tc_sact_hint.assign(h->slots.size(), 0);
    h->tc_sact_hint[0] = (d->flags & 1) ? 1 : 0;
    h->tc_sact_hint[osl] = (d->flags & 2) ? 1 : 0;
    h->tc_sact_hint[3] = (d->flags & 4) ? 1 : 0;
  }
  int rc = tc_pack_weights(h);
  if (rc) { tc_free(h); return rc; }
  Plan* pl = (Plan*)h->tc;
  const int B = d->B;
  std::vector<Buf> bufs(h->slots.size());
  std::vector<void*> to_free;
  auto mk = [&](int sl, const float* host) -> int {
    Buf& b = bufs[sl];
    b.cp = pl->cp[sl]; b.n = h->slots[sl].per_frame; b.ld = b.n; b.hist = 0; b.lo = 0;
    const size_t elems = (size_t)B * b.n * b.cp;
    RB_CUDA(cudaMalloc((void**)&b.p, elems * 2));
    to_free.push_back(b.p);
    RB_CUDA(cudaMemset(b.p, 0xFF, elems * 2));     // NaN fill: unwritten outputs show up
    if (host) {
      const int C = h->slots[sl].channels;
      std::vector<__nv_bfloat16> tmp(elems, __float2bfloat16_rn(0.f));
      for (int bi = 0; bi < B; ++bi)
        for (int c = 0; c < C; ++c)
          for (int t = 0; t < b.n; ++t)
            tmp[(((size_t)bi * (b.cp >> 3) + (c >> 3)) * b.n + t) * 8 + (c & 7)] = __float2bfloat16_rn(host[((size_t)bi * C + c) * b.n + t]);
      RB_CUDA(cudaMemcpy(b.p, tmp.data(), elems * 2, cudaMemcpyHostToDevice));
    }
    return RB_OK;
  };
  {
    const int osl = d->n_ops == 2 ? 2 : 1;
    pl->sact.assign(h->slots.size(), 0);
    pl->sact[0] = (d->flags & 1) ? 1 : 0;         // x is given (and stored) LeakyReLU-activated
    pl->sact[osl] = (d->flags & 2) ? 1 : 0;       // y is returned LeakyReLU-activated
    pl->sact[3] = (d->flags & 4) ? 1 : 0;         // the explicit residual is given activated
  }
  rc = mk(0, d->x);
  if (!rc) rc = mk(1, nullptr);
  if (!rc && d->n_ops == 2) rc = mk(2, nullptr);
  if (!rc && d->res) rc = mk(3, d->res);
  if (!rc && d->trace) {
    RB_CUDA(cudaMalloc((void**)&tc::g_trace, 64 * 16 * 8));
    RB_CUDA(cudaMemset(tc::g_trace, 0, 64 * 16 * 8));
  }
  for (int i = 0; i < d->n_ops && !rc; ++i) {
    if (pl->skip[i]) continue;
    rc = tc_exec_op(h, pl, i, bufs, B, nullptr, 0, 0);
  }
  if (tc::g_trace) {
    cudaDeviceSynchronize();
    cudaMemcpy(d->trace, tc::g_trace, 64 * 16 * 8, cudaMemcpyDeviceToHost);
    cudaFree(tc::g_trace);
    tc::g_trace = nullptr;
  }
  if (!rc) {
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { set_error("debug conv failed: %s", cudaGetErrorString(e)); rc = RB_ERR_CUDA; }
  }
  if (!rc) {
    const int osl = d->n_ops == 2 ? 2 : 1;
    const Buf& b = bufs[osl];
    const int C = h->slots[osl].channels;
    std::vector<__nv_bfloat16> tmp((size_t)B * b.n * b.cp);
    cudaMemcpy(tmp.data(), b.p, tmp.size() * 2, cudaMemcpyDeviceToHost);
    for (int bi = 0; bi < B; ++bi)
      for (int c = 0; c < C; ++c)
        for (int t = 0; t < b.n; ++t)
          d->y[((size_t)bi * C + c) * b.n + t] = __bfloat162float(tmp[(((size_t)bi * (b.cp >> 3) + (c >> 3)) * b.n + t) * 8 + (c & 7)]);
    if (d->fused_out) *d->fused_out = pl->layers[0].pairdu ? 2 : (pl->layers[0].fused ? 1 : 0);
  }
  for (void* p : to_free) cudaFree(p);
  tc_free(h);
  return rc;
}
