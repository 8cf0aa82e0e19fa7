// engine.cu -- C-ABI entry points, weight packing, workspace planning and the op
// interpreter of the fp32 engine.  Host C++; every device operation goes through the
// launch_* functions of kernels_f32.cu / rvq.cu / (bf16) tc_engine.cu.
#include <stdarg.h>
#include <string.h>

#include <algorithm>

#include "engine.h"

namespace rb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

struct BusyGuard {
  rb_handle* h;
  bool ok;
  explicit BusyGuard(rb_handle* h_) : h(h_), ok(!h_->busy.test_and_set()) {}
  ~BusyGuard() { if (ok) h->busy.clear(); }
};

static int upload(const float* host, size_t n, float** dev) {
  if (*dev) { cudaFree(*dev); *dev = nullptr; }
  RB_CUDA(cudaMalloc((void**)dev, std::max<size_t>(n, 1) * sizeof(float)));
  RB_CUDA(cudaMemcpy(*dev, host, n * sizeof(float), cudaMemcpyHostToDevice));
  return RB_OK;
}


// ------------------------------------------------------------------ offline workspace pool
int pool_get(rb_handle* h, size_t bytes, int* idx) {
  int best = -1;
  for (size_t i = 0; i < h->pool.size(); ++i)
    if (!h->pool[i].busy && h->pool[i].bytes >= bytes &&
        (best < 0 || h->pool[i].bytes < h->pool[best].bytes)) best = (int)i;
  if (best < 0) {
    // grow the largest free buffer if there is one, else add a new one
    for (size_t i = 0; i < h->pool.size(); ++i)
      if (!h->pool[i].busy && (best < 0 || h->pool[i].bytes > h->pool[best].bytes)) best = (int)i;
    if (best >= 0) {
      RB_CUDA(cudaFree(h->pool[best].p));
      h->pool[best].p = nullptr;
    } else {
      h->pool.push_back(PoolBuf());
      best = (int)h->pool.size() - 1;
    }
    RB_CUDA(cudaMalloc(&h->pool[best].p, bytes));
    h->pool[best].bytes = bytes;
  }
  h->pool[best].busy = true;
  *idx = best;
  return RB_OK;
}

OpTimer::OpTimer(rb_handle* h_, int oi_, cudaStream_t s_) : h(h_), s(s_), oi(oi_), on(h_->profile) {
  if (!on) return;
  rb_handle::Ev e{oi, nullptr, nullptr};
  cudaEventCreate(&e.a);
  cudaEventCreate(&e.b);
  cudaEventRecord(e.a, s);
  h->events.push_back(e);
}
OpTimer::~OpTimer() {
  if (on) cudaEventRecord(h->events.back().b, s);
}

// ------------------------------------------------------------------------- op interpreter
int exec_op_f32(RunCtx& c, int oi) {
  const rb_op& op = c.h->ops[oi];
  const OpWeights& w = c.h->w[oi];
  const View& x = c.v[op.src];
  const View& y = c.v[op.dst];
  const int lo = c.streaming ? -x.hist : 0;
  switch (op.kind) {
    case RB_OP_CONV: {
      ConvArgs a;
      a.x = x; a.y = y;
      a.res = op.res >= 0 ? c.v[op.res] : View();
      a.w = w.w; a.bias = op.has_bias ? w.bias : nullptr; a.alpha = w.alpha;
      a.B = c.B; a.c_in = op.c_in; a.c_out = op.c_out; a.k = op.k; a.stride = op.stride;
      a.dil = op.dilation; a.off = op.off; a.ileave = op.ileave; a.pre_act = op.pre_act;
      a.res_delay = op.res_delay; a.lo = lo;
      RB_CHECK(op.pre_act != RB_ACT_SNAKE || w.alpha, RB_ERR_STATE, "op %d: Snake alpha not loaded", oi);
      return launch_conv_f32(a, c.s);
    }
    case RB_OP_PQMF_FWD:
      return launch_pqmf_fwd_f32(x, y, w.w, c.B, op.c_out, op.n_band, op.k, op.off, lo, c.s);
    case RB_OP_PQMF_INV:
      return launch_pqmf_inv_f32(x, y, w.w, c.B, op.n_band, op.k, op.off, lo, c.s);
    case RB_OP_AFFINE: {
      RB_CHECK(w.aff_a && w.aff_b, RB_ERR_STATE, "op %d: AdaIN statistics not loaded", oi);
      const int b0 = c.h->cur_b0;
      RB_CHECK(b0 + c.B <= w.aff_rows, RB_ERR_INVALID,
               "batch %d exceeds the %d rows of AdaIN statistics (cc.MAX_BATCH_SIZE)", b0 + c.B, w.aff_rows);
      bool apply = true;
      OpWeights& wm = c.h->w[oi];
      const size_t ro = (size_t)b0 * op.c_in;
      int rc = adain_step(c.h, wm, op.c_in, c.s, [&](float* mean_t, float* std_t, int count) -> int {
        return launch_adain_stats_f32(x, c.B, mean_t + ro, std_t + ro, count, c.s);
      }, &apply);
      if (rc) return rc;
      return apply ? launch_affine_f32(x, w.aff_a + ro, w.aff_b + ro, c.B, c.s) : RB_OK;
    }
    case RB_OP_HEAD_V2:
      return launch_head_v2_f32(x, op.src2 >= 0 ? c.v[op.src2] : View(), y, c.B, op.c_out, op.k, c.s);
    case RB_OP_HEAD_V1:
      return launch_head_v1_f32(x, c.v[op.src2], y, c.B, op.c_out, c.s);
    case RB_OP_NOISE_V2:
      return launch_noise_v2_f32(x, c.noise, c.noise_bstride, y, c.B, op.c_out,
                                 c.h->noise_seed++, c.s);
    case RB_OP_CONCAT_COND: {
      CondArgs ca;
      int rc = cond_args(c.h, oi, c.B, y.n, &ca);
      if (rc) return rc;
      return launch_concat_f32(x, y, ca, c.B, op.c_in, op.c_out, c.s);
    }
  }
  set_error("op %d: unknown kind %d", oi, op.kind);
  return RB_ERR_INVALID;
}

// bf16 tensor-core engine (tc_engine.cu)
int tc_run_program(rb_handle* h, int op_begin, int op_end, const void* in, void* out, int B, int F,
                   cudaStream_t s);
int tc_pack_weights(rb_handle* h);
void tc_free(rb_handle* h);

static size_t slot_bytes_f32(const rb_handle* h, int slot, int B, int F) {
  const rb_slot& s = h->slots[slot];
  return (size_t)B * s.channels * round4(s.per_frame * F) * sizeof(float);
}

// Runs ops [op_begin, op_end) offline on fp32 device buffers. `in` is the program input
// slot's tensor, `out` the output slot's, both dense [B][C][n].
static int run_offline_f32(rb_handle* h, int op_begin, int op_end, int in_slot, int out_slot,
                           const float* in, float* out, int B, int F, cudaStream_t s) {
  const int ns = (int)h->slots.size();
  std::vector<int> last_use(ns, -1), buf_of(ns, -1);
  for (int i = op_begin; i < op_end; ++i) {
    const rb_op& op = h->ops[i];
    last_use[op.src] = i;
    if (op.res >= 0) last_use[op.res] = i;
    if (op.src2 >= 0) last_use[op.src2] = i;
    last_use[op.dst] = std::max(last_use[op.dst], i);
  }
  // microbatch so that the live activations stay within a budget (and mostly inside L2)
  int mb = h->microbatch;
  if (mb <= 0) {
    size_t per_stream = 0;
    for (int sl = 0; sl < ns; ++sl) per_stream = std::max(per_stream, slot_bytes_f32(h, sl, 1, F));
    const size_t budget = (size_t)768 << 20;            // per live buffer
    mb = (int)std::max<size_t>(1, std::min<size_t>(B, budget / std::max<size_t>(per_stream, 1)));
  }
  for (int b0 = 0; b0 < B; b0 += mb) {
    const int nb = std::min(mb, B - b0);
    RunCtx c{h, nb, F, std::vector<View>(ns), s, false, nullptr, 0};
    h->cur_b0 = b0;
    for (auto& pb : h->pool) pb.busy = false;
    std::fill(buf_of.begin(), buf_of.end(), -1);
    auto make_view = [&](int sl, float* p) {
      View v;
      v.p = p; v.C = h->slots[sl].channels; v.n = h->slots[sl].per_frame * F;
      v.ld = v.n; v.hist = 0;
      return v;
    };
    c.v[in_slot] = make_view(in_slot, const_cast<float*>(in) +
                                          (size_t)b0 * h->slots[in_slot].channels * h->slots[in_slot].per_frame * F);
    c.v[out_slot] = make_view(out_slot, out + (size_t)b0 * h->slots[out_slot].channels *
                                                  h->slots[out_slot].per_frame * F);
    if (h->spec.noise_slot >= 0 && h->noise) {
      const rb_slot& nsl = h->slots[h->spec.noise_slot];
      c.noise_bstride = (int64_t)nsl.channels * nsl.per_frame * F;
      RB_CHECK(h->noise_numel >= (int64_t)B * c.noise_bstride, RB_ERR_INVALID,
               "noise tensor has %lld elements, need %lld", (long long)h->noise_numel,
               (long long)B * c.noise_bstride);
      c.noise = h->noise + (size_t)b0 * c.noise_bstride;
    }
    for (int i = op_begin; i < op_end; ++i) {
      const rb_op& op = h->ops[i];
      if (!c.v[op.dst].p) {
        int bi;
        int rc = pool_get(h, slot_bytes_f32(h, op.dst, nb, F), &bi);
        if (rc) return rc;
        buf_of[op.dst] = bi;
        c.v[op.dst] = make_view(op.dst, (float*)h->pool[bi].p);
        c.v[op.dst].ld = round4(c.v[op.dst].n);
      }
      if (op.src2 >= 0 && op.kind != RB_OP_NOISE_V2)
        RB_CHECK(c.v[op.src2].p, RB_ERR_STATE, "op %d reads slot %d before it is written", i, op.src2);
      RB_CHECK(c.v[op.src].p, RB_ERR_STATE, "op %d reads slot %d before it is written", i, op.src);
      int rc;
      {
        OpTimer tm(h, i, s);
        rc = exec_op_f32(c, i);
      }
      if (rc) return rc;
      const int used[3] = {op.src, op.res, op.src2};
      for (int u : used)
        if (u >= 0 && last_use[u] == i && buf_of[u] >= 0) {
          h->pool[buf_of[u]].busy = false;
          buf_of[u] = -1;
        }
    }
  }
  h->cur_b0 = 0;
  adain_finalize(h, op_begin, op_end);
  return RB_OK;
}

// ---------------------------------------------------------------------------- AdaIN learning
int adain_step(rb_handle* h, OpWeights& w, int C, cudaStream_t s,
               const std::function<int(float*, float*, int)>& stats_fn, bool* apply) {
  *apply = true;
  if (!w.stats) return RB_OK;                    // plain a / b tables (RB_W_AFFINE_A / _B): always applied
  const size_t n = (size_t)w.aff_rows * C;
  int rc = RB_OK;
  if (h->adain_learn_y) {                        // blocks.py:897-905: learn the target, pass x through
    if ((rc = stats_fn(w.stats + 2 * n, w.stats + 3 * n, w.ny))) return rc;
    w.touch_y = true;
    *apply = false;
  } else if (h->adain_learn_x) {                 // :908-914, then the transfer with the new statistics
    if ((rc = stats_fn(w.stats, w.stats + n, w.nx))) return rc;
    w.touch_x = true;
  }
  const bool active = (w.nx + (w.touch_x ? 1 : 0)) > 0 && (w.ny + (w.touch_y ? 1 : 0)) > 0;
  if (w.touch_x || w.touch_y)
    if ((rc = launch_adain_ab(w.stats, w.aff_rows, C, active ? 1 : 0, w.aff_a, w.aff_b, s))) return rc;
  if (!active) *apply = false;                   // :916: identity until both sides have statistics
  return RB_OK;
}

void adain_finalize(rb_handle* h, int ob, int oe) {
  for (int i = ob; i < oe && i < (int)h->w.size(); ++i) {
    OpWeights& w = h->w[i];
    if (w.touch_x) ++w.nx;
    if (w.touch_y) ++w.ny;
    w.touch_x = w.touch_y = false;
  }
}

}  // namespace rb

using namespace rb;

// ================================================================================ C-ABI
namespace rb {
int cond_args(rb_handle* h, int oi, int B, int n, CondArgs* out) {
  const rb_op& op = h->ops[oi];
  const OpWeights& w = h->w[oi];
  const int c_spk = h->spec.cond_channels, f0_C = op.c_out - op.c_in - c_spk;
  RB_CHECK(w.cond, RB_ERR_STATE, "conditioning vector (speaker embedding) not loaded");
  RB_CHECK(w.cond_rows == 1 || h->cur_b0 + B <= w.cond_rows, RB_ERR_INVALID,
           "batch %d exceeds the %d rows of the speaker table (rb_set_conditioning)", h->cur_b0 + B, w.cond_rows);
  out->spk = w.cond;
  out->spk_stride = w.cond_rows == 1 ? 0 : c_spk;
  out->c_spk = c_spk;
  out->f0 = nullptr;
  out->f0_C = f0_C;
  out->f0_F = n;
  out->b0 = h->cur_b0;
  if (h->f0_B > 0) {
    RB_CHECK(f0_C > 0, RB_ERR_INVALID, "pitch channels are set but the decoder takes none (latent_size %d = %d + %d)",
             op.c_out, op.c_in, c_spk);
    RB_CHECK(h->f0_C == f0_C && h->f0_F == n && h->cur_b0 + B <= h->f0_B, RB_ERR_INVALID,
             "pitch channels are [%d, %d, %d], this call needs [>= %d, %d, %d]", h->f0_B, h->f0_C, h->f0_F,
             h->cur_b0 + B, f0_C, n);
    out->f0 = h->f0;
  }
  return RB_OK;
}
}  // namespace rb

extern "C" {

const char* rb_last_error(void) { return g_err; }
int rb_abi_version(void) { return RB_ABI_VERSION; }

int rb_create(const rb_netspec* spec, int device, rb_handle** out) {
  RB_CHECK(spec && out, RB_ERR_INVALID, "rb_create: null argument");
  RB_CHECK(spec->abi_version == RB_ABI_VERSION, RB_ERR_INVALID, "ABI version %d, library has %d",
           spec->abi_version, RB_ABI_VERSION);
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  RB_CHECK(e == cudaSuccess && ndev > 0, RB_ERR_CUDA,
           "no CUDA device: rave_b200 has no CPU fallback (%s)", cudaGetErrorString(e));
  RB_CHECK(device >= 0 && device < ndev, RB_ERR_INVALID, "device %d out of range", device);
  RB_CUDA(cudaSetDevice(device));
  RB_CHECK(spec->n_slots > 0 && spec->n_ops > 0 && spec->n_enc_ops <= spec->n_ops, RB_ERR_INVALID,
           "empty or inconsistent netspec");
  rb_handle* h = new rb_handle();
  h->device = device;
  h->spec = *spec;
  h->slots.assign(spec->slots, spec->slots + spec->n_slots);
  h->ops.assign(spec->ops, spec->ops + spec->n_ops);
  h->spec.slots = h->slots.data();
  h->spec.ops = h->ops.data();
  h->w.resize(h->ops.size());
  for (size_t i = 0; i < h->ops.size(); ++i) {
    const rb_op& op = h->ops[i];
    bool ok = op.src >= 0 && op.src < spec->n_slots && op.dst >= 0 && op.dst < spec->n_slots &&
              op.res < spec->n_slots && op.src2 < spec->n_slots && op.kind >= RB_OP_CONV &&
              op.kind <= RB_OP_CONCAT_COND;
    if (!ok) {
      delete h;
      set_error("op %zu is malformed", i);
      return RB_ERR_INVALID;
    }
  }
  *out = h;
  return RB_OK;
}

int rb_destroy(rb_handle* h) {
  if (!h) return RB_OK;
  cudaSetDevice(h->device);
  tc_free(h);
  for (auto& w : h->w) {
    cudaFree(w.w); cudaFree(w.bias); cudaFree(w.alpha); cudaFree(w.aff_a); cudaFree(w.aff_b); cudaFree(w.stats);
    cudaFree(w.cond);
  }
  cudaFree(h->f0);
  cudaFree(h->codebooks); cudaFree(h->cb_norms); cudaFree(h->rvq_ws);
  for (auto& p : h->pool) cudaFree(p.p);
  for (void* p : h->sbuf) cudaFree(p);
  if (h->stage_in) cudaFreeHost(h->stage_in);
  if (h->stage_out) cudaFreeHost(h->stage_out);
  cudaFree(h->dev_in); cudaFree(h->dev_out);
  if (h->pipe.ready) {
    cudaStreamSynchronize(h->pipe.s_out); cudaStreamSynchronize(h->pipe.s_cmp); cudaStreamSynchronize(h->pipe.s_in);
    for (int i = 0; i < rb_handle::HostPipe::NBUF; ++i) {
      cudaFree(h->pipe.din[i]); cudaFree(h->pipe.dout[i]);
      cudaEventDestroy(h->pipe.in_done[i]); cudaEventDestroy(h->pipe.cmp_done[i]); cudaEventDestroy(h->pipe.out_done[i]);
    }
    cudaEventDestroy(h->pipe.join_ev);
    cudaStreamDestroy(h->pipe.s_in); cudaStreamDestroy(h->pipe.s_cmp); cudaStreamDestroy(h->pipe.s_out);
  }
  delete h;
  return RB_OK;
}

int rb_load_weights(rb_handle* h, const rb_tensor_desc* t, int n) {
  RB_CHECK(h && (t || n == 0), RB_ERR_INVALID, "rb_load_weights: null argument");
  RB_CUDA(cudaSetDevice(h->device));
  for (int i = 0; i < n; ++i) {
    const rb_tensor_desc& d = t[i];
    RB_CHECK(d.data && d.numel > 0, RB_ERR_INVALID, "tensor %d is empty", i);
    if (d.role == RB_W_CODEBOOK) {
      const int64_t want = (int64_t)h->spec.rvq_quantizers * h->spec.rvq_codebook_size * h->spec.rvq_dim;
      RB_CHECK(want > 0 && d.numel == want, RB_ERR_INVALID, "codebooks: %lld elements, expected %lld",
               (long long)d.numel, (long long)want);
      int rc = upload(d.data, d.numel, &h->codebooks);
      if (rc) return rc;
      const int rows = h->spec.rvq_quantizers * h->spec.rvq_codebook_size;
      cudaFree(h->cb_norms);
      RB_CUDA(cudaMalloc((void**)&h->cb_norms, rows * sizeof(float)));
      rc = launch_rvq_norms(h->codebooks, rows, h->spec.rvq_dim, h->cb_norms, 0);
      if (rc) return rc;
      RB_CUDA(cudaDeviceSynchronize());
      continue;
    }
    RB_CHECK(d.op >= 0 && d.op < (int)h->ops.size(), RB_ERR_INVALID, "tensor %d: op %d out of range", i, d.op);
    const rb_op& op = h->ops[d.op];
    OpWeights& w = h->w[d.op];
    int rc = RB_OK;
    switch (d.role) {
      case RB_W_CONV: {
        if (op.kind == RB_OP_PQMF_FWD) {
          // analysis weight [M][1][k] -> [k][8] (bands kept, zero padded)
          RB_CHECK(d.numel == (int64_t)op.n_band * op.k, RB_ERR_INVALID, "pqmf analysis weight size");
          std::vector<float> p((size_t)op.k * 8, 0.f);
          for (int m = 0; m < op.c_out && m < 8; ++m)
            for (int j = 0; j < op.k; ++j) p[(size_t)j * 8 + m] = d.data[(size_t)m * op.k + j];
          rc = upload(p.data(), p.size(), &w.w);
          w.w_host.assign(d.data, d.data + d.numel);
        } else if (op.kind == RB_OP_PQMF_INV) {
          // synthesis weight [m][c][j] -> [c][j][m]
          const int M = op.n_band;
          RB_CHECK(d.numel == (int64_t)M * M * op.k, RB_ERR_INVALID, "pqmf synthesis weight size");
          std::vector<float> p((size_t)M * M * op.k);
          for (int m = 0; m < M; ++m)
            for (int c = 0; c < M; ++c)
              for (int j = 0; j < op.k; ++j)
                p[((size_t)c * op.k + j) * M + m] = d.data[((size_t)m * M + c) * op.k + j];
          rc = upload(p.data(), p.size(), &w.w);
          w.w_host.assign(d.data, d.data + d.numel);
        } else {
          RB_CHECK(op.kind == RB_OP_CONV, RB_ERR_INVALID, "op %d takes no conv weight", d.op);
          RB_CHECK(d.numel == (int64_t)op.c_out * op.c_in * op.k, RB_ERR_INVALID,
                   "op %d: conv weight has %lld elements, expected %d*%d*%d", d.op, (long long)d.numel,
                   op.c_out, op.c_in, op.k);
          std::vector<float> p((size_t)d.numel);
          for (int co = 0; co < op.c_out; ++co)
            for (int ci = 0; ci < op.c_in; ++ci)
              for (int j = 0; j < op.k; ++j)
                p[((size_t)ci * op.k + j) * op.c_out + co] = d.data[((size_t)co * op.c_in + ci) * op.k + j];
          rc = upload(p.data(), p.size(), &w.w);
          w.w_host.assign(d.data, d.data + d.numel);
        }
        break;
      }
      case RB_W_BIAS:
        RB_CHECK(d.numel == op.c_out / std::max(op.ileave, 1), RB_ERR_INVALID, "op %d: bias size", d.op);
        rc = upload(d.data, d.numel, &w.bias);
        w.bias_host.assign(d.data, d.data + d.numel);
        break;
      case RB_W_ALPHA:
        RB_CHECK(d.numel == op.c_in, RB_ERR_INVALID, "op %d: alpha size", d.op);
        rc = upload(d.data, d.numel, &w.alpha);
        w.alpha_host.assign(d.data, d.data + d.numel);
        break;
      case RB_W_AFFINE_A:
      case RB_W_AFFINE_B: {
        RB_CHECK(op.kind == RB_OP_AFFINE && d.numel % op.c_in == 0, RB_ERR_INVALID, "op %d: affine size", d.op);
        w.aff_rows = (int)(d.numel / op.c_in);
        rc = upload(d.data, d.numel, d.role == RB_W_AFFINE_A ? &w.aff_a : &w.aff_b);
        break;
      }
      case RB_W_ADAIN_STATS: {
        // [4][rows][C] (mean_x, std_x, mean_y, std_y) followed by num_update_x, num_update_y
        RB_CHECK(op.kind == RB_OP_AFFINE && d.numel > 2 && (d.numel - 2) % (4 * op.c_in) == 0, RB_ERR_INVALID,
                 "op %d: AdaIN statistics size", d.op);
        w.aff_rows = (int)((d.numel - 2) / (4 * op.c_in));
        const int64_t n = (int64_t)w.aff_rows * op.c_in;
        if ((rc = upload(d.data, 4 * n, &w.stats))) break;
        w.nx = (int)d.data[4 * n];
        w.ny = (int)d.data[4 * n + 1];
        w.touch_x = w.touch_y = false;
        std::vector<float> zero((size_t)n, 0.f);
        if ((rc = upload(zero.data(), n, &w.aff_a))) break;
        if ((rc = upload(zero.data(), n, &w.aff_b))) break;
        rc = launch_adain_ab(w.stats, w.aff_rows, op.c_in, (w.nx > 0 && w.ny > 0) ? 1 : 0, w.aff_a, w.aff_b, 0);
        if (!rc) RB_CUDA(cudaStreamSynchronize(0));
        break;
      }
      case RB_W_COND:
        RB_CHECK(op.kind == RB_OP_CONCAT_COND && d.numel == h->spec.cond_channels, RB_ERR_INVALID,
                 "conditioning vector size");
        rc = upload(d.data, d.numel, &w.cond);
        w.cond_rows = 1;
        ++h->cond_version;
        break;
      default:
        set_error("tensor %d: unknown role %d", i, d.role);
        return RB_ERR_INVALID;
    }
    if (rc) return rc;
  }
  // every op that needs a weight must have one before the model can run
  bool all = true;
  for (size_t i = 0; i < h->ops.size(); ++i) {
    const rb_op& op = h->ops[i];
    if ((op.kind == RB_OP_CONV || op.kind == RB_OP_PQMF_FWD || op.kind == RB_OP_PQMF_INV) && !h->w[i].w) all = false;
    if (op.kind == RB_OP_CONV && op.has_bias && !h->w[i].bias) all = false;
  }
  h->weights_loaded = all;
  if (all) {
    int rc = tc_pack_weights(h);
    if (rc) return rc;
  }
  return RB_OK;
}

static int check_ready(rb_handle* h) {
  RB_CHECK(h, RB_ERR_INVALID, "null handle");
  RB_CHECK(h->weights_loaded, RB_ERR_STATE, "weights are not (fully) loaded");
  RB_CUDA(cudaSetDevice(h->device));
  return RB_OK;
}

static int run_any(rb_handle* h, int ob, int oe, int in_slot, int out_slot, const void* in, void* out,
                   int B, int F, int dtype, cudaStream_t s) {
  const long long before = g_launches;
  int rc;
  if (dtype == RB_F32)
    rc = run_offline_f32(h, ob, oe, in_slot, out_slot, (const float*)in, (float*)out, B, F, s);
  else if (dtype == RB_BF16)
    rc = tc_run_program(h, ob, oe, in, out, B, F, s);
  else {
    set_error("unknown dtype %d", dtype);
    rc = RB_ERR_INVALID;
  }
  h->launches += g_launches - before;
  return rc;
}

int rb_encode(rb_handle* h, const void* x, int B, int T, int dtype, void* z, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  RB_CHECK(x && z && B > 0 && T > 0, RB_ERR_INVALID, "rb_encode: bad arguments");
  RB_CHECK(T % h->spec.samples_per_frame == 0, RB_ERR_INVALID,
           "T=%d is not a multiple of the model ratio %d", T, h->spec.samples_per_frame);
  RB_CHECK(!h->spec.streaming, RB_ERR_STATE, "model was built with streaming=True: use rb_stream_*");
  BusyGuard g(h);
  RB_CHECK(g.ok, RB_ERR_STATE, "handle is busy on another thread");
  return run_any(h, 0, h->spec.n_enc_ops, h->spec.enc_in_slot, h->spec.enc_out_slot, x, z, B,
                 T / h->spec.samples_per_frame, dtype, (cudaStream_t)stream);
}

int rb_decode(rb_handle* h, const void* z, int B, int F, int dtype, void* y, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  RB_CHECK(z && y && B > 0 && F > 0, RB_ERR_INVALID, "rb_decode: bad arguments");
  RB_CHECK(!h->spec.streaming, RB_ERR_STATE, "model was built with streaming=True: use rb_stream_*");
  BusyGuard g(h);
  RB_CHECK(g.ok, RB_ERR_STATE, "handle is busy on another thread");
  return run_any(h, h->spec.n_enc_ops, h->spec.n_ops, h->spec.dec_in_slot, h->spec.dec_out_slot, z, y, B,
                 F, dtype, (cudaStream_t)stream);
}

int rb_forward(rb_handle* h, const void* x, int B, int T, int dtype, void* y, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  RB_CHECK(x && y && B > 0 && T > 0, RB_ERR_INVALID, "rb_forward: bad arguments");
  RB_CHECK(T % h->spec.samples_per_frame == 0, RB_ERR_INVALID,
           "T=%d is not a multiple of the model ratio %d", T, h->spec.samples_per_frame);
  RB_CHECK(!h->spec.streaming, RB_ERR_STATE, "model was built with streaming=True: use rb_stream_*");
  const rb_slot& zo = h->slots[h->spec.enc_out_slot];
  const rb_slot& zi = h->slots[h->spec.dec_in_slot];
  RB_CHECK(zo.channels == zi.channels, RB_ERR_INVALID,
           "encode yields %d channels but the decoder takes %d (model.py:620 vs core.py:78)",
           zo.channels, zi.channels);
  BusyGuard g(h);
  RB_CHECK(g.ok, RB_ERR_STATE, "handle is busy on another thread");
  const int F = T / h->spec.samples_per_frame;
  if (dtype == RB_BF16)   // one program: the latent stays on the device in the engine's layout
    return run_any(h, 0, h->spec.n_ops, h->spec.enc_in_slot, h->spec.dec_out_slot, x, y, B, F, dtype,
                   (cudaStream_t)stream);
  const size_t esz = dtype == RB_BF16 ? 2 : 4;
  // latent scratch lives in a dedicated buffer (small)
  const size_t zbytes = (size_t)B * zo.channels * ((F + 3) & ~3) * esz;
  if (h->rvq_ws_bytes < zbytes) {
    cudaFree(h->rvq_ws);
    h->rvq_ws = nullptr;
    RB_CUDA(cudaMalloc((void**)&h->rvq_ws, zbytes));
    h->rvq_ws_bytes = zbytes;
  }
  rc = run_any(h, 0, h->spec.n_enc_ops, h->spec.enc_in_slot, h->spec.enc_out_slot, x, h->rvq_ws, B, F,
               dtype, (cudaStream_t)stream);
  if (rc) return rc;
  return run_any(h, h->spec.n_enc_ops, h->spec.n_ops, h->spec.dec_in_slot, h->spec.dec_out_slot,
                 h->rvq_ws, y, B, F, dtype, (cudaStream_t)stream);
}

int rb_forward_host(rb_handle* h, const float* x, int B, int T, int dtype, float* y) {
  int rc = check_ready(h);
  if (rc) return rc;
  RB_CHECK(x && y && B > 0 && T > 0, RB_ERR_INVALID, "rb_forward_host: bad arguments");
  const size_t n = (size_t)B * T;
  if (h->stage_elems < n) {
    if (h->stage_in) cudaFreeHost(h->stage_in);
    if (h->stage_out) cudaFreeHost(h->stage_out);
    cudaFree(h->dev_in); cudaFree(h->dev_out);
    h->stage_in = h->stage_out = nullptr; h->dev_in = h->dev_out = nullptr; h->stage_elems = 0;
    RB_CUDA(cudaMallocHost((void**)&h->stage_in, n * sizeof(float)));
    RB_CUDA(cudaMallocHost((void**)&h->stage_out, n * sizeof(float)));
    RB_CUDA(cudaMalloc(&h->dev_in, n * sizeof(float)));
    RB_CUDA(cudaMalloc(&h->dev_out, n * sizeof(float)));
    h->stage_elems = n;
  }
  RB_CHECK(dtype == RB_F32, RB_ERR_UNSUPPORTED, "rb_forward_host takes fp32 host buffers; for the bf16 "
           "engine convert on the device (rave_b200.RAVE does)");
  memcpy(h->stage_in, x, n * sizeof(float));
  RB_CUDA(cudaMemcpyAsync(h->dev_in, h->stage_in, n * sizeof(float), cudaMemcpyHostToDevice, 0));
  rc = rb_forward(h, h->dev_in, B, T, dtype, h->dev_out, nullptr);
  if (rc) return rc;
  RB_CUDA(cudaMemcpyAsync(h->stage_out, h->dev_out, n * sizeof(float), cudaMemcpyDeviceToHost, 0));
  RB_CUDA(cudaStreamSynchronize(0));
  memcpy(y, h->stage_out, n * sizeof(float));
  return RB_OK;
}

// Block k of a call: H2D on s_in -> forward on s_cmp -> D2H on s_out.  Ring slot i = seq % NBUF:
// its input block may be overwritten once the forward that read it is done (cmp_done[i]), its
// output block once the copy that drained it is done (out_done[i]).  Waiting on an event that
// was never recorded is a no-op, which covers the first lap.
int rb_forward_host_async(rb_handle* h, const void* x, int B, int T, int dtype, void* y, int n_chunks) {
  int rc = check_ready(h);
  if (rc) return rc;
  RB_CHECK(x && y && B > 0 && T > 0 && n_chunks > 0, RB_ERR_INVALID, "rb_forward_host_async: bad arguments");
  RB_CHECK(B % n_chunks == 0, RB_ERR_INVALID, "B=%d is not a multiple of n_chunks=%d", B, n_chunks);
  RB_CHECK(dtype == RB_F32 || dtype == RB_BF16, RB_ERR_INVALID, "unknown dtype %d", dtype);
  RB_CUDA(cudaSetDevice(h->device));
  auto& pp = h->pipe;
  constexpr int NBUF = rb_handle::HostPipe::NBUF;
  if (!pp.ready) {
    RB_CUDA(cudaStreamCreateWithFlags(&pp.s_in, cudaStreamNonBlocking));
    RB_CUDA(cudaStreamCreateWithFlags(&pp.s_cmp, cudaStreamNonBlocking));
    RB_CUDA(cudaStreamCreateWithFlags(&pp.s_out, cudaStreamNonBlocking));
    for (int i = 0; i < NBUF; ++i) {
      RB_CUDA(cudaEventCreateWithFlags(&pp.in_done[i], cudaEventDisableTiming));
      RB_CUDA(cudaEventCreateWithFlags(&pp.cmp_done[i], cudaEventDisableTiming));
      RB_CUDA(cudaEventCreateWithFlags(&pp.out_done[i], cudaEventDisableTiming));
    }
    RB_CUDA(cudaEventCreateWithFlags(&pp.join_ev, cudaEventDisableTiming));
    pp.ready = true;
  }
  const int Bc = B / n_chunks;
  const size_t bytes = (size_t)Bc * T * (dtype == RB_BF16 ? 2 : 4);
  if (pp.block_bytes < bytes) {
    RB_CUDA(cudaStreamSynchronize(pp.s_out)); RB_CUDA(cudaStreamSynchronize(pp.s_cmp)); RB_CUDA(cudaStreamSynchronize(pp.s_in));
    for (int i = 0; i < NBUF; ++i) {
      cudaFree(pp.din[i]); cudaFree(pp.dout[i]);
      pp.din[i] = pp.dout[i] = nullptr;
    }
    pp.block_bytes = 0;
    for (int i = 0; i < NBUF; ++i) {
      RB_CUDA(cudaMalloc(&pp.din[i], bytes));
      RB_CUDA(cudaMalloc(&pp.dout[i], bytes));
    }
    pp.block_bytes = bytes;
  }
  for (int k = 0; k < n_chunks; ++k, ++pp.seq) {
    const int i = (int)(pp.seq % NBUF);
    RB_CUDA(cudaStreamWaitEvent(pp.s_in, pp.cmp_done[i], 0));
    RB_CUDA(cudaMemcpyAsync(pp.din[i], (const char*)x + (size_t)k * bytes, bytes, cudaMemcpyHostToDevice, pp.s_in));
    RB_CUDA(cudaEventRecord(pp.in_done[i], pp.s_in));
    RB_CUDA(cudaStreamWaitEvent(pp.s_cmp, pp.in_done[i], 0));
    RB_CUDA(cudaStreamWaitEvent(pp.s_cmp, pp.out_done[i], 0));
    rc = rb_forward(h, pp.din[i], Bc, T, dtype, pp.dout[i], pp.s_cmp);
    if (rc) return rc;
    RB_CUDA(cudaEventRecord(pp.cmp_done[i], pp.s_cmp));
    RB_CUDA(cudaStreamWaitEvent(pp.s_out, pp.cmp_done[i], 0));
    RB_CUDA(cudaMemcpyAsync((char*)y + (size_t)k * bytes, pp.dout[i], bytes, cudaMemcpyDeviceToHost, pp.s_out));
    RB_CUDA(cudaEventRecord(pp.out_done[i], pp.s_out));
  }
  return RB_OK;
}

int rb_host_join(rb_handle* h, void* stream) {
  RB_CHECK(h, RB_ERR_INVALID, "null handle");
  if (!h->pipe.ready) return RB_OK;
  RB_CUDA(cudaEventRecord(h->pipe.join_ev, h->pipe.s_out));   // s_out is ordered after s_cmp after s_in
  RB_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, h->pipe.join_ev, 0));
  return RB_OK;
}

int rb_host_sync(rb_handle* h) {
  RB_CHECK(h, RB_ERR_INVALID, "null handle");
  if (!h->pipe.ready) return RB_OK;
  RB_CUDA(cudaStreamSynchronize(h->pipe.s_out));
  return RB_OK;
}

// model.py:615-620 / scripts/export.py:381-397 (see include/rave_b200.h)
int rb_set_conditioning(rb_handle* h, const float* spk, int spk_rows, const float* f0, int B, int f0_channels,
                        int F, void* stream) {
  RB_CHECK(h, RB_ERR_INVALID, "null handle");
  BusyGuard g(h);
  RB_CHECK(g.ok, RB_ERR_STATE, "handle is busy");
  cudaStream_t s = (cudaStream_t)stream;
  int oi = -1;
  for (size_t i = 0; i < h->ops.size(); ++i)
    if (h->ops[i].kind == RB_OP_CONCAT_COND) oi = (int)i;
  RB_CHECK(oi >= 0, RB_ERR_INVALID, "the net has no conditioning channels");
  RB_CUDA(cudaSetDevice(h->device));
  if (spk) {
    OpWeights& w = h->w[oi];
    RB_CHECK(spk_rows >= 1, RB_ERR_INVALID, "spk_rows must be >= 1");
    const size_t n = (size_t)spk_rows * h->spec.cond_channels;
    if (spk_rows != w.cond_rows || !w.cond) {
      if (w.cond) { RB_CUDA(cudaStreamSynchronize(s)); cudaFree(w.cond); w.cond = nullptr; }
      RB_CUDA(cudaMalloc((void**)&w.cond, n * sizeof(float)));
      w.cond_rows = spk_rows;
    }
    RB_CUDA(cudaMemcpyAsync(w.cond, spk, n * sizeof(float), cudaMemcpyDeviceToDevice, s));
    ++h->cond_version;
  }
  if (f0) {
    RB_CHECK(B >= 1 && f0_channels >= 1 && F >= 1, RB_ERR_INVALID, "f0 must be [B, f0_channels, F]");
    const size_t n = (size_t)B * f0_channels * F;
    if (n > h->f0_cap) {
      if (h->f0) { RB_CUDA(cudaStreamSynchronize(s)); cudaFree(h->f0); h->f0 = nullptr; h->f0_cap = 0; }
      RB_CUDA(cudaMalloc((void**)&h->f0, n * sizeof(float)));
      h->f0_cap = n;
    }
    RB_CUDA(cudaMemcpyAsync(h->f0, f0, n * sizeof(float), cudaMemcpyDeviceToDevice, s));
    h->f0_B = B; h->f0_C = f0_channels; h->f0_F = F;
  } else if (B < 0) {
    h->f0_B = h->f0_C = h->f0_F = 0;
  }
  return RB_OK;
}

int rb_set_noise(rb_handle* h, const float* noise, int64_t numel) {
  RB_CHECK(h, RB_ERR_INVALID, "null handle");
  h->noise = noise;
  h->noise_numel = noise ? numel : 0;
  return RB_OK;
}

int rb_profile_ops(rb_handle* h, int enable) {
  RB_CHECK(h, RB_ERR_INVALID, "null handle");
  h->profile = enable != 0;
  return RB_OK;
}

int rb_op_times(rb_handle* h, float* ms, int n) {
  RB_CHECK(h && ms && n == (int)h->ops.size(), RB_ERR_INVALID, "rb_op_times: n must equal the op count");
  for (int i = 0; i < n; ++i) ms[i] = 0.f;
  for (auto& e : h->events) {
    float t = 0.f;
    RB_CUDA(cudaEventSynchronize(e.b));
    RB_CUDA(cudaEventElapsedTime(&t, e.a, e.b));
    ms[e.op] += t;
    cudaEventDestroy(e.a);
    cudaEventDestroy(e.b);
  }
  h->events.clear();
  return RB_OK;
}

// ScriptedRAVE.update_adain (scripts/export.py:248-265): learn_source -> learn_x, learn_target -> learn_y
// of every AdaIN site; resets as AdaptiveInstanceNormalization.reset_x / reset_y (blocks.py:875-883).
int rb_adain_update(rb_handle* h, int learn_source, int learn_target, int reset_source, int reset_target) {
  RB_CHECK(h, RB_ERR_INVALID, "rb_adain_update: null handle");
  RB_CUDA(cudaSetDevice(h->device));
  h->adain_learn_x = learn_source != 0;
  h->adain_learn_y = learn_target != 0;
  for (size_t i = 0; i < h->w.size(); ++i) {
    OpWeights& w = h->w[i];
    if (h->ops[i].kind != RB_OP_AFFINE) continue;
    RB_CHECK(w.stats || !(learn_source || learn_target || reset_source || reset_target), RB_ERR_STATE,
             "op %d: AdaIN statistics were loaded as fixed a / b tables (RB_W_ADAIN_STATS is needed to learn)", (int)i);
    if (!w.stats) continue;
    const int C = h->ops[i].c_in;
    const size_t n = (size_t)w.aff_rows * C;
    std::vector<float> zeros(n, 0.f), ones(n, 1.f);
    if (reset_source) {
      RB_CUDA(cudaMemcpy(w.stats, zeros.data(), n * 4, cudaMemcpyHostToDevice));
      RB_CUDA(cudaMemcpy(w.stats + n, ones.data(), n * 4, cudaMemcpyHostToDevice));
      w.nx = 0;
    }
    if (reset_target) {
      RB_CUDA(cudaMemcpy(w.stats + 2 * n, zeros.data(), n * 4, cudaMemcpyHostToDevice));
      RB_CUDA(cudaMemcpy(w.stats + 3 * n, ones.data(), n * 4, cudaMemcpyHostToDevice));
      w.ny = 0;
    }
    if (reset_source || reset_target) {
      int rc = launch_adain_ab(w.stats, w.aff_rows, C, (w.nx > 0 && w.ny > 0) ? 1 : 0, w.aff_a, w.aff_b, 0);
      if (rc) return rc;
      RB_CUDA(cudaStreamSynchronize(0));
    }
  }
  return RB_OK;
}

// Reads back the statistics of AdaIN op `op` (host buffer of 4 * rows * C floats) and its counters.
int rb_adain_get_stats(rb_handle* h, int op, float* stats, int64_t numel, int* num_update_x, int* num_update_y) {
  RB_CHECK(h && op >= 0 && op < (int)h->ops.size() && h->ops[op].kind == RB_OP_AFFINE, RB_ERR_INVALID,
           "rb_adain_get_stats: op %d is not an AdaIN site", op);
  const OpWeights& w = h->w[op];
  RB_CHECK(w.stats, RB_ERR_STATE, "op %d: no AdaIN statistics loaded", op);
  const int64_t n = 4ll * w.aff_rows * h->ops[op].c_in;
  RB_CHECK(!stats || numel == n, RB_ERR_INVALID, "rb_adain_get_stats: buffer of %lld floats, need %lld",
           (long long)numel, (long long)n);
  RB_CUDA(cudaSetDevice(h->device));
  RB_CUDA(cudaDeviceSynchronize());
  if (stats) RB_CUDA(cudaMemcpy(stats, w.stats, (size_t)n * 4, cudaMemcpyDeviceToHost));
  if (num_update_x) *num_update_x = w.nx;
  if (num_update_y) *num_update_y = w.ny;
  return RB_OK;
}

int rb_set_microbatch(rb_handle* h, int streams) {
  RB_CHECK(h && streams >= 0, RB_ERR_INVALID, "rb_set_microbatch: bad arguments");
  h->microbatch = streams;
  return RB_OK;
}

int rb_kernel_launches(rb_handle* h, int64_t* count) {
  RB_CHECK(h && count, RB_ERR_INVALID, "null argument");
  *count = h->launches;
  return RB_OK;
}

int rb_workspace_bytes(rb_handle* h, int B, int T, int dtype, size_t* bytes) {
  RB_CHECK(h && bytes && B > 0 && T > 0, RB_ERR_INVALID, "rb_workspace_bytes: bad arguments");
  RB_CHECK(T % h->spec.samples_per_frame == 0, RB_ERR_INVALID, "T is not a multiple of the model ratio");
  const int F = T / h->spec.samples_per_frame;
  size_t mx = 0;
  for (size_t s = 0; s < h->slots.size(); ++s) mx = std::max(mx, slot_bytes_f32(h, (int)s, 1, F));
  if (dtype == RB_BF16) mx /= 2;
  // at most four activations are live at once (x, conv3 out, residual, dst)
  *bytes = 4 * mx * (size_t)std::min(B, std::max(1, h->microbatch > 0 ? h->microbatch : B));
  return RB_OK;
}

// ------------------------------------------------------------------------------ RVQ
int rb_rvq_encode(rb_handle* h, const float* z, int B, int F, int64_t* idx, void* stream) {
  RB_CHECK(h && z && idx && B > 0 && F > 0, RB_ERR_INVALID, "rb_rvq_encode: bad arguments");
  RB_CHECK(h->codebooks, RB_ERR_STATE, "RVQ codebooks are not loaded (not a discrete model?)");
  RB_CUDA(cudaSetDevice(h->device));
  const size_t need = (size_t)B * F * h->spec.rvq_dim * sizeof(float);
  if (h->rvq_ws_bytes < need) {
    cudaFree(h->rvq_ws);
    h->rvq_ws = nullptr;
    RB_CUDA(cudaMalloc((void**)&h->rvq_ws, need));
    h->rvq_ws_bytes = need;
  }
  const long long before = g_launches;
  int rc = launch_rvq_encode(z, B, h->spec.rvq_dim, F, h->codebooks, h->cb_norms, h->spec.rvq_quantizers,
                             h->spec.rvq_codebook_size, idx, h->rvq_ws, (cudaStream_t)stream);
  h->launches += g_launches - before;
  return rc;
}

int rb_rvq_decode(rb_handle* h, const int64_t* idx, int B, int F, float* zq, void* stream) {
  RB_CHECK(h && zq && idx && B > 0 && F > 0, RB_ERR_INVALID, "rb_rvq_decode: bad arguments");
  RB_CHECK(h->codebooks, RB_ERR_STATE, "RVQ codebooks are not loaded (not a discrete model?)");
  RB_CUDA(cudaSetDevice(h->device));
  const long long before = g_launches;
  int rc = launch_rvq_decode(idx, B, h->spec.rvq_dim, F, h->codebooks, h->spec.rvq_quantizers,
                             h->spec.rvq_codebook_size, zq, (cudaStream_t)stream);
  h->launches += g_launches - before;
  return rc;
}

}  // extern "C"
