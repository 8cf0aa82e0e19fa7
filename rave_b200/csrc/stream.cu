// stream.cu -- streaming entry points (cached_conv semantics, scripts/export.py:542-594).
//
// Every activation slot owns one persistent buffer [max_streams][C][hist4 + max_n].  The
// first `hist` columns of a row are the left context the consumers of that slot need
// (CachedPadding1d state of each conv, the AlignBranches delay line of a Residual, the
// one-sample overlap of a transposed conv).  A step writes the new samples behind the
// context, runs the same fused kernels as the offline path with `lo = -hist`, then rolls the
// tail of each row to the front: every state element is read once and written once per step.
#include <algorithm>

#include "engine.h"

using namespace rb;

namespace rb {
// bf16 streaming on the tensor-core engine (tc_engine.cu)
int tc_stream_alloc(rb_handle* h);
int tc_stream_reset(rb_handle* h, const int32_t* ids, int n);
int tc_stream_run(rb_handle* h, int ob, int oe, int B, int F, const void* in, void* out, cudaStream_t s);
int tc_stream_handoff(rb_handle* h, int B, int F, cudaStream_t s);
}  // namespace rb

namespace {

int slot_view(rb_handle* h, int sl, int F, View* v) {
  const rb_slot& s = h->slots[sl];
  v->p = (float*)h->sbuf[sl];
  v->C = s.channels;
  v->ld = h->s_ld[sl];
  v->hist = h->s_hist[sl];
  v->n = s.per_frame * F;
  return RB_OK;
}

int run_stream_ops(rb_handle* h, int ob, int oe, int B, int F, int out_slot, float* out, cudaStream_t s) {
  const int ns = (int)h->slots.size();
  RunCtx c{h, B, F, std::vector<View>(ns), s, true, nullptr, 0};
  for (int sl = 0; sl < ns; ++sl) slot_view(h, sl, F, &c.v[sl]);
  if (out_slot >= 0 && out && h->s_hist[out_slot] == 0) {   // write the result straight to the caller
    c.v[out_slot].p = out;
    c.v[out_slot].ld = c.v[out_slot].n;
  }
  if (h->spec.noise_slot >= 0 && h->noise) {
    const rb_slot& nsl = h->slots[h->spec.noise_slot];
    c.noise_bstride = (int64_t)nsl.channels * nsl.per_frame * F;
    RB_CHECK(h->noise_numel >= (int64_t)B * c.noise_bstride, RB_ERR_INVALID, "noise tensor too small");
    c.noise = h->noise;
  }
  for (int i = ob; i < oe; ++i) {
    int rc = exec_op_f32(c, i);
    if (rc) return rc;
  }
  // state update for every slot this program reads with history
  std::vector<char> seen(ns, 0);
  for (int i = ob; i < oe; ++i) {
    const rb_op& op = h->ops[i];
    const int used[3] = {op.src, op.res, op.src2};
    for (int u : used) {
      if (u < 0 || seen[u] || h->s_hist[u] == 0 || u == h->spec.noise_slot) continue;
      seen[u] = 1;
      const View& v = c.v[u];
      // rows of streams [0, B) are contiguous: B*C rows of pitch ld
      int rc = launch_roll_f32(v.p + (v.hist - h->slots[u].hist), B * v.C, v.ld, h->slots[u].hist, v.n, s);
      if (rc) return rc;
    }
  }
  adain_finalize(h, ob, oe);
  return RB_OK;
}

int copy_in(rb_handle* h, int sl, const float* src, int B, int F, cudaStream_t s) {
  View v;
  slot_view(h, sl, F, &v);
  return launch_copy_rows_f32(src, v.n, v.p + v.hist, v.ld, B * v.C, v.n, s);
}

int stream_check(rb_handle* h, int B, int n_frames) {
  RB_CHECK(h, RB_ERR_INVALID, "null handle");
  RB_CHECK(h->weights_loaded, RB_ERR_STATE, "weights are not (fully) loaded");
  RB_CHECK(h->spec.streaming, RB_ERR_STATE, "model was not built with streaming=True");
  RB_CHECK(h->stream_ready, RB_ERR_STATE, "call rb_stream_alloc first");
  RB_CHECK(B > 0 && B <= h->max_streams, RB_ERR_INVALID, "B=%d exceeds the %d allocated streams", B,
           h->max_streams);
  RB_CHECK(n_frames > 0 && n_frames <= h->max_frames, RB_ERR_INVALID,
           "chunk of %d frames exceeds the allocated %d", n_frames, h->max_frames);
  RB_CUDA(cudaSetDevice(h->device));
  return RB_OK;
}

struct Busy {
  rb_handle* h;
  bool ok;
  explicit Busy(rb_handle* h_) : h(h_), ok(!h_->busy.test_and_set()) {}
  ~Busy() { if (ok) h->busy.clear(); }
};

}  // namespace

extern "C" {

int rb_stream_alloc(rb_handle* h, int max_streams, int max_chunk, int dtype) {
  RB_CHECK(h && max_streams > 0 && max_chunk > 0, RB_ERR_INVALID, "rb_stream_alloc: bad arguments");
  RB_CHECK(h->spec.streaming, RB_ERR_STATE, "model was not built with streaming=True");
  RB_CHECK(dtype == RB_F32 || dtype == RB_BF16, RB_ERR_INVALID, "rb_stream_alloc: bad dtype");
  RB_CHECK(max_chunk % h->spec.samples_per_frame == 0, RB_ERR_INVALID,
           "chunk %d is not a multiple of the model ratio %d", max_chunk, h->spec.samples_per_frame);
  RB_CUDA(cudaSetDevice(h->device));
  // not ready until every buffer of the new arena exists: a failed (re)allocation must not leave
  // rb_stream_* running on freed or partially allocated state
  h->stream_ready = false;
  RB_CUDA(cudaDeviceSynchronize());               // no step of the old arena may still be in flight
  for (void* p : h->sbuf) cudaFree(p);
  const int ns = (int)h->slots.size();
  h->sbuf.assign(ns, nullptr);
  h->s_ld.assign(ns, 0);
  h->s_hist.assign(ns, 0);
  h->max_streams = max_streams;
  h->max_frames = max_chunk / h->spec.samples_per_frame;
  h->stream_dtype = dtype;
  auto release = [&]() {
    for (void*& p : h->sbuf) { cudaFree(p); p = nullptr; }
  };
  if (dtype == RB_BF16) {
    int rc = tc_stream_alloc(h);
    if (rc) { release(); return rc; }
    h->stream_ready = true;
    return RB_OK;
  }
  for (int sl = 0; sl < ns; ++sl) {
    const rb_slot& s = h->slots[sl];
    if (sl == h->spec.noise_slot) continue;
    h->s_hist[sl] = round4(s.hist);
    h->s_ld[sl] = h->s_hist[sl] + round4(s.per_frame * h->max_frames);
    const size_t bytes = (size_t)max_streams * s.channels * h->s_ld[sl] * sizeof(float);
    cudaError_t e = cudaMalloc(&h->sbuf[sl], bytes);
    if (e == cudaSuccess) e = cudaMemset(h->sbuf[sl], 0, bytes);
    if (e != cudaSuccess) {
      release();
      set_error("rb_stream_alloc: %s (%zu bytes for slot %d)", cudaGetErrorString(e), bytes, sl);
      return RB_ERR_CUDA;
    }
  }
  h->stream_ready = true;
  return RB_OK;
}

int rb_stream_reset(rb_handle* h, const int32_t* ids, int n) {
  RB_CHECK(h && h->stream_ready, RB_ERR_STATE, "streams are not allocated");
  RB_CUDA(cudaSetDevice(h->device));
  // the memsets below run on the legacy stream, which does not order against cudaStreamNonBlocking
  // caller streams: drain the device so a reset can never race with a step still in flight
  RB_CUDA(cudaDeviceSynchronize());
  if (h->stream_dtype == RB_BF16) return tc_stream_reset(h, ids, n);
  for (size_t sl = 0; sl < h->slots.size(); ++sl) {
    if (!h->sbuf[sl]) continue;
    const size_t row = (size_t)h->slots[sl].channels * h->s_ld[sl] * sizeof(float);
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

int rb_stream_encode(rb_handle* h, const void* x, int B, int n, void* z, void* stream) {
  RB_CHECK(h && x && z, RB_ERR_INVALID, "rb_stream_encode: null argument");
  RB_CHECK(n > 0 && n % h->spec.samples_per_frame == 0, RB_ERR_INVALID,
           "chunk %d is not a multiple of the model ratio %d", n, h->spec.samples_per_frame);
  const int F = n / h->spec.samples_per_frame;
  int rc = stream_check(h, B, F);
  if (rc) return rc;
  Busy g(h);
  RB_CHECK(g.ok, RB_ERR_STATE, "rb_stream_* calls on one handle must be serialised");
  const long long before = g_launches;
  cudaStream_t s = (cudaStream_t)stream;
  if (h->stream_dtype == RB_BF16) {
    rc = tc_stream_run(h, 0, h->spec.n_enc_ops, B, F, x, z, s);
  } else {
    rc = copy_in(h, h->spec.enc_in_slot, (const float*)x, B, F, s);
    if (!rc) rc = run_stream_ops(h, 0, h->spec.n_enc_ops, B, F, h->spec.enc_out_slot, (float*)z, s);
  }
  h->launches += g_launches - before;
  return rc;
}

int rb_stream_decode(rb_handle* h, const void* z, int B, int frames, void* y, void* stream) {
  RB_CHECK(h && z && y, RB_ERR_INVALID, "rb_stream_decode: null argument");
  int rc = stream_check(h, B, frames);
  if (rc) return rc;
  Busy g(h);
  RB_CHECK(g.ok, RB_ERR_STATE, "rb_stream_* calls on one handle must be serialised");
  const long long before = g_launches;
  cudaStream_t s = (cudaStream_t)stream;
  if (h->stream_dtype == RB_BF16) {
    rc = tc_stream_run(h, h->spec.n_enc_ops, h->spec.n_ops, B, frames, z, y, s);
  } else {
    rc = copy_in(h, h->spec.dec_in_slot, (const float*)z, B, frames, s);
    if (!rc) rc = run_stream_ops(h, h->spec.n_enc_ops, h->spec.n_ops, B, frames, h->spec.dec_out_slot, (float*)y, s);
  }
  h->launches += g_launches - before;
  return rc;
}

int rb_stream_step(rb_handle* h, const void* x, int B, int n, void* y, void* stream) {
  RB_CHECK(h && x && y, RB_ERR_INVALID, "rb_stream_step: null argument");
  RB_CHECK(n > 0 && n % h->spec.samples_per_frame == 0, RB_ERR_INVALID,
           "chunk %d is not a multiple of the model ratio %d", n, h->spec.samples_per_frame);
  const int F = n / h->spec.samples_per_frame;
  int rc = stream_check(h, B, F);
  if (rc) return rc;
  const rb_slot& zo = h->slots[h->spec.enc_out_slot];
  const rb_slot& zi = h->slots[h->spec.dec_in_slot];
  RB_CHECK(zo.channels == zi.channels, RB_ERR_INVALID, "encode yields %d channels, decoder takes %d",
           zo.channels, zi.channels);
  Busy g(h);
  RB_CHECK(g.ok, RB_ERR_STATE, "rb_stream_* calls on one handle must be serialised");
  const long long before = g_launches;
  cudaStream_t s = (cudaStream_t)stream;
  if (h->stream_dtype == RB_BF16) {
    rc = tc_stream_run(h, 0, h->spec.n_enc_ops, B, F, x, nullptr, s);
    if (!rc) rc = tc_stream_handoff(h, B, F, s);
    if (!rc) rc = tc_stream_run(h, h->spec.n_enc_ops, h->spec.n_ops, B, F, nullptr, y, s);
    h->launches += g_launches - before;
    return rc;
  }
  rc = copy_in(h, h->spec.enc_in_slot, (const float*)x, B, F, s);
  if (!rc) rc = run_stream_ops(h, 0, h->spec.n_enc_ops, B, F, -1, nullptr, s);
  if (!rc) {
    View zv, dv;
    slot_view(h, h->spec.enc_out_slot, F, &zv);
    slot_view(h, h->spec.dec_in_slot, F, &dv);
    rc = launch_copy_rows_f32(zv.p + zv.hist, zv.ld, dv.p + dv.hist, dv.ld, B * zv.C, zv.n, s);
  }
  if (!rc) rc = run_stream_ops(h, h->spec.n_enc_ops, h->spec.n_ops, B, F, h->spec.dec_out_slot, (float*)y, s);
  h->launches += g_launches - before;
  return rc;
}

int rb_stream_state_elements(rb_handle* h, int64_t* per_stream) {
  RB_CHECK(h && per_stream, RB_ERR_INVALID, "null argument");
  int64_t n = 0;
  for (size_t sl = 0; sl < h->slots.size(); ++sl) n += (int64_t)h->slots[sl].channels * h->slots[sl].hist;
  *per_stream = n;
  return RB_OK;
}

}  // extern "C"
