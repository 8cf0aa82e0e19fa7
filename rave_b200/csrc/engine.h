// engine.h -- host-side state of one rb_handle (not part of the C-ABI).
#pragma once
#include <atomic>
#include <map>
#include <functional>
#include <vector>

#include "common.cuh"

struct rb_handle;

namespace rb {

static inline int round4(int v) { return (v + 3) & ~3; }

struct RunCtx {
  rb_handle* h;
  int B, F;                      // streams in this (micro)batch, latent frames
  std::vector<View> v;           // per slot
  cudaStream_t s;
  bool streaming;
  const float* noise;            // for this microbatch
  int64_t noise_bstride;
};

int exec_op_f32(RunCtx& c, int oi);
struct OpWeights;
// AdaIN.forward in eval mode around the transfer (blocks.py:893-919); `stats_fn(mean, std, count)`
// updates the running statistics of rows [b0, b0 + B) from the op's input.  Sets *apply when the
// transfer a*x + b has to run.  adain_finalize counts the updates of ops [ob, oe) once per call.
int adain_step(rb_handle* h, OpWeights& w, int C, cudaStream_t s,
               const std::function<int(float*, float*, int)>& stats_fn, bool* apply);
void adain_finalize(rb_handle* h, int ob, int oe);
int pool_get(rb_handle* h, size_t bytes, int* idx);   // offline workspace pool (engine.cu)
// the conditioning channels CONCAT_COND op `oi` appends for B streams x n frames (rb_set_conditioning)
int cond_args(rb_handle* h, int oi, int B, int n, CondArgs* out);

// RAII: records a start/stop event pair for op `oi` when profiling is on
struct OpTimer {
  rb_handle* h; cudaStream_t s; int oi; bool on;
  OpTimer(rb_handle* h, int oi, cudaStream_t s);
  ~OpTimer();
};

struct OpWeights {
  float* w = nullptr;       // fp32 engine packing ([c_in][k][c_out]; PQMF: see engine.cu)
  float* bias = nullptr;
  float* alpha = nullptr;
  float* aff_a = nullptr;   // [rows][C]
  float* aff_b = nullptr;
  int aff_rows = 0;
  // AdaIN statistics (blocks.py:860-868): [4][aff_rows][C] = mean_x, std_x, mean_y, std_y; the
  // update counters live on the host (their increments never depend on data)
  float* stats = nullptr;
  int nx = 0, ny = 0;
  bool touch_x = false, touch_y = false;   // learned during the running call: counted once at its end
  float* cond = nullptr;    // CONCAT_COND: speaker table [cond_rows][cond_channels]
  int cond_rows = 0;        // 1: the same vector for every stream (RB_W_COND); else row b <-> stream b
  // host copies kept for the bf16 tensor-core packer (tc_engine.cu)
  std::vector<float> w_host;   // folded fp32 [c_out][c_in][k] (PQMF: as loaded)
  std::vector<float> bias_host, alpha_host;
};

struct PoolBuf {
  void* p = nullptr;
  size_t bytes = 0;
  bool busy = false;
};

}  // namespace rb

struct rb_handle {
  int device = 0;
  rb_netspec spec{};
  std::vector<rb_slot> slots;
  std::vector<rb_op> ops;
  std::vector<rb::OpWeights> w;
  bool weights_loaded = false;

  float* codebooks = nullptr;   // [Q][K][D]
  float* cb_norms = nullptr;    // [Q][K]
  float* rvq_ws = nullptr;
  size_t rvq_ws_bytes = 0;

  const float* noise = nullptr;  // user noise, device
  int64_t noise_numel = 0;
  unsigned long long noise_seed = 0x1234;

  // rb_set_conditioning: pitch channels of the coming calls, device copy [f0_B][f0_C][f0_F]
  float* f0 = nullptr;
  size_t f0_cap = 0;
  int f0_B = 0, f0_C = 0, f0_F = 0;

  std::vector<rb::PoolBuf> pool;  // offline activation workspace
  unsigned long long cond_version = 0;   // bumped whenever the speaker table changes (folded conditioning cache)
  int microbatch = 0;
  int cur_b0 = 0;               // first stream of the microbatch being executed (per-stream tables)
  bool adain_learn_x = false, adain_learn_y = false;   // export.py:248-265 learn_source / learn_target

  // streaming state arena: one persistent buffer per slot
  bool stream_ready = false;
  int stream_dtype = RB_F32;
  int max_streams = 0, max_frames = 0;
  std::vector<void*> sbuf;
  std::vector<int> s_ld, s_hist;

  // pinned staging for rb_forward_host
  float* stage_in = nullptr;
  float* stage_out = nullptr;
  void* dev_in = nullptr;
  void* dev_out = nullptr;
  size_t stage_elems = 0;

  // rb_forward_host_async: three streams, rings of device in / out blocks guarded by events
  struct HostPipe {
    static constexpr int NBUF = 3;
    cudaStream_t s_in = nullptr, s_cmp = nullptr, s_out = nullptr;
    void* din[NBUF] = {nullptr, nullptr, nullptr};
    void* dout[NBUF] = {nullptr, nullptr, nullptr};
    cudaEvent_t in_done[NBUF], cmp_done[NBUF], out_done[NBUF], join_ev;
    size_t block_bytes = 0;
    unsigned long long seq = 0;
    bool ready = false;
  } pipe;

  // per-op profiling (rb_profile_ops)
  bool profile = false;
  struct Ev { int op; cudaEvent_t a, b; };
  std::vector<Ev> events;

  void* tc = nullptr;           // rb::tc::Plan of the bf16 engine (tc_engine.cu)
  std::vector<int> tc_sact_hint;  // testing hook only: per-slot "stored activated" flags that replace the planner's

  std::atomic_flag busy = ATOMIC_FLAG_INIT;
  long long launches = 0;
};
