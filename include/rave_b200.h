/* rave_b200.h -- C-ABI of the B200-native RAVE encode/decode engine.
 *
 * Drop-in boundary for the hot path of abargum/RAVE (paths relative to the reference):
 *   rave/model.py:594-634      RAVE.encode / decode / forward
 *   rave/pqmf.py:234-284       CachedPQMF.forward / inverse
 *   rave/blocks.py:508-710     EncoderV2 / GeneratorV2 (and :321-497 v1 Encoder / Generator)
 *   rave/quantization.py:302-318  ResidualVectorQuantization.encode / decode
 *   cached_conv (un-vendored)  streaming state of every conv (scripts/export.py:542-594)
 *
 * The reference is pure Python, so there is no FFI to match symbol for symbol; the
 * Python shim rave_b200/model.py is the one caller and mirrors the reference's method
 * names.  One C call per public method.  Plain pointers and sizes only -- no torch
 * types.  All functions return 0 on success or a negative rb_status; none throws or
 * exits; rb_last_error() gives the thread-local message.  All device work is enqueued
 * on the caller's stream (a cudaStream_t passed as void*); device-pointer entry points
 * never synchronise.  There is no CPU fallback: every entry point needs a CUDA device.
 */
#ifndef RAVE_B200_H
#define RAVE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RB_ABI_VERSION 1

typedef enum rb_status {
  RB_OK = 0,
  RB_ERR_INVALID = -1,     /* bad argument / shape / config  -> ValueError in the shim   */
  RB_ERR_CUDA = -2,        /* CUDA runtime error             -> RuntimeError in the shim */
  RB_ERR_STATE = -3,       /* call order (weights missing, streams not allocated, busy)  */
  RB_ERR_UNSUPPORTED = -4
} rb_status;

typedef enum rb_dtype { RB_F32 = 0, RB_BF16 = 1 } rb_dtype;

/* Activation applied to a conv's INPUT (the reference always activates before a conv:
 * blocks.py:94-107, :557, :644, :669).  act(0) == 0 for both, so it commutes with
 * zero padding. */
typedef enum rb_act { RB_ACT_NONE = 0, RB_ACT_LEAKY = 1, RB_ACT_SNAKE = 2 } rb_act;

typedef enum rb_op_kind {
  RB_OP_CONV = 1,        /* cc.Conv1d / cc.ConvTranspose1d (as r phase-convs), fused pre-act,
                            bias, residual add                                            */
  RB_OP_PQMF_FWD = 2,    /* CachedPQMF.forward  pqmf.py:269-273                            */
  RB_OP_PQMF_INV = 3,    /* CachedPQMF.inverse  pqmf.py:275-284                            */
  RB_OP_AFFINE = 4,      /* AdaIN transfer blocks.py:885-891 as per-(stream,channel) a*x+b */
  RB_OP_HEAD_V2 = 5,     /* GeneratorV2.forward tail blocks.py:699-707                     */
  RB_OP_HEAD_V1 = 6,     /* Generator.forward tail blocks.py:412-417                       */
  RB_OP_NOISE_V2 = 7,    /* NoiseGeneratorV2 tail blocks.py:281-291 (+ core.py:66,95-129)  */
  RB_OP_CONCAT_COND = 8  /* model.py:618-620: append the constant conditioning channels    */
} rb_op_kind;

/* A tensor slot: activations [B, channels, frames * per_frame] (time innermost in the
 * fp32 engine, channels innermost in the bf16 engine).  `hist` = samples of left
 * context kept across rb_stream_step calls (the cached_conv state), 0 when offline. */
typedef struct rb_slot {
  int32_t channels;
  int32_t per_frame;     /* samples of this tensor per latent frame                       */
  int32_t hist;
  int32_t reserved;
} rb_slot;

typedef struct rb_op {
  int32_t kind;          /* rb_op_kind                                                     */
  int32_t src;           /* input slot                                                     */
  int32_t dst;           /* output slot                                                    */
  int32_t res;           /* residual slot added to the output, or -1 (blocks.py:44-46)     */
  int32_t src2;          /* second input (HEAD_V2: noise, HEAD_V1: loudness) or -1          */
  int32_t c_in, c_out;   /* conv channels (c_out counts phase-stacked channels if ileave>1) */
  int32_t k, stride, dilation;
  int32_t off;           /* left offset: out[t] = sum_j W[j] x[t*stride + j*dilation - off] */
  int32_t ileave;        /* >1: out channel c = ph*(c_out/ileave)+co is stored at
                            [co][t*ileave+ph] (transposed conv phases)                     */
  int32_t pre_act;       /* rb_act on the input                                            */
  int32_t res_delay;     /* streaming: residual read at t - res_delay (AlignBranches)      */
  int32_t has_bias;
  int32_t n_band;        /* PQMF ops: number of bands M; PQMF_FWD: c_out = bands kept      */
  int32_t src_ch_off;    /* first input channel read from src (noise / split heads)        */
  int32_t reserved[3];
} rb_op;

typedef struct rb_netspec {
  int32_t abi_version;   /* RB_ABI_VERSION                                                 */
  int32_t n_slots;
  const rb_slot* slots;
  int32_t n_enc_ops;     /* ops [0, n_enc_ops) = encode program, the rest = decode program */
  int32_t n_ops;
  const rb_op* ops;
  int32_t enc_in_slot, enc_out_slot;   /* audio [B,1,T] -> latent [B,Cz,F]                 */
  int32_t dec_in_slot, dec_out_slot;   /* latent [B,Cz',F] -> audio [B,1,T]                */
  int32_t samples_per_frame;           /* total ratio: PQMF bands * prod(ratios)           */
  int32_t streaming;                   /* cc.use_cached_conv flag at construction          */
  int32_t cond_channels;               /* speaker (+pitch) channels appended by encode     */
  int32_t noise_slot;                  /* slot holding injected U(-1,1) noise, or -1       */
  /* RVQ (discrete config), 0 when absent */
  int32_t rvq_quantizers, rvq_codebook_size, rvq_dim;
  int32_t reserved[4];
} rb_netspec;

/* Parameter roles for rb_load_weights. */
typedef enum rb_role {
  RB_W_CONV = 0,      /* folded conv weight, fp32 [c_out][c_in][k] of the op as declared    */
  RB_W_BIAS = 1,      /* [c_out]                                                            */
  RB_W_ALPHA = 2,     /* Snake alpha of the op's pre-activation, [c_in]                     */
  RB_W_AFFINE_A = 3,  /* AFFINE scale  [max_streams][C]                                     */
  RB_W_AFFINE_B = 4,  /* AFFINE shift  [max_streams][C]                                     */
  RB_W_COND = 5,      /* conditioning vector [cond_channels] (op = CONCAT_COND)             */
  RB_W_CODEBOOK = 6,  /* RVQ codebooks [Q][codebook_size][dim] (op index ignored)           */
  RB_W_ADAIN_STATS = 7 /* AdaIN buffers of an AFFINE op (blocks.py:860-868): [4][max_streams][C] =
                          mean_x, std_x, mean_y, std_y, then num_update_x, num_update_y (2 floats);
                          replaces AFFINE_A / _B and allows statistic learning (rb_adain_update) */
} rb_role;

typedef struct rb_tensor_desc {
  int32_t op;          /* index into rb_netspec.ops                                         */
  int32_t role;        /* rb_role                                                           */
  const float* data;   /* HOST pointer, fp32, dense                                         */
  int64_t numel;
} rb_tensor_desc;

typedef struct rb_handle rb_handle;

int rb_create(const rb_netspec* spec, int device, rb_handle** out);
int rb_destroy(rb_handle* h);
int rb_load_weights(rb_handle* h, const rb_tensor_desc* tensors, int n);

/* Offline entry points.  x: [B,1,T], z: [B,Cz,F], y: [B,1,T], device pointers, dense,
 * dtype = element type of BOTH pointers and the engine that runs (fp32 CUDA-core engine
 * or bf16 tensor-core engine).  T must be a multiple of samples_per_frame. */
int rb_encode(rb_handle* h, const void* x, int B, int T, int dtype, void* z, void* stream);
int rb_decode(rb_handle* h, const void* z, int B, int F, int dtype, void* y, void* stream);
int rb_forward(rb_handle* h, const void* x, int B, int T, int dtype, void* y, void* stream);

/* Same path with HOST buffers (fp32): host->device copy, forward, device->host copy,
 * synchronous.  This is what a non-torch caller binds. */
int rb_forward_host(rb_handle* h, const float* x, int B, int T, int dtype, float* y);

/* Pipelined host-buffer path (what a serving loop calls).  x / y are PINNED host buffers of
 * `dtype` elements, [B,1,T].  The batch is cut into n_chunks blocks of streams (B % n_chunks == 0);
 * each block's host->device copy, forward and device->host copy run on three internal streams,
 * so the copies of one block overlap the compute of its neighbours -- and, because the call only
 * enqueues, the copies of the next call overlap the tail of this one.  x must stay unchanged and
 * y is undefined until rb_host_sync (host blocks) or until work ordered after rb_host_join
 * (`stream` waits for everything enqueued so far; nothing blocks). */
int rb_forward_host_async(rb_handle* h, const void* x, int B, int T, int dtype, void* y, int n_chunks);
int rb_host_join(rb_handle* h, void* stream);
int rb_host_sync(rb_handle* h);

/* Uniform noise U(-1,1) for the noise branch, [B, L/8, n_band, 8] fp32 device pointer
 * (the reference draws it unseeded, blocks.py:287); NULL = engine-generated (Philox). */
int rb_set_noise(rb_handle* h, const float* noise, int64_t numel);

/* Per-stream conditioning of the calls that follow (rave/model.py:615-620: emb = speaker.repeat(B, 1, F),
 * z = cat(z, emb); scripts/export.py:381-397: z = cat(z, target_emb, f0_norm)).
 *   spk: DEVICE fp32 [spk_rows][cond_channels]; spk_rows = 1 conditions every stream with the same vector
 *        (what RB_W_COND loads, the reference's single `speaker` buffer), spk_rows >= B conditions stream b
 *        with row b (offline batch row / streaming state row).  NULL keeps the current table.
 *   f0:  DEVICE fp32 [B][f0_channels][F], the pitch channels appended behind the speaker channels;
 *        f0_channels must equal decoder latent_size - latent - cond_channels as the config binds it, and B / F
 *        the streams and latent frames of the calls that use it (a streaming caller sets each chunk's f0
 *        before the step).  NULL with B = 0 keeps the current ones, NULL with B < 0 clears them: nets with
 *        pitch channels then feed zeros.
 * Both are copied (enqueued on `stream`), so the caller's buffers are free after the call returns to the stream. */
int rb_set_conditioning(rb_handle* h, const float* spk, int spk_rows, const float* f0, int B, int f0_channels,
                        int F, void* stream);

/* Discrete boundary (scripts/export.py:503-517). z fp32 [B,D,F]; idx int64 [B,Q,F]. */
int rb_rvq_encode(rb_handle* h, const float* z, int B, int F, int64_t* idx, void* stream);
int rb_rvq_decode(rb_handle* h, const int64_t* idx, int B, int F, float* zq, void* stream);

/* Streaming (cached_conv semantics): state row i <-> stream i, zero after alloc/reset. */
int rb_stream_alloc(rb_handle* h, int max_streams, int max_chunk, int dtype);
int rb_stream_reset(rb_handle* h, const int32_t* stream_ids, int n);   /* NULL = all */
int rb_stream_encode(rb_handle* h, const void* x, int B, int n, void* z, void* stream);
int rb_stream_decode(rb_handle* h, const void* z, int B, int frames, void* y, void* stream);
int rb_stream_step(rb_handle* h, const void* x, int B, int n, void* y, void* stream);
int rb_stream_state_elements(rb_handle* h, int64_t* per_stream);

int rb_workspace_bytes(rb_handle* h, int B, int T, int dtype, size_t* bytes);
int rb_kernel_launches(rb_handle* h, int64_t* count);   /* launches issued so far */
int rb_set_microbatch(rb_handle* h, int streams);       /* 0 = automatic */

/* AdaIN statistic learning (rave/blocks.py:870-883,893-919), driven like ScriptedRAVE.update_adain
 * (scripts/export.py:248-265): learn_source sets learn_x, learn_target sets learn_y on every AdaIN
 * site; reset_source / reset_target zero the means, set the stds to one and clear the counters.
 * While a flag is set every encode / decode / forward / stream step folds the per-(stream, channel)
 * mean and unbiased std of the site's input into the running statistics of rows [0, B).  Needs
 * the statistics loaded with RB_W_ADAIN_STATS. */
int rb_adain_update(rb_handle* h, int learn_source, int learn_target, int reset_source, int reset_target);
/* stats: host buffer of 4 * max_streams * C floats (mean_x, std_x, mean_y, std_y) or NULL. */
int rb_adain_get_stats(rb_handle* h, int op, float* stats, int64_t numel, int* num_update_x, int* num_update_y);
/* Speaker encoder `SpeakerRAVE` (rave/CombinedRave.py:200-328; bound by v2.gin:98 / onnx.gin:20), the
 * step before the path: model.py:246-247 computes speaker = speaker_encoder(pqmf(utterance)), the
 * vector RB_W_COND carries.  Offline, fp32.  Layers in forward order; weights are the module's own
 * tensors, [c_out][c_in][k] (fc6: [256][1536]). */
typedef enum rb_speaker_layer {
  RB_SPK_IN = 0,                                       /* in_layer        16 -> 128, k7           */
  RB_SPK_L2_DIL = 1, RB_SPK_L2_PW = 2, RB_SPK_L2_DOWN = 3,   /* layer2: DilatedUnit d=1, 128 -> 256 /4 */
  RB_SPK_L3_DIL = 4, RB_SPK_L3_PW = 5, RB_SPK_L3_DOWN = 6,   /* layer3: d=3, 256 -> 256 /4             */
  RB_SPK_L4_DIL = 7, RB_SPK_L4_PW = 8, RB_SPK_L4_DOWN = 9,   /* layer4: d=5, 256 -> 256 /2             */
  RB_SPK_CAT = 10,                                     /* cat_layer       256 -> 256, k1          */
  RB_SPK_OUT = 11,                                     /* out_layer       768 -> 768, k3          */
  RB_SPK_ATT1 = 12, RB_SPK_ATT2 = 13,                  /* attention.0 2304 -> 128, attention.3 128 -> 768 */
  RB_SPK_FC6 = 14,                                     /* fc6 Linear(1536, 256)                   */
  RB_SPK_N_LAYERS = 15
} rb_speaker_layer;
typedef struct rb_speaker rb_speaker;
int rb_speaker_create(int device, rb_speaker** out);
int rb_speaker_destroy(rb_speaker* s);
int rb_speaker_load(rb_speaker* s, int layer, const float* weight, int64_t w_numel, const float* bias,
                    int64_t b_numel);                  /* HOST pointers */
/* which = 0: attention.2 BatchNorm1d(128); 1: bn5 BatchNorm1d(1536); eval mode (running statistics) */
int rb_speaker_load_bn(rb_speaker* s, int which, const float* gamma, const float* beta, const float* mean,
                       const float* var, int n, float eps);
/* x: device fp32 [B,16,T] PQMF bands of the utterance (T % 32 == 0), emb: device fp32 [B,256];
 * enqueued on `stream`, does not synchronise (except to grow its workspace). */
int rb_speaker_embed(rb_speaker* s, const float* x, int B, int T, float* emb, void* stream);
/* The utterance itself: w = CachedPQMF.forward_conv.weight [16][1][k] (host), then audio: device fp32
 * [B,1,T], T % 512 == 0 -> pqmf (pqmf.py:269-273) -> rb_speaker_embed. */
int rb_speaker_load_pqmf(rb_speaker* s, const float* w, int n_band, int k);
int rb_speaker_embed_audio(rb_speaker* s, const float* audio, int B, int T, float* emb, void* stream);
int rb_speaker_launches(rb_speaker* s, int64_t* count);
/* rb_speaker_embed without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same result; the workspace is a file-scope static device
 * array of RB_SPEAKER_NOALLOC_MAX_FLOATS floats (9.4 MB, held by every GPU that loads the library until
 * exit).  Sizes whose workspace exceeds it (about B * T > 4096 band samples) return RB_ERR_INVALID.
 * Calls no cudaMalloc / cudaFree / allocating library routine and never synchronises.  All calls on one
 * GPU share the array: calls that overlap on the device (different streams) race -- serialise them. */
#define RB_SPEAKER_NOALLOC_MAX_FLOATS (576 * 4096 + 1536 * 16)
int rb_speaker_embed_noalloc(rb_speaker* s, const float* x, int B, int T, float* emb, void* stream);

/* `rave.resampler.Resampler` (rave/resampler.py:9-66), wrapped around encode / decode by the exported
 * model (scripts/export.py:377-378, :407-408).  down_w [k_down] = downsample.weight, up_w [ratio][k_up] =
 * upsample.weight (host).  streaming = cc.USE_BUFFER_CONV at construction: per-stream history replaces the
 * centred zero padding (cached_conv.CachedConv1d), rows [0, B) <-> streams, max_streams rows allocated.
 * x / y: device, [B,1,n] of `dtype` (rb_dtype); to_model: n % ratio == 0 -> n / ratio samples;
 * from_model: -> n * ratio samples.  Enqueued on `stream`; never synchronises. */
typedef struct rb_resampler rb_resampler;
int rb_resampler_create(int device, int ratio, const float* down_w, int k_down, const float* up_w, int k_up,
                        int streaming, int max_streams, rb_resampler** out);
int rb_resampler_destroy(rb_resampler* r);
int rb_resampler_reset(rb_resampler* r, void* stream);
int rb_resampler_to_model(rb_resampler* r, const void* x, int B, int n, int dtype, void* y, void* stream);
int rb_resampler_from_model(rb_resampler* r, const void* x, int B, int n, int dtype, void* y, void* stream);
int rb_resampler_launches(rb_resampler* r, int64_t* count);

/* Pitch front-end (rave/pitch_utils.py): the f0 channels the export wrapper concatenates to the latent
 * (scripts/export.py:348-359, :397).  rb_yin_estimate = estimate() :15-42 (YIN: framing, difference
 * function, cumulative-mean normalisation, threshold + first-minimum search); x: device fp32 [B][T],
 * f0: device fp32 [B][n_frames] in Hz, 0 = not periodic; n_frames from rb_yin_frames.  get_pitch()
 * :100-106 is this call with frame_stride_s = (T - 2*int(fs/pitch_min)) / (T/block - 1) / fs.
 * rb_f0_one_hot = get_f0_norm() :116-135 ('abs' normalisation) for f0 from any estimator: one_hot device
 * fp32 [B][F][bins+1] (channels_first: [B][bins+1][F], the layout the wrapper permutes to), class `bins`
 * = unvoiced; log_f0_norm [B][F] (NaN where unvoiced) or NULL. */
int rb_yin_frames(int T, int sample_rate, double pitch_min, double frame_stride_s, int* n_frames);
int rb_yin_estimate(const float* x, int B, int T, int sample_rate, double pitch_min, double pitch_max,
                    double frame_stride_s, double threshold, float* f0, int n_frames, void* stream);
int rb_f0_one_hot(const float* f0, int B, int F, int num_bins, int channels_first, float* one_hot,
                  float* log_f0_norm, void* stream);

/* Export-time latent post / pre-processing (scripts/export.py:482-539; defined upstream but not on this fork's
 * live `myforward` path).  rb_latent_*: VariationalScriptedRAVE :484-499 after reparametrize -- pca [C][C] =
 * latent_pca, mean [C] = latent_mean (host); post: z device fp32 [B,C,F] -> out [B,n_keep,F] = pca[:n_keep] (z - mean);
 * pre: z [B,n_keep,F], noise [B,C-n_keep,F] (the reference draws randn; NULL = zeros) -> out [B,C,F] =
 * pca^T cat(z, noise) + mean.  rb_unit_norm_to_angles / rb_angles_to_unit_norm: rave/blocks.py:926-956
 * (SphericalScriptedRAVE :533-539): x [B,C,F] <-> angles [B,C-1,F]. */
typedef struct rb_latent rb_latent;
int rb_latent_create(int device, int C, const float* pca, const float* mean, rb_latent** out);
int rb_latent_destroy(rb_latent* h);
int rb_latent_post(rb_latent* h, const float* z, int B, int F, int n_keep, float* out, void* stream);
int rb_latent_pre(rb_latent* h, const float* z, const float* noise, int B, int F, int n_keep, float* out, void* stream);
int rb_unit_norm_to_angles(const float* x, int B, int C, int F, float* angles, void* stream);
int rb_angles_to_unit_norm(const float* angles, int B, int C, int F, float* x, void* stream);

/* Per-op device timing (CUDA events on the launching stream around every op of the
 * program).  enable!=0 starts recording; rb_op_times synchronises, writes the accumulated
 * milliseconds per op (n = number of ops) and clears the record. */
int rb_profile_ops(rb_handle* h, int enable);
int rb_op_times(rb_handle* h, float* ms, int n);
const char* rb_last_error(void);

/* Testing hook (tests/test_gpu_tc_conv.py): one conv op of the bf16 tensor-core engine, or
 * a conv + 1x1-conv pair (a DilatedUnit, fused when the planner fuses it), on HOST fp32
 * data.  Not part of the drop-in surface. */
typedef struct rb_debug_conv {
  rb_op op[2];            /* kind/src/dst/res are filled in by the library                  */
  int32_t n_ops;          /* 1, or 2: op[1] is the 1x1 conv whose residual is op[0]'s input */
  int32_t B, n_in;        /* streams, input samples per stream                              */
  const float* w[2];      /* folded fp32 [c_out][c_in][k]                                   */
  const float* bias[2];   /* [c_out / ileave] or NULL                                       */
  const float* alpha[2];  /* Snake alpha [c_in] or NULL                                     */
  const float* x;         /* [B][c_in][n_in]                                                */
  const float* res;       /* n_ops == 1: optional residual [B][c_out][n_out]                */
  float* y;               /* [B][c_out / ileave][n_out]                                     */
  int32_t* fused_out;     /* optional: set to 1 when the pair ran as one fused kernel       */
  int32_t flags;          /* activated storage: 1 x is leaky(x), 2 return leaky(y), 4 res   */
  uint64_t* trace;        /* optional [64][16]: clock stamps of CTA 0's pipeline events     */
} rb_debug_conv;
int rb_debug_conv_bf16(const rb_debug_conv* d);
int rb_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RAVE_B200_H */
