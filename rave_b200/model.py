"""`RAVE`: drop-in for the reference's `rave.RAVE` on the encode/decode inference path.

Same three methods and shapes as rave/model.py:594-634 --
    encode(x[B,1,T]) -> z[B, latent+speaker, T/ratio];  decode(z) -> y[B,1,T];  forward(x)
-- and the attribute names its callers read (scripts/export.py:162-164,394-405,503-517:
`.pqmf`, `.encoder`, `.decoder`, `.sr`, `.latent_size`, `.encoder.rvq.encode/decode`,
`.latent_pca/.latent_mean/.fidelity`).  No Lightning, no training code.  Construction
follows the reference: parse gin files, then build (`RAVE.from_gin`), with
`streaming=True` standing for `cc.use_cached_conv(True)` before construction
(export.py:543), then `load_state_dict(ckpt["state_dict"])` (export.py:566-569).

All arithmetic runs in librave_b200.so through the C-ABI (include/rave_b200.h); torch is
used for device memory and the current CUDA stream only.  CPU tensors raise: there is
no fallback path.
"""
import ctypes as C
from types import SimpleNamespace

import numpy as np
import torch

from . import capi, gin_lite
from .netspec import NetSpec, build_netspec
from .pqmf_design import design as pqmf_design
from .weights import collect

MAX_BATCH_SIZE = 64   # cached_conv.MAX_BATCH_SIZE: default number of stream rows


def _stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class _RVQ:
    """`model.encoder.rvq` (quantization.ResidualVectorQuantization inference surface)."""

    def __init__(self, owner):
        self._o = owner

    def encode(self, z: torch.Tensor) -> torch.Tensor:
        o = self._o
        z = o._check_input(z, o.spec.rvq["dim"], "rvq.encode", force_f32=True)
        B, _, F = z.shape
        idx = torch.empty(B, o.spec.rvq["num_quantizers"], F, dtype=torch.int64, device=z.device)
        capi.check(o._lib.rb_rvq_encode(o._h, z.data_ptr(), B, F, idx.data_ptr(), _stream_ptr()))
        return idx

    def decode(self, q_indices: torch.Tensor) -> torch.Tensor:
        o = self._o
        if not q_indices.is_cuda:
            raise ValueError("rvq.decode: indices must live on the CUDA device (no CPU path)")
        if q_indices.dtype != torch.int64 or q_indices.dim() != 3 or \
                q_indices.shape[1] != o.spec.rvq["num_quantizers"]:
            raise ValueError("rvq.decode expects int64 [B, num_quantizers, F]")
        q_indices = q_indices.contiguous()
        B, _, F = q_indices.shape
        zq = torch.empty(B, o.spec.rvq["dim"], F, dtype=torch.float32, device=q_indices.device)
        capi.check(o._lib.rb_rvq_decode(o._h, q_indices.data_ptr(), B, F, zq.data_ptr(), _stream_ptr()))
        return zq


class RAVE:
    def __init__(self, spec: NetSpec, device="cuda:0", dtype=torch.float32):
        if not torch.cuda.is_available():
            raise RuntimeError("rave_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        if dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("dtype must be torch.float32 or torch.bfloat16")
        self.spec = spec
        self.device = torch.device(device)
        self.dtype = dtype
        self._rb_dtype = capi.RB_F32 if dtype == torch.float32 else capi.RB_BF16
        self._lib = capi.load_library()
        self._h = C.c_void_p()
        self._keep = []
        self._adain_rows = {}
        slots = (capi.RbSlot * len(spec.slots))(*[capi.RbSlot(s.channels, s.per_frame, s.hist, 0) for s in spec.slots])
        ops = (capi.RbOp * len(spec.ops))()
        for i, o in enumerate(spec.ops):
            for name, _ in capi.RbOp._fields_[:-1]:
                setattr(ops[i], name, int(getattr(o, name)))
        ns = capi.RbNetSpec()
        ns.abi_version = capi.ABI_VERSION
        ns.n_slots, ns.slots = len(spec.slots), slots
        ns.n_enc_ops, ns.n_ops, ns.ops = spec.n_enc_ops, len(spec.ops), ops
        ns.enc_in_slot, ns.enc_out_slot = spec.enc_in, spec.enc_out
        ns.dec_in_slot, ns.dec_out_slot = spec.dec_in, spec.dec_out
        ns.samples_per_frame = spec.samples_per_frame
        ns.streaming = int(spec.streaming)
        ns.cond_channels = spec.cond_channels
        ns.noise_slot = spec.noise_slot
        if spec.rvq:
            ns.rvq_quantizers = spec.rvq["num_quantizers"]
            ns.rvq_codebook_size = spec.rvq["codebook_size"]
            ns.rvq_dim = spec.rvq["dim"]
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        capi.check(self._lib.rb_create(C.byref(ns), idx, C.byref(self._h)))
        # attribute surface of the reference model
        self.sr = spec.sr
        self.latent_size = spec.latent_size
        self.cond_channels = spec.cond_channels
        h, hk, wa, ws = pqmf_design(spec.attenuation, spec.n_band)
        self.pqmf = SimpleNamespace(n_band=spec.n_band, h=h, hk=hk, forward_conv_weight=wa,
                                    inverse_conv_weight=ws)
        self.encoder = SimpleNamespace(noise_augmentation=0, enabled=int(spec.rvq is not None),
                                       rvq=_RVQ(self) if spec.rvq else None, kind=spec.encoder_type)
        self.decoder = SimpleNamespace(latent_size=spec.dec_latent_size)
        self.speaker_encoder = None   # see load_speaker_encoder (rave_b200/speaker.py)
        self.latent_pca = np.eye(spec.latent_size, dtype=np.float32)
        self.latent_mean = np.zeros(spec.latent_size, dtype=np.float32)
        self.fidelity = np.zeros(spec.latent_size, dtype=np.float32)
        self._streams_ready = False
        self._noise = None
        # PQMF filters do not depend on the checkpoint: load them now
        self._load([(i, capi.RB_W_CONV, w) for i, o in enumerate(spec.ops)
                    for w in ([wa] if o.kind == 2 else [ws] if o.kind == 3 else [])])

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_gin(cls, config_files, bindings=None, streaming=False, device="cuda:0",
                 dtype=torch.float32, search_paths=None):
        cfg = gin_lite.parse_config_files_and_bindings(config_files, bindings, search_paths)
        return cls(build_netspec(cfg, streaming=streaming), device=device, dtype=dtype)

    def __del__(self):
        try:
            if getattr(self, "_h", None) and self._h.value:
                self._lib.rb_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _load(self, items):
        if not items:
            return
        arr = (capi.RbTensorDesc * len(items))()
        keep = []
        for j, (op, role, a) in enumerate(items):
            a = np.ascontiguousarray(a, dtype=np.float32)
            keep.append(a)
            arr[j].op, arr[j].role = int(op), int(role)
            arr[j].data = a.ctypes.data_as(C.POINTER(C.c_float))
            arr[j].numel = a.size
            if int(role) == capi.RB_W_ADAIN_STATS:           # rows of the statistics tables (cc.MAX_BATCH_SIZE)
                self._adain_rows[int(op)] = (a.size - 2) // (4 * self.spec.ops[int(op)].c_in)
        capi.check(self._lib.rb_load_weights(self._h, arr, len(items)))

    def load_state_dict(self, state_dict, strict=True, max_batch_size=None):
        """Accepts a reference checkpoint's `state_dict` (weight-norm pairs or folded).  `max_batch_size`:
        what `cc.MAX_BATCH_SIZE` would be set to before construction -- the per-stream AdaIN statistics tables
        (blocks.py:860-866, 64 rows in a shipped checkpoint) are extended to that many rows with the
        constructor's values, so more than 64 streams can run (and learn their own statistics)."""
        self.skipped_keys = []          # strict=False: what the state_dict did not provide (upstream's missing_keys)
        self._load(collect(self.spec, state_dict, (self.pqmf.forward_conv_weight,
                                                   self.pqmf.inverse_conv_weight), strict=strict,
                           skipped=self.skipped_keys, max_batch_size=max_batch_size))
        return self

    @classmethod
    def from_run(cls, run, streaming=False, ema_weights=False, device="cuda:0", dtype=torch.float32,
                 search_paths=None):
        """scripts/export.py:541-568: `run/config.gin` + the newest `*last*.ckpt` under `run`."""
        import os
        from . import checkpoint as ck
        m = cls.from_gin([os.path.join(run, "config.gin")], None, streaming=streaming, device=device, dtype=dtype,
                         search_paths=search_paths)
        path = ck.search_for_run(run)
        if path is None:
            print("No checkpoint found, RAVE will remain randomly initialized")      # export.py:570
            return m
        m.load_checkpoint(path, ema_weights=ema_weights)
        return m

    def load_checkpoint(self, path, ema_weights=False):
        """scripts/export.py:557-568 (strict=False as upstream); also builds the speaker encoder when the
        checkpoint carries `speaker_encoder.*` (model.py:164-174)."""
        from . import checkpoint as ck
        own, spk = ck.split_state_dict(ck.load_checkpoint(path, ema_weights=ema_weights))
        self.load_state_dict(own, strict=False)
        if spk and all(k + ".weight" in spk or k + ".weight_g" in spk for k in ("in_layer", "fc6")) \
                and "bn5.running_var" in spk:
            self.load_speaker_encoder(spk)
        return self

    def set_speaker(self, emb):
        """`self.speaker` of the reference (model.py:246-247): the fixed [1, S, 1] embedding -- or one row
        per stream, [B, S] / [B, S, 1]: stream b of the following calls is conditioned on row b (the reference
        repeats its single vector over the batch, model.py:615-618; the export wrapper swaps it per call,
        scripts/export.py:381-394)."""
        S = self.spec.cond_channels
        if isinstance(emb, torch.Tensor):
            t = emb.detach().to(device=self.device, dtype=torch.float32)
        else:
            t = torch.as_tensor(np.asarray(emb, dtype=np.float32), device=self.device)
        if t.numel() % S or t.numel() == 0:
            raise ValueError(f"set_speaker: expected [rows, {S}] values, got {tuple(t.shape)}")
        t = t.reshape(-1, S).contiguous()
        capi.check(self._lib.rb_set_conditioning(self._h, t.data_ptr(), t.shape[0], None, 0, 0, 0, _stream_ptr()))

    def set_f0(self, f0):
        """The pitch channels `f0_norm` the export wrapper appends behind the speaker channels
        (scripts/export.py:358-359,397): [B, f0_channels, F] for the following calls of B streams x F frames
        (each streaming chunk needs its own); None clears them (zeros).  Only nets whose decoder is bound
        wider than latent + speaker take any (`spec.f0_channels`)."""
        if f0 is None:
            capi.check(self._lib.rb_set_conditioning(self._h, None, 0, None, -1, 0, 0, _stream_ptr()))
            return
        if not isinstance(f0, torch.Tensor) or not f0.is_cuda:
            raise ValueError("set_f0: f0 must be a CUDA tensor")
        if f0.dim() != 3 or f0.shape[1] != self.spec.f0_channels or self.spec.f0_channels == 0:
            raise ValueError(f"set_f0: expected [B, {self.spec.f0_channels}, F], got {tuple(f0.shape)}")
        t = f0.detach().to(torch.float32).contiguous()
        capi.check(self._lib.rb_set_conditioning(self._h, None, 0, t.data_ptr(), t.shape[0], t.shape[1], t.shape[2],
                                                 _stream_ptr()))

    def load_speaker_encoder(self, state_dict):
        """model.py:164-174: build the speaker encoder (CombinedRave.SpeakerRAVE) from its state_dict."""
        from .speaker import N_BANDS, SpeakerRAVE
        wa = np.asarray(self.pqmf.forward_conv_weight, dtype=np.float32)
        wa = wa.reshape(self.spec.n_band, 1, -1) if self.spec.n_band == N_BANDS else None
        self.speaker_encoder = SpeakerRAVE(state_dict, pqmf_weight=wa, device=self.device)
        return self.speaker_encoder

    def set_speaker_from_audio(self, audio):
        """model.py:246-247: speaker = speaker_encoder(pqmf(utterance)).unsqueeze(2), on the device.
        audio: [1, 1, T].  Returns the [1, 256] embedding it installed."""
        if self.speaker_encoder is None:
            raise RuntimeError("no speaker encoder: call load_speaker_encoder(state_dict) first")
        if audio.dim() != 3 or audio.shape[0] != 1:
            raise ValueError("the reference keeps one speaker embedding: audio must be [1, 1, T]")
        emb = self.speaker_encoder.embed_audio(audio)
        self.set_speaker(emb.reshape(-1))
        return emb

    def eval(self):
        return self

    def set_noise(self, noise):
        """Inject the U(-1,1) draw of the noise branch ([B, L/8, n_band, 8]); None = internal RNG."""
        if noise is None:
            self._noise = None
            capi.check(self._lib.rb_set_noise(self._h, None, 0))
            return
        if not noise.is_cuda:
            raise ValueError("noise must live on the CUDA device")
        self._noise = noise.contiguous().float()
        capi.check(self._lib.rb_set_noise(self._h, self._noise.data_ptr(), self._noise.numel()))

    # ------------------------------------------------------------------ checks
    def _check_input(self, x, channels, what, force_f32=False):
        if not isinstance(x, torch.Tensor):
            raise ValueError(f"{what}: expected a torch.Tensor")
        if not x.is_cuda:
            raise ValueError(f"{what}: tensor must live on the CUDA device (no CPU path)")
        if x.dim() != 3 or x.shape[1] != channels:
            raise ValueError(f"{what}: expected shape [B, {channels}, T], got {tuple(x.shape)}")
        if not x.is_contiguous():
            raise ValueError(f"{what}: tensor must be contiguous")
        want = torch.float32 if force_f32 else self.dtype
        return x if x.dtype == want else x.to(want)

    def _ensure_streams(self, B, n):
        if not self._streams_ready:
            self.alloc_streams(max(B, MAX_BATCH_SIZE), n)

    # ------------------------------------------------------------------ public surface
    def alloc_streams(self, max_streams, max_chunk):
        capi.check(self._lib.rb_stream_alloc(self._h, int(max_streams), int(max_chunk), self._rb_dtype))
        self._streams_ready = True

    def reset_streams(self, stream_ids=None):
        if stream_ids is None:
            capi.check(self._lib.rb_stream_reset(self._h, None, 0))
        else:
            ids = (C.c_int32 * len(stream_ids))(*stream_ids)
            capi.check(self._lib.rb_stream_reset(self._h, ids, len(stream_ids)))

    @torch.no_grad()
    def encode(self, x, f0=None):
        """model.py:594-622; `f0` ([B, f0_channels, F], optional) = set_f0(f0) for this call."""
        if f0 is not None:
            self.set_f0(f0)
        x = self._check_input(x, 1, "encode")
        B, _, T = x.shape
        spf = self.spec.samples_per_frame
        if T % spf:
            raise ValueError(f"encode: T={T} must be a multiple of the model ratio {spf}")
        cz = self.spec.slots[self.spec.enc_out].channels
        z = torch.empty(B, cz, T // spf, dtype=self.dtype, device=x.device)
        if self.spec.streaming:
            self._ensure_streams(B, T)
            capi.check(self._lib.rb_stream_encode(self._h, x.data_ptr(), B, T, z.data_ptr(), _stream_ptr()))
        else:
            capi.check(self._lib.rb_encode(self._h, x.data_ptr(), B, T, self._rb_dtype, z.data_ptr(), _stream_ptr()))
        return z

    @torch.no_grad()
    def decode(self, z):
        z = self._check_input(z, self.spec.dec_latent_size, "decode")
        B, _, F = z.shape
        y = torch.empty(B, 1, F * self.spec.samples_per_frame, dtype=self.dtype, device=z.device)
        if self.spec.streaming:
            self._ensure_streams(B, F * self.spec.samples_per_frame)
            capi.check(self._lib.rb_stream_decode(self._h, z.data_ptr(), B, F, y.data_ptr(), _stream_ptr()))
        else:
            capi.check(self._lib.rb_decode(self._h, z.data_ptr(), B, F, self._rb_dtype, y.data_ptr(), _stream_ptr()))
        return y

    @torch.no_grad()
    def forward(self, x, f0=None):
        """model.py:631-634; with `f0` the export wrapper's `myforward` core (scripts/export.py:380-404)."""
        if f0 is not None:
            self.set_f0(f0)
        x = self._check_input(x, 1, "forward")
        B, _, T = x.shape
        if T % self.spec.samples_per_frame:
            raise ValueError(f"forward: T={T} must be a multiple of the model ratio {self.spec.samples_per_frame}")
        y = torch.empty_like(x)
        if self.spec.streaming:
            self._ensure_streams(B, T)
            capi.check(self._lib.rb_stream_step(self._h, x.data_ptr(), B, T, y.data_ptr(), _stream_ptr()))
        else:
            capi.check(self._lib.rb_forward(self._h, x.data_ptr(), B, T, self._rb_dtype, y.data_ptr(), _stream_ptr()))
        return y

    __call__ = forward

    @torch.no_grad()
    def forward_stereo(self, x, f0=None):
        """The export wrapper's fake stereo (scripts/export.py:399-411): the latent is duplicated along the batch,
        decoded, and the two halves become the two output channels: [B, 1, T] -> [B, 2, T] (the channels differ only
        through the noise branch)."""
        z = self.encode(x, f0=f0)
        y = self.decode(torch.cat([z, z], 0))
        return torch.cat(y.chunk(2, 0), 1)

    def forward_host(self, x: np.ndarray) -> np.ndarray:
        """Host-buffer entry (rb_forward_host): H2D copy, forward, D2H copy, synchronous."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 3 or x.shape[1] != 1:
            raise ValueError("forward_host expects [B, 1, T]")
        y = np.empty_like(x)
        capi.check(self._lib.rb_forward_host(self._h, x.ctypes.data_as(C.c_void_p), x.shape[0], x.shape[2],
                                             capi.RB_F32, y.ctypes.data_as(C.c_void_p)))
        return y

    def forward_host_async(self, x_host: torch.Tensor, y_host: torch.Tensor, n_chunks: int = 4):
        """Pipelined host-buffer forward (rb_forward_host_async): x_host / y_host are PINNED CPU tensors
        [B, 1, T] of the model dtype.  Only enqueues: blocks of B / n_chunks streams are copied in,
        run and copied out on three internal streams, so copies overlap compute within the call and
        across consecutive calls.  y_host is valid after host_sync() (or, for GPU work, after
        host_join())."""
        for t, what in ((x_host, "x_host"), (y_host, "y_host")):
            if t.device.type != "cpu" or not t.is_pinned() or not t.is_contiguous() or t.dtype != self.dtype:
                raise ValueError(f"forward_host_async: {what} must be a pinned contiguous CPU tensor of {self.dtype}")
        if x_host.dim() != 3 or x_host.shape[1] != 1 or y_host.shape != x_host.shape:
            raise ValueError("forward_host_async expects x_host and y_host of shape [B, 1, T]")
        B, _, T = x_host.shape
        if T % self.spec.samples_per_frame:
            raise ValueError(f"forward: T={T} must be a multiple of the model ratio {self.spec.samples_per_frame}")
        if n_chunks < 1 or B % n_chunks:
            raise ValueError(f"forward_host_async: B={B} is not a multiple of n_chunks={n_chunks}")
        capi.check(self._lib.rb_forward_host_async(self._h, x_host.data_ptr(), B, T, self._rb_dtype,
                                                   y_host.data_ptr(), n_chunks))

    def host_join(self):
        """The current torch stream waits for everything forward_host_async has enqueued."""
        capi.check(self._lib.rb_host_join(self._h, _stream_ptr()))

    def host_sync(self):
        """Blocks until every forward_host_async result has landed in its host buffer."""
        capi.check(self._lib.rb_host_sync(self._h))

    # ------------------------------------------------------------------ introspection
    def kernel_launches(self):
        n = C.c_int64()
        capi.check(self._lib.rb_kernel_launches(self._h, C.byref(n)))
        return n.value

    def stream_state_elements(self):
        n = C.c_int64()
        capi.check(self._lib.rb_stream_state_elements(self._h, C.byref(n)))
        return n.value

    def profile_ops(self, enable=True):
        capi.check(self._lib.rb_profile_ops(self._h, int(enable)))

    def op_times_ms(self):
        n = len(self.spec.ops)
        buf = (C.c_float * n)()
        capi.check(self._lib.rb_op_times(self._h, buf, n))
        return list(buf)

    # ------------------------------------------------------------------ AdaIN statistics
    def update_adain(self, learn_source=False, learn_target=False, reset_source=False, reset_target=False):
        """ScriptedRAVE.update_adain (scripts/export.py:248-265): learn_source / learn_target set
        learn_x / learn_y of every AdaptiveInstanceNormalization site (blocks.py:893-919), the
        resets restore mean 0 / std 1 / zero updates.  While a flag is set every call folds the
        statistics of its input into the running buffers, on the device."""
        capi.check(self._lib.rb_adain_update(self._h, int(bool(learn_source)), int(bool(learn_target)),
                                             int(bool(reset_source)), int(bool(reset_target))))

    def adain_state(self):
        """{reference buffer name: tensor} of every AdaIN site (mean_x, std_x, mean_y, std_y
        [MAX_BATCH_SIZE, C, 1] and the two update counters), read back from the device."""
        from .netspec import OP_AFFINE
        out = {}
        for i, op in enumerate(self.spec.ops):
            if op.kind != OP_AFFINE:
                continue
            nx, ny = C.c_int(), C.c_int()
            rows = self._adain_rows[i]
            buf = np.empty(4 * rows * op.c_in, np.float32)
            capi.check(self._lib.rb_adain_get_stats(self._h, i, buf.ctypes.data_as(C.POINTER(C.c_float)), buf.size,
                                                    C.byref(nx), C.byref(ny)))
            p = op.params["adain"]
            for j, name in enumerate(("mean_x", "std_x", "mean_y", "std_y")):
                out[p + name] = torch.from_numpy(buf[j * rows * op.c_in:(j + 1) * rows * op.c_in]
                                                 .reshape(rows, op.c_in, 1).copy())
            out[p + "num_update_x"] = torch.tensor([float(nx.value)])
            out[p + "num_update_y"] = torch.tensor([float(ny.value)])
        return out

    def set_microbatch(self, streams):
        capi.check(self._lib.rb_set_microbatch(self._h, int(streams)))
