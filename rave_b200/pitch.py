"""Host-side mirror of the reference's pitch front-end (rave/pitch_utils.py): `estimate` / `get_pitch`
(YIN, :15-106) and the f0 -> normalised log-f0 -> one-hot conditioning channels (`get_f0_norm`,
:116-135).  Same function names and arguments; all compute in librave_b200.so (csrc/pitch.cu), no CPU
or PyTorch fallback.  The FCPE network the export path queries for f0 (torchfcpe, :10-12,139-146) is a
pretrained third-party model outside the tree: any f0 source can feed `f0_norm_one_hot`.
"""
import ctypes as C

import torch

from . import capi


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


def estimate(signal, sample_rate: int = 44100, pitch_min: float = 20.0, pitch_max: float = 20000.0,
             frame_stride: float = 0.01, threshold: float = 0.1) -> torch.Tensor:
    """pitch_utils.py:15-42.  signal: CUDA tensor [..., T] -> f0 [..., n_frames] (Hz, 0 = not periodic)."""
    signal = torch.as_tensor(signal)
    if not signal.is_cuda:
        raise RuntimeError("rave_b200.pitch.estimate needs a CUDA tensor (there is no CPU fallback)")
    lib = capi.load_library()
    lead, T = signal.shape[:-1], signal.shape[-1]
    x = signal.reshape(-1, T).to(torch.float32).contiguous()
    n = C.c_int()
    capi.check(lib.rb_yin_frames(T, int(sample_rate), float(pitch_min), float(frame_stride), C.byref(n)))
    f0 = torch.empty(x.shape[0], n.value, device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        capi.check(lib.rb_yin_estimate(x.data_ptr(), x.shape[0], T, int(sample_rate), float(pitch_min), float(pitch_max),
                                       float(frame_stride), float(threshold), f0.data_ptr(), n.value, _stream(x.device)))
    return f0.reshape(*lead, n.value).to(signal.dtype)


def get_pitch(x, block_size: int, fs: int = 44100, pitch_min: float = 70.0, pitch_max: float = 400.0):
    """pitch_utils.py:100-106: one frame per block of `block_size` samples."""
    desired_num_frames = x.shape[-1] / block_size
    tau_max = int(fs / pitch_min)
    frame_length = 2 * tau_max
    frame_stride = (x.shape[-1] - frame_length) / (desired_num_frames - 1) / fs
    return estimate(x, sample_rate=fs, pitch_min=pitch_min, pitch_max=pitch_max, frame_stride=frame_stride)


def f0_norm_one_hot(f0, num_f0_bins: int = 256, channels_first: bool = False):
    """get_f0_norm (pitch_utils.py:116-135, norm_mode 'abs'): f0 [B, F] in Hz (0 = unvoiced) ->
    (one_hot [B, F, bins + 1] -- or [B, bins + 1, F], what scripts/export.py:359 permutes to --,
    log_f0_norm [B, F] with NaN where unvoiced)."""
    if not f0.is_cuda or f0.dim() != 2:
        raise ValueError("f0 must be a CUDA tensor [B, F]")
    lib = capi.load_library()
    f = f0.to(torch.float32).contiguous()
    B, F = f.shape
    shape = (B, num_f0_bins + 1, F) if channels_first else (B, F, num_f0_bins + 1)
    oh = torch.empty(shape, device=f.device, dtype=torch.float32)
    ln = torch.empty(B, F, device=f.device, dtype=torch.float32)
    with torch.cuda.device(f.device):
        capi.check(lib.rb_f0_one_hot(f.data_ptr(), B, F, int(num_f0_bins), int(channels_first), oh.data_ptr(),
                                     ln.data_ptr(), _stream(f.device)))
    return oh, ln
