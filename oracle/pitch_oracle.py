"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the pitch front-end that produces the f0 channels.

  rave/pitch_utils.py:15-42     estimate (YIN): framing, difference function, search, period -> Hz
  rave/pitch_utils.py:45-96     _frame, _diff (FFT autocorrelation, eq. 6 and 8), _search
  rave/pitch_utils.py:100-106   get_pitch (frame stride chosen so that one frame lands per block)
  rave/pitch_utils.py:116-135   extract_utterance_log_f0 / quantize_f0_norm ('abs') / get_f0_norm:
                                log-f0 normalisation, 256 bins + one "unvoiced" class, one-hot

The live export path takes f0 from FCPE (torchfcpe, a pretrained third-party network that is not in
the tree: pitch_utils.py:10-12,139-146); `get_pitch` is the self-contained estimator the file keeps
beside it (pitch_utils.py:117).  f0 -> one-hot is restated for any f0 source.

PINNING: tests/test_oracle_pins.py::test_pitch_oracle_matches_live_reference (functions loaded by
path, FCPE stubbed) and tests/golden/pitch.npz.  rave_b200/ never imports this file.
"""
import math

import torch
import torch.nn.functional as F


def yin_params(sample_rate, pitch_min, pitch_max, frame_stride):
    tau_min = int(sample_rate / pitch_max)
    tau_max = int(sample_rate / pitch_min)
    return tau_min, tau_max, 2 * tau_max, int(frame_stride * sample_rate)


def cmdf(frames, tau_max):
    """pitch_utils.py:52-75: cumulative-mean-normalised difference, lags 1 .. tau_max-1."""
    fft_size = int(2 ** (-int(-math.log(frames.shape[-1]) // math.log(2)) + 1))
    fft = torch.fft.rfft(frames, fft_size, dim=-1)
    corr = torch.fft.irfft(fft * fft.conj())[..., :tau_max]
    sqrcs = F.pad((frames * frames).cumsum(-1), [1, 0])
    corr_0 = sqrcs[..., -1:]
    corr_tau = sqrcs.flip(-1)[..., :tau_max] - sqrcs[..., :tau_max]
    diff = corr_0 + corr_tau - 2 * corr
    return (diff[..., 1:] * torch.arange(1, diff.shape[-1])
            / torch.clamp(diff[..., 1:].cumsum(-1), min=1e-5))


def search(c, tau_max, threshold):
    """pitch_utils.py:77-91."""
    first_below = (c < threshold).int().argmax(-1, keepdim=True)
    first_below = torch.where(first_below > 0, first_below, tau_max)
    beyond = torch.arange(c.shape[-1]) >= first_below
    rising = F.pad(c.diff() >= 0.0, [0, 1], value=1.0)
    return (beyond & rising).int().argmax(-1)


def estimate(signal, sample_rate=44100, pitch_min=20.0, pitch_max=20000.0, frame_stride=0.01, threshold=0.1):
    """pitch_utils.py:15-42.  signal [..., T] -> f0 [..., n_frames] in Hz, 0 = not periodic."""
    tau_min, tau_max, frame_length, stride = yin_params(sample_rate, pitch_min, pitch_max, frame_stride)
    if signal.shape[-1] < frame_length:
        signal = F.pad(signal, [0, frame_length - signal.shape[-1]])
    frames = signal.unfold(-1, frame_length, stride)
    c = cmdf(frames, tau_max)[..., tau_min:]
    tau = search(c, tau_max, threshold)
    return torch.where(tau > 0, sample_rate / (tau + tau_min + 1).type(signal.dtype),
                       torch.tensor(0).type(signal.dtype))


def get_pitch(x, block_size, fs=44100, pitch_min=70.0, pitch_max=400.0):
    """pitch_utils.py:100-106."""
    n_frames = x.shape[-1] / block_size
    tau_max = int(fs / pitch_min)
    frame_length = 2 * tau_max
    frame_stride = (x.shape[-1] - frame_length) / (n_frames - 1) / fs
    return estimate(x, sample_rate=fs, pitch_min=pitch_min, pitch_max=pitch_max, frame_stride=frame_stride)


def f0_norm_one_hot(f0, num_f0_bins=256):
    """pitch_utils.py:116-135 with norm_mode 'abs' (the default the export path uses): f0 [B, F] in Hz
    (0 = unvoiced) -> (one_hot [B, F, bins + 1], log_f0_norm [B, F]).  Unvoiced frames become NaN,
    which bucketize sends past the last boundary: class `bins`."""
    f0 = f0.clone()
    f0[f0 == 0] = float("nan")
    log_f0 = torch.log(f0)
    t40, t400 = torch.tensor([40.]).to(f0), torch.tensor([400.]).to(f0)
    log_f0_norm = (log_f0 - torch.log(t40)) / (torch.log(t400 - torch.log(t40)))     # (sic) upstream parenthesisation
    log_f0_norm = log_f0_norm + 0.5
    bins = torch.linspace(0, 1, num_f0_bins + 1).to(f0)
    idx = torch.bucketize(log_f0_norm, bins, right=True) - 1
    eye = torch.eye(num_f0_bins + 1).to(f0)                    # one_hot() :107-113: a row lookup, so -1 wraps
    return eye[idx.reshape(-1)].reshape(idx.shape[0], idx.shape[1], num_f0_bins + 1), log_f0_norm
