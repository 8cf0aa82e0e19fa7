"""Host-side mirror of the export wrapper's latent post / pre-processing (scripts/export.py:482-539) and
of rave/blocks.py:926-956 (`unit_norm_vector_to_angles`, `angles_to_unit_norm_vector`).  All compute in
librave_b200.so (csrc/latent.cu); CUDA tensors only, no CPU fallback."""
import ctypes as C

import numpy as np
import torch

from . import capi


def _check(t, what):
    if not t.is_cuda or t.dim() != 3:
        raise ValueError(f"{what} must be a CUDA tensor [B, C, F]")
    return t.to(torch.float32).contiguous()


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


class LatentPCA:
    """`latent_pca` [C, C] / `latent_mean` [C] of a variational model (export.py:80-92, :484-499)."""

    def __init__(self, latent_pca, latent_mean, device="cuda:0"):
        self._lib = capi.load_library()
        pca = np.ascontiguousarray(torch.as_tensor(latent_pca).detach().cpu().numpy(), dtype=np.float32)
        mean = np.ascontiguousarray(torch.as_tensor(latent_mean).detach().cpu().numpy(), dtype=np.float32)
        if pca.ndim != 2 or pca.shape[0] != pca.shape[1] or mean.shape != (pca.shape[0],):
            raise ValueError("latent_pca must be [C, C] and latent_mean [C]")
        self.C = pca.shape[0]
        dev = torch.device(device)
        self.device = torch.device("cuda", dev.index if dev.index is not None else torch.cuda.current_device())
        fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))  # noqa: E731
        h = C.c_void_p()
        capi.check(self._lib.rb_latent_create(self.device.index, self.C, fp(pca), fp(mean), C.byref(h)))
        self._h = h

    def post_process_latent(self, z, latent_size):
        """export.py:486-489 on an already reparametrized z [B, C, F] -> [B, latent_size, F]."""
        z = _check(z, "z")
        if z.shape[1] != self.C:
            raise ValueError(f"z has {z.shape[1]} channels, the PCA {self.C}")
        out = torch.empty(z.shape[0], latent_size, z.shape[2], device=z.device, dtype=torch.float32)
        capi.check(self._lib.rb_latent_post(self._h, z.data_ptr(), z.shape[0], z.shape[2], int(latent_size),
                                            out.data_ptr(), _stream(z.device)))
        return out

    def pre_process_latent(self, z, noise=None):
        """export.py:491-499; `noise` [B, C - latent_size, F] is the randn draw (None: zeros)."""
        z = _check(z, "z")
        n = z.shape[1]
        if noise is not None:
            noise = _check(noise, "noise")
            if noise.shape != (z.shape[0], self.C - n, z.shape[2]):
                raise ValueError("noise must be [B, C - latent_size, F]")
        out = torch.empty(z.shape[0], self.C, z.shape[2], device=z.device, dtype=torch.float32)
        capi.check(self._lib.rb_latent_pre(self._h, z.data_ptr(), noise.data_ptr() if noise is not None else None,
                                           z.shape[0], z.shape[2], n, out.data_ptr(), _stream(z.device)))
        return out

    def close(self):
        if getattr(self, "_h", None):
            self._lib.rb_latent_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def unit_norm_vector_to_angles(x):
    """blocks.py:926-939: [B, C, F] -> [B, C - 1, F]."""
    x = _check(x, "x")
    out = torch.empty(x.shape[0], x.shape[1] - 1, x.shape[2], device=x.device, dtype=torch.float32)
    capi.check(capi.load_library().rb_unit_norm_to_angles(x.data_ptr(), x.shape[0], x.shape[1], x.shape[2],
                                                          out.data_ptr(), _stream(x.device)))
    return out


def angles_to_unit_norm_vector(angles):
    """blocks.py:942-956: [B, C - 1, F] -> [B, C, F]."""
    a = _check(angles, "angles")
    out = torch.empty(a.shape[0], a.shape[1] + 1, a.shape[2], device=a.device, dtype=torch.float32)
    capi.check(capi.load_library().rb_angles_to_unit_norm(a.data_ptr(), a.shape[0], a.shape[1] + 1, a.shape[2],
                                                          out.data_ptr(), _stream(a.device)))
    return out


class DiscreteLatent:
    """`DiscreteScriptedRAVE.post_process_latent / pre_process_latent` (scripts/export.py:503-517) on the model's
    own RVQ (`model.encoder.rvq`, csrc/rvq.cu): the latent leaves as the residual-VQ indices cast to float and
    comes back through a clamp to the codebook range, the index -> embedding sum and, when the encoder was
    trained with noise augmentation, extra noise channels (the reference draws `randn`; here an argument)."""

    def __init__(self, model):
        if model.spec.rvq is None:
            raise ValueError("DiscreteLatent needs a discrete config (encoder with an RVQ)")
        self._m = model

    def post_process_latent(self, z):
        """z [B, D, F] (the encoder's latent channels, any float dtype) -> indices as float [B, Q, F]."""
        return self._m.encoder.rvq.encode(z.float().contiguous()).float()

    def pre_process_latent(self, z, noise=None):
        """z [B, Q, F] float indices -> quantised latent [B, D (+ noise channels), F] fp32."""
        k = self._m.spec.rvq["codebook_size"]
        idx = torch.clamp(z, 0, k - 1).long()
        zq = self._m.encoder.rvq.decode(idx)
        if noise is not None:
            zq = torch.cat([zq, noise.to(zq)], 1)
        return zq
