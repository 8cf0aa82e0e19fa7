"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the export-time latent post / pre-processing.

  scripts/export.py:484-499   VariationalScriptedRAVE.post_process_latent (after reparametrize) /
                              pre_process_latent (the randn noise is an argument here)
  rave/blocks.py:926-956      unit_norm_vector_to_angles / angles_to_unit_norm_vector
                              (SphericalScriptedRAVE, export.py:533-539)

PINNING: the angle maps are compared with the live rave/blocks.py functions
(tests/test_oracle_pins.py::test_latent_oracle_angles_match_live_reference).  scripts/export.py itself
cannot be imported here (absl flags, cached_conv, nn_tilde), so its `*ScriptedRAVE` class definitions are
cut out of the file's syntax tree and executed over a stub base (oracle/ref_loader.py:
load_export_latent_classes): the two PCA methods below are compared bit for bit with the reference's own
method bodies, randn draw included (::test_latent_oracle_pca_matches_live_export_methods).  Pinned.
rave_b200/ never imports this file.
"""
import numpy as np
import torch
import torch.nn.functional as F


def post_process_latent(z, latent_pca, latent_mean, latent_size):
    """export.py:486-489 (z already reparametrized)."""
    z = z - latent_mean.unsqueeze(-1)
    z = F.conv1d(z, latent_pca.unsqueeze(-1))
    return z[:, :latent_size]


def pre_process_latent(z, noise, latent_pca, latent_mean):
    """export.py:491-499 with the randn draw passed in."""
    z = torch.cat([z, noise], 1)
    z = F.conv1d(z, latent_pca.T.unsqueeze(-1))
    return z + latent_mean.unsqueeze(-1)


def unit_norm_vector_to_angles(x):
    """blocks.py:926-939."""
    norms = x.flip(1).pow(2)
    norms[:, 1] += norms[:, 0]
    norms = norms[:, 1:]
    norms = norms.cumsum(1).flip(1).sqrt()
    angles = torch.arccos(x[:, :-1] / norms)
    angles[:, -1] = torch.where(x[:, -1] >= 0, angles[:, -1], 2 * np.pi - angles[:, -1])
    angles[:, :-1] = angles[:, :-1] / np.pi
    angles[:, -1] = angles[:, -1] / (2 * np.pi)
    return 2 * (angles - .5)


def angles_to_unit_norm_vector(angles):
    """blocks.py:942-956."""
    angles = (angles / 2 + .5) % 1
    angles[:, :-1] = angles[:, :-1] * np.pi
    angles[:, -1] = angles[:, -1] * (2 * np.pi)
    cos = angles.cos()
    sin = angles.sin().cumprod(dim=1)
    cos = torch.cat([cos, torch.ones(cos.shape[0], 1, cos.shape[-1]).type_as(cos)], 1)
    sin = torch.cat([torch.ones(sin.shape[0], 1, sin.shape[-1]).type_as(sin), sin], 1)
    return cos * sin
