"""-m gpu: rb_speaker_embed_noalloc, the allocation-free variant of rb_speaker_embed (no counterpart in the
reference; added for callers that cannot call cudaMalloc).  Its
workspace is a file-scope static device array shared by every call, so the test runs two calls in a row on
different inputs and sizes -- the second must not be disturbed by what the first left behind -- and compares
each with the existing entry point (same kernels, same order: bit for bit, well inside the 1e-4 the fp32
tests use).  Collected last on purpose (pytest -x): a failure here hides nothing after it."""
import ctypes as C

import pytest
import torch

from .util import golden_state_dict, load_golden, rel_l2

pytestmark = pytest.mark.gpu
F32_TOL = 1e-4


def _encoder():
    from rave_b200.speaker import SpeakerRAVE
    meta, _ = load_golden("speaker")
    return SpeakerRAVE(golden_state_dict(meta))


def _embed(enc, fn, x, check_no_alloc=False):
    emb = torch.full((x.shape[0], 256), float("nan"), device=x.device, dtype=torch.float32)
    stream = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]              # the DEVICE's free memory: sees a cudaMalloc of the library
    rc = fn(enc._h, x.data_ptr(), x.shape[0], x.shape[2], emb.data_ptr(), stream)
    torch.cuda.synchronize()
    if check_no_alloc:
        assert torch.cuda.mem_get_info()[0] == free0, "the no-alloc entry point changed the device's free memory"
    return rc, emb


def test_speaker_embed_noalloc_matches_allocating_entry_point_twice():
    from oracle.weights_recipe import recipe_input
    from rave_b200 import capi
    enc = _encoder()
    lib = capi.load_library()
    # two calls in a row through the C-ABI, different shapes: the second reuses (and overwrites) the scratch
    for shape, seed in (((2, 16, 2048), 11), ((3, 16, 256), 12), ((1, 16, 4096), 13)):
        x = torch.from_numpy(recipe_input(shape, seed, scale=1.0)).cuda()
        rc, want = _embed(enc, lib.rb_speaker_embed, x)
        assert rc == 0, lib.rb_last_error()
        rc, got = _embed(enc, lib.rb_speaker_embed_noalloc, x, check_no_alloc=True)
        assert rc == 0, lib.rb_last_error()
        assert torch.isfinite(got).all()
        assert rel_l2(got.cpu().numpy(), want.cpu().numpy()) < F32_TOL
        assert torch.equal(got, want)                       # same kernels in the same order
    # beyond the static scratch: a clear error, nothing launched
    big = torch.zeros(4, 16, 2048, device="cuda")
    rc, _ = _embed(enc, lib.rb_speaker_embed_noalloc, big)
    assert rc == capi.RB_ERR_INVALID
    assert b"scratch" in lib.rb_last_error()
    with pytest.raises(ValueError):
        capi.check(rc)
    # and the allocating entry point still takes that size
    rc, emb = _embed(enc, lib.rb_speaker_embed, big)
    assert rc == 0 and torch.isfinite(emb).all()
