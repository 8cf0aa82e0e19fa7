"""-m gpu parity tests: the CUDA path (through the C-ABI) against the CPU oracle on the same
seeded inputs and against the committed golden fixtures.  Tolerances are the ones
BASELINE.json states: rel-L2 <= 1e-3 for the fp32 engine (the tests hold it to 1e-4),
<= 1e-2 for bf16, RVQ indices bit-exact."""
import numpy as np
import pytest
import torch

from .util import golden_input, golden_state_dict, load_golden, rel_l2

pytestmark = pytest.mark.gpu

GIN = {"v2_c8": (["v2.gin"], ["CAPACITY=8"]), "v2causal_c8": (["v2.gin", "causal.gin"], ["CAPACITY=8"]),
       "v3_c8": (["v3.gin"], ["CAPACITY=8"]), "v3causal_c8": (["v3.gin", "causal.gin"], ["CAPACITY=8"]),
       "discrete_c8": (["discrete.gin"], ["CAPACITY=8"]), "v2noise_c8": (["v2.gin", "noise.gin"], ["CAPACITY=8"]),
       "v1_c8": (["v1.gin"], ["CAPACITY=8"]), "raspberry_c16": (["raspberry.gin"], []),
       "raspberry_causal_c16": (["raspberry.gin", "causal.gin"], []),
       # BASELINE.json configs at their real channel widths (oracle/make_golden.py SIZED): configs[0]/[1] v2
       # CAPACITY 64, configs[2] causal v2, configs[3] discrete (CAPACITY 96, LATENT_SIZE 128, discrete.gin:13-19),
       # configs[4] v3 (Snake + AdaIN, CAPACITY 64)
       "v2_c64": (["v2.gin"], []), "v2causal_c64": (["v2.gin", "causal.gin"], []),
       "discrete_c96": (["discrete.gin"], []), "v3_c64": (["v3.gin"], []),
       "v3causal_c64": (["v3.gin", "causal.gin"], [])}
F32_TOL = 1e-4


def _noise(meta, a, x):
    """The injected U(-1,1) draw of the noise branch for input x ([B, L/8, n_band, 8]), or None."""
    if not meta["cfg"].get("noise"):
        return None
    from oracle.weights_recipe import recipe_input
    B, _, T = x.shape
    return torch.from_numpy(recipe_input((B, T // 16 // 8, 16, 8), int(a["noise_seed"].item()), uniform=True)).cuda()


def _model(name, streaming=False, dtype=torch.float32):
    from rave_b200 import RAVE
    meta, a = load_golden(name)
    files, binds = GIN[name]
    m = RAVE.from_gin(files, binds, streaming=streaming, dtype=dtype)
    m.load_state_dict(golden_state_dict(meta))
    return m, meta, a


@pytest.mark.parametrize("name", list(GIN))
def test_offline_matches_golden(name):
    m, meta, a = _model(name)
    x = torch.from_numpy(golden_input(a)).cuda()
    if meta["cfg"].get("noise"):
        from oracle.weights_recipe import recipe_input
        B, _, T = x.shape
        m.set_noise(torch.from_numpy(recipe_input((B, T // 16 // 8, 16, 8), int(a["noise_seed"].item()),
                                                  uniform=True)).cuda())
    z = m.encode(x)
    assert tuple(z.shape) == a["z"].shape
    assert rel_l2(z.cpu().numpy(), a["z"]) < F32_TOL
    y = m.decode(torch.from_numpy(a["z"]).cuda())
    assert y.shape == x.shape                       # reference tests/test_configs.py:68
    assert rel_l2(y.cpu().numpy(), a["y"]) < F32_TOL
    y2 = m(x)                                        # forward == decode(encode(x)), model.py:634
    # random-init Snake + AdaIN nets are chaotic (BF16_CHAOTIC below): the decoder amplifies the 7e-6 of the
    # encoded latent ~45x (v3_c64: 3.0e-4 measured), so their chained forward is held to BASELINE.json's
    # stated fp32 bar, 1e-3; encode and decode above keep 1e-4
    assert rel_l2(y2.cpu().numpy(), a["y"]) < (1e-3 if name.startswith("v3") else 2 * F32_TOL)


@pytest.mark.parametrize("name", list(GIN))
def test_streaming_matches_golden_chunk_by_chunk(name):
    m, meta, a = _model(name, streaming=True)
    x = torch.from_numpy(golden_input(a)).cuda()
    n = int(a["chunk"].item())
    noise = _noise(meta, a, x)
    q = n // 16 // 8

    def set_noise(c):
        if noise is not None:
            m.set_noise(noise[:, c * q:(c + 1) * q].contiguous())
    zs, ys = [], []
    for c in range(x.shape[-1] // n):
        z = m.encode(x[..., c * n:(c + 1) * n].contiguous())
        zs.append(z)
        set_noise(c)
        ys.append(m.decode(z))
    assert rel_l2(torch.cat(zs, -1).cpu().numpy(), a["z_stream"]) < F32_TOL
    assert rel_l2(torch.cat(ys, -1).cpu().numpy(), a["y_stream"]) < F32_TOL
    # fused encode->decode step on a fresh state gives the same audio
    m.reset_streams()
    ys2 = []
    for c in range(x.shape[-1] // n):
        set_noise(c)
        ys2.append(m(x[..., c * n:(c + 1) * n].contiguous()))
    assert rel_l2(torch.cat(ys2, -1).cpu().numpy(), a["y_stream"]) < F32_TOL


def test_streaming_ragged_chunks_and_reset():
    """Chunk sizes may vary call to call (multiples of the ratio); reset zeroes one stream."""
    from oracle.model import OracleRAVE
    m, meta, a = _model("v2causal_c8", streaming=True)
    o = OracleRAVE(meta["cfg"], golden_state_dict(meta), streaming=True)
    x = torch.from_numpy(golden_input(a))
    m.alloc_streams(4, 4096)
    pos = 0
    for n in (1024, 3072, 2048, 1024, 4096):
        xc = x[..., pos:pos + n].contiguous()
        pos += n
        y = m(xc.cuda())
        yo = o.forward(xc)
        assert rel_l2(y.cpu().numpy(), yo.numpy()) < F32_TOL
    m.reset_streams([1])
    o2 = OracleRAVE(meta["cfg"], golden_state_dict(meta), streaming=True)
    xc = x[..., :2048].contiguous()
    y = m(xc.cuda())
    assert rel_l2(y[1].cpu().numpy(), o2.forward(xc)[1].numpy()) < F32_TOL      # fresh state
    assert rel_l2(y[0].cpu().numpy(), o.forward(xc)[0].numpy()) < F32_TOL       # carried state


def test_causal_stream_equals_offline_after_delay_gpu():
    m, meta, a = _model("v2causal_c8", streaming=True)
    x = torch.from_numpy(golden_input(a)).cuda()
    ys = torch.cat([m(x[..., c * 2048:(c + 1) * 2048].contiguous()) for c in range(8)], -1).cpu().numpy()
    d = 928
    assert rel_l2(ys[..., 14336:], a["y"][..., 14336 - d:-d]) < 1e-4


@pytest.mark.parametrize("name", ["discrete_c8", "discrete_c96"])
def test_rvq_indices_bit_exact_and_decode(name):
    from oracle.weights_recipe import recipe_input
    m, meta, a = _model(name)
    zq = torch.from_numpy(recipe_input((4, 128, 64), int(a["rvq_z_seed"].item()), scale=1.0)).cuda()
    idx = m.encoder.rvq.encode(zq)
    assert idx.dtype == torch.int64
    assert np.array_equal(idx.cpu().numpy(), a["rvq_idx"])
    dec = m.encoder.rvq.decode(idx)
    assert rel_l2(dec.cpu().numpy(), a["rvq_dec"]) < 1e-6


def test_rvq_bit_exact_vs_oracle_large_and_ragged():
    """131072 vectors (the C4 size) plus ragged shapes, live against the oracle."""
    from oracle.model import OracleRAVE
    from oracle.weights_recipe import recipe_input
    m, meta, _ = _model("discrete_c8")
    o = OracleRAVE(meta["cfg"], golden_state_dict(meta))
    for shape, seed in (((1, 128, 1), 1), ((3, 128, 7), 2), ((5, 128, 65), 3), ((512, 128, 256), 4)):
        z = torch.from_numpy(recipe_input(shape, seed, scale=1.0))
        ref = o.rvq_encode(z)
        got = m.encoder.rvq.encode(z.cuda()).cpu()
        assert torch.equal(got, ref), f"{shape}: {(got != ref).sum().item()} mismatching indices"
        assert rel_l2(m.encoder.rvq.decode(got.cuda()).cpu().numpy(), o.rvq_decode(ref).numpy()) < 1e-6


@pytest.mark.parametrize("B,T", [(1, 1024), (3, 5120), (2, 131072)])
def test_v2_ragged_sizes_vs_oracle(B, T):
    """Smallest legal input (one frame), odd batch, and BASELINE configs[0] length 2^17."""
    from oracle.model import OracleRAVE
    from oracle.weights_recipe import recipe_input
    m, meta, _ = _model("v2_c8")
    o = OracleRAVE(meta["cfg"], golden_state_dict(meta))
    x = torch.from_numpy(recipe_input((B, 1, T), 77))
    y = m(x.cuda())
    assert rel_l2(y.cpu().numpy(), o.forward(x).numpy()) < F32_TOL


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_plain_pqmf_binding_vs_oracle(dtype):
    """`rave.RAVE.pqmf = @pqmf.PQMF` (pqmf.py:179-231; no shipped config binds it): same analysis, synthesis
    16 samples earlier than CachedPQMF.  Oracle branch pinned live in tests/test_oracle_pins.py."""
    from oracle.model import OracleRAVE
    from rave_b200 import RAVE
    meta, a = load_golden("v2_c8")
    sd = golden_state_dict(meta)
    m = RAVE.from_gin(["v2.gin"], ["CAPACITY=8", "rave.RAVE.pqmf = @pqmf.PQMF", "pqmf.PQMF.n_band = 16",
                                   "pqmf.PQMF.attenuation = 100"], dtype=dtype)
    m.load_state_dict(sd)
    o = OracleRAVE(dict(meta["cfg"], pqmf="PQMF"), sd)
    x = torch.from_numpy(golden_input(a))
    y = m(x.cuda().to(dtype)).float().cpu().numpy()
    yo = o.forward(x).numpy()
    assert rel_l2(y, yo) < (F32_TOL if dtype == torch.float32 else 2e-2)
    assert rel_l2(yo[..., :-16], a["y"][..., 16:]) < 1e-4          # the 16-sample advance against CachedPQMF


def test_error_conventions():
    from rave_b200 import RAVE
    m, meta, a = _model("v2_c8")
    with pytest.raises(ValueError):
        m.encode(torch.zeros(1, 1, 1000).cuda())                 # not a multiple of the ratio
    with pytest.raises(ValueError):
        m.encode(torch.zeros(1, 1, 1024))                        # CPU tensor: no fallback
    with pytest.raises(ValueError):
        m.decode(torch.zeros(1, 64, 4).cuda())                   # wrong latent width
    with pytest.raises(ValueError):
        m.encode(torch.zeros(2, 1, 4096).cuda()[:, :, ::2])      # non-contiguous
    raw = RAVE.from_gin(["v2.gin"], ["CAPACITY=8"])
    with pytest.raises(RuntimeError):
        raw.encode(torch.zeros(1, 1, 1024).cuda())               # weights not loaded
    with pytest.raises(KeyError):
        raw.load_state_dict({})


def test_streams_are_independent_at_full_length():
    """Size-independent property at the BASELINE configs[1] length (2^18 samples): a stream's
    output does not depend on its batch neighbours or its row (bitwise), which is what lets
    the batch be sharded across GPUs with no exchange."""
    from oracle.weights_recipe import recipe_input
    m, meta, _ = _model("v2_c8")
    x = torch.from_numpy(recipe_input((3, 1, 2 ** 18), 5)).cuda()
    y = m(x)
    y1 = m(x[1:2].contiguous())
    assert torch.equal(y[1:2], y1)
    yp = m(x.flip(0).contiguous())
    assert torch.equal(yp.flip(0), y)


BF16_TOL = 1e-2
# Nets whose bf16 FLOOR (tests/util.py:bf16_floor -- the reference algorithm itself with bf16
# weights / activations) is far above 1e-2: random-init Snake + AdaIN is chaotic in bf16.
BF16_CHAOTIC = ("v3_c8", "v3causal_c8", "v3_c64", "v3causal_c64")


@pytest.mark.parametrize("name", list(GIN))
def test_offline_bf16_matches_golden(name):
    """bf16 tensor-core engine (tcgen05) against the fp32 oracle goldens: rel-L2 <= 1e-2
    (BASELINE.json).  For the chaotic v3 fixtures the bar is 2x the bf16 floor of the reference
    algorithm instead (and the floor itself is asserted to be above 1e-2, so the exemption
    cannot hide a regression on a well-conditioned net)."""
    from .util import bf16_floor
    m, meta, a = _model(name, dtype=torch.bfloat16)
    tol_z = tol_y = tol_f = BF16_TOL
    if name in BF16_CHAOTIC:
        fz, fy, ff = bf16_floor(name)
        assert fy > BF16_TOL and ff > BF16_TOL
        tol_z, tol_y, tol_f = max(BF16_TOL, 2 * fz), 2 * fy, 2 * ff
    x = torch.from_numpy(golden_input(a)).cuda().bfloat16()
    if meta["cfg"].get("noise"):
        from oracle.weights_recipe import recipe_input
        B, _, T = x.shape
        m.set_noise(torch.from_numpy(recipe_input((B, T // 16 // 8, 16, 8), int(a["noise_seed"].item()),
                                                  uniform=True)).cuda())
    z = m.encode(x)
    assert z.dtype == torch.bfloat16 and tuple(z.shape) == a["z"].shape
    nl = m.latent_size
    assert rel_l2(z[:, :nl].float().cpu().numpy(), a["z"][:, :nl]) < tol_z
    assert rel_l2(z.float().cpu().numpy(), a["z"]) < tol_z
    y = m.decode(torch.from_numpy(a["z"]).cuda().bfloat16())
    assert y.shape == x.shape
    assert rel_l2(y.float().cpu().numpy(), a["y"]) < tol_y
    y2 = m(x)
    assert rel_l2(y2.float().cpu().numpy(), a["y"]) < (tol_f if name in BF16_CHAOTIC else 2 * BF16_TOL)


@pytest.mark.parametrize("name", [n for n in GIN if n not in ("v2noise_c8",) + BF16_CHAOTIC])
def test_streaming_bf16_matches_golden_chunk_by_chunk(name):
    """cached_conv streaming on the bf16 tensor-core engine: chunk-by-chunk outputs against the
    oracle's fp32 streaming run (same bar as the offline bf16 path)."""
    m, meta, a = _model(name, streaming=True, dtype=torch.bfloat16)
    x = torch.from_numpy(golden_input(a)).cuda().bfloat16()
    n = int(a["chunk"].item())
    zs, ys, ys2 = [], [], []
    zg = torch.from_numpy(a["z_stream"]).cuda().bfloat16()
    fpc = zg.shape[-1] // (x.shape[-1] // n)
    for c in range(x.shape[-1] // n):
        zs.append(m.encode(x[..., c * n:(c + 1) * n].contiguous()))
        ys.append(m.decode(zg[..., c * fpc:(c + 1) * fpc].contiguous()))   # decoder fed the oracle's latents
    assert rel_l2(torch.cat(zs, -1).float().cpu().numpy(), a["z_stream"]) < BF16_TOL
    assert rel_l2(torch.cat(ys, -1).float().cpu().numpy(), a["y_stream"]) < 1.2 * BF16_TOL
    # whole-step entry (rb_stream_step) on fresh state: encode -> decode with the latent kept on device
    m.reset_streams()
    for c in range(x.shape[-1] // n):
        ys2.append(m(x[..., c * n:(c + 1) * n].contiguous()))
    assert rel_l2(torch.cat(ys2, -1).float().cpu().numpy(), a["y_stream"]) < 2 * BF16_TOL


@pytest.mark.parametrize("name", ["v2_c8", "v2causal_c8", "raspberry_causal_c16", "v2_c64", "v2causal_c64",
                                  "discrete_c96"])
def test_streaming_bf16_stream_as_row_layers(name, monkeypatch, capfd):
    """Deep layers of a streaming step (<= 8 rows per stream and chunk) run with the stream index as the row
    axis and the taps unrolled into K once there are >= 128 streams; RB_TC_SR_MINB lowers that threshold so
    the 1-2 stream fixtures take the path.  Same bars as the plain streaming test, and the two paths agree."""
    m, meta, a = _model(name, streaming=True, dtype=torch.bfloat16)
    x = torch.from_numpy(golden_input(a)).cuda().bfloat16()
    n = int(a["chunk"].item())
    chunks = [x[..., c * n:(c + 1) * n].contiguous() for c in range(x.shape[-1] // n)]
    plain = torch.cat([m(c) for c in chunks], -1).float().cpu().numpy()
    monkeypatch.setenv("RB_TC_SR_MINB", "1")
    m.reset_streams()
    sr = torch.cat([m(c) for c in chunks], -1).float().cpu().numpy()
    monkeypatch.setenv("RB_TC_SR_DEBUG", "1")              # the engine reports its choice per op on stderr
    capfd.readouterr()
    m(chunks[0])
    torch.cuda.synchronize()
    assert "sr_valid 1" in capfd.readouterr().err, "stream-as-row path not taken"
    monkeypatch.delenv("RB_TC_SR_DEBUG")
    assert rel_l2(sr, a["y_stream"]) < 2 * BF16_TOL
    assert rel_l2(sr, plain) < BF16_TOL
    m.reset_streams()
    zs = torch.cat([m.encode(c) for c in chunks], -1).float().cpu().numpy()
    assert rel_l2(zs, a["z_stream"]) < BF16_TOL


@pytest.mark.parametrize("name", ["v2causal_c8", "v2causal_c64"])
def test_streaming_bf16_many_streams_fill_row_tiles(name):
    """256 streams (128 copies of the 2-stream fixture): above the default threshold the deep layers run
    stream-as-row over two 128-stream row tiles; every copy must equal the first bit for bit and match
    the oracle's streaming run.  v2causal_c64 is BASELINE configs[2] at its real width: the C = 512 / 1024
    stream-as-row twins, the 4-block weight ring and N_TILE = max(64, N / 8) of the C3 bench leg."""
    m, meta, a = _model(name, streaming=True, dtype=torch.bfloat16)
    x = torch.from_numpy(golden_input(a)).cuda().bfloat16().repeat(128, 1, 1).contiguous()
    n = int(a["chunk"].item())
    m.alloc_streams(256, n)
    y = torch.cat([m(x[..., c * n:(c + 1) * n].contiguous()) for c in range(x.shape[-1] // n)], -1)
    assert torch.equal(y[2:], y[:2].repeat(127, 1, 1))
    assert rel_l2(y[:2].float().cpu().numpy(), a["y_stream"]) < 2 * BF16_TOL
    assert rel_l2(y[-2:].float().cpu().numpy(), a["y_stream"]) < 2 * BF16_TOL
    # a stream count that is not a multiple of the 128-row tile, two frames per chunk (twice the rows per layer)
    m2, _, _ = _model(name, streaming=True, dtype=torch.bfloat16)
    m2.alloc_streams(130, 2 * n)
    x2 = x[:130].contiguous()
    y2 = torch.cat([m2(x2[..., c * 2 * n:(c + 1) * 2 * n].contiguous()) for c in range(x.shape[-1] // (2 * n))], -1)
    assert torch.equal(y2[2:], y2[:2].repeat(64, 1, 1))
    assert rel_l2(y2[:2].float().cpu().numpy(), a["y_stream"]) < 2 * BF16_TOL
    if name == "v2causal_c64":
        # the size of the C3 bench leg: 1024 streams -- the CTAs of the stream-as-row units run SEVERAL work
        # items each, so every stage of their rings is reused (two GEMM1 issuers once read A stages there that had
        # not been filled: six K chunks per item, five stages)
        m3, _, _ = _model(name, streaming=True, dtype=torch.bfloat16)
        m3.alloc_streams(1024, n)
        x3 = x.repeat(4, 1, 1).contiguous()
        y3 = torch.cat([m3(x3[..., c * n:(c + 1) * n].contiguous()) for c in range(x.shape[-1] // n)], -1)
        assert torch.equal(y3[2:], y3[:2].repeat(511, 1, 1))
        assert rel_l2(y3[:2].float().cpu().numpy(), a["y_stream"]) < 2 * BF16_TOL


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_host_pipeline_equals_device_forward(dtype):
    """rb_forward_host_async (blocks of streams on three streams, two calls back to back so the ring
    of device blocks wraps) returns exactly what forward() returns for the same inputs."""
    m, meta, a = _model("v2_c8", dtype=dtype)
    x0 = torch.from_numpy(golden_input(a))
    B0, _, T = x0.shape
    reps = 8 // B0 if B0 < 8 else 1
    xs = [torch.cat([x0 * (1.0 + 0.1 * (i + r)) for r in range(reps)]).to(dtype) for i in range(2)]
    want = [m(x.cuda()).cpu() for x in xs]
    xh = [x.contiguous().pin_memory() for x in xs]
    yh = [torch.empty_like(x).pin_memory() for x in xh]
    B = xh[0].shape[0]
    for i in range(2):
        m.forward_host_async(xh[i], yh[i], n_chunks=B if i == 0 else 2)
    m.host_sync()
    for i in range(2):
        assert torch.equal(yh[i], want[i])
    with pytest.raises(ValueError):
        m.forward_host_async(xh[0], yh[0], n_chunks=3)
    with pytest.raises(ValueError):
        m.forward_host_async(xs[0], yh[0])            # not pinned


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_adain_statistic_learning_vs_oracle(dtype):
    """AdaIN learn_y / learn_x running statistics and the transfer that follows (blocks.py:893-919),
    driven like ScriptedRAVE.update_adain (export.py:248-265); the oracle side of this sequence is
    pinned against the live reference modules in tests/test_oracle_pins.py."""
    from oracle.model import OracleRAVE
    from oracle.weights_recipe import recipe_input
    m, meta, _ = _model("v3_c8", dtype=dtype)
    o = OracleRAVE(meta["cfg"], golden_state_dict(meta))
    plan = [dict(reset_source=True, reset_target=True, learn_target=True), dict(learn_target=True),
            dict(learn_source=True), dict(learn_source=True), dict()]
    for i, kw in enumerate(plan):
        x = torch.from_numpy(recipe_input((3, 1, 8192), 40 + i))
        o.update_adain(**kw)
        m.update_adain(**kw)
        yo = o.forward(x)
        y = m(x.cuda().to(dtype)).float().cpu()
        if dtype == torch.float32:
            assert rel_l2(y.numpy(), yo.numpy()) < 1e-3, (i, kw)
    st = m.adain_state()
    tol = 2e-3 if dtype == torch.float32 else 0.25       # bf16: random-init Snake net is chaotic (see BF16_CHAOTIC)
    n_checked = 0
    for k, v in st.items():
        ref = o.sd[k]
        if "num_update" in k:
            assert float(v) == float(ref), k
        else:
            assert rel_l2(v[:3].numpy(), ref[:3].numpy()) < tol, k
        n_checked += 1
    assert n_checked >= 6


# ----------------------------------------------------------------------------- speaker encoder
def _speaker():
    from oracle import rave_oracle as O
    from oracle.weights_recipe import recipe_input
    from rave_b200.speaker import SpeakerRAVE
    meta, a = load_golden("speaker")
    sd = golden_state_dict(meta)
    _, hk = O.pqmf_bank(100, 16)
    wf, _ = O.pqmf_conv_weights(hk)
    enc = SpeakerRAVE(sd, pqmf_weight=wf)
    bands = recipe_input(tuple(int(v) for v in a["bands_shape"]), int(a["bands_seed"].item()), scale=1.0)
    audio = recipe_input(tuple(int(v) for v in a["audio_shape"]), int(a["audio_seed"].item()), scale=0.3)
    return enc, {k: torch.from_numpy(v) for k, v in sd.items()}, wf, a, bands, audio


def test_speaker_encoder_matches_golden_and_oracle():
    """CombinedRave.py:297-328 on the device vs the reference's outputs (golden) and the oracle."""
    from oracle import speaker_oracle as S
    from oracle.weights_recipe import recipe_input
    enc, sd, wf, a, bands, audio = _speaker()
    emb = enc(torch.from_numpy(bands).cuda())
    assert tuple(emb.shape) == (2, 256) and emb.dtype == torch.float32
    assert rel_l2(emb.cpu().numpy(), a["emb"]) < F32_TOL
    emb_a = enc.embed_audio(torch.from_numpy(audio).cuda())
    assert rel_l2(emb_a.cpu().numpy(), a["emb_audio"]) < F32_TOL
    # sizes the fixture does not hold: oracle on the same seeded input
    for B, T in ((1, 64), (3, 4096), (2, 8192 + 32)):
        x = torch.from_numpy(recipe_input((B, 16, T), 100 + T, scale=1.0))
        assert rel_l2(enc(x.cuda()).cpu().numpy(), S.speaker_embed(sd, x).numpy()) < F32_TOL, (B, T)
    x = torch.from_numpy(recipe_input((2, 1, 131072), 77, scale=0.3))     # model.py:30: 131072 samples
    assert rel_l2(enc.embed_audio(x.cuda()).cpu().numpy(), S.speaker_from_audio(sd, wf, x).numpy()) < F32_TOL
    assert enc.kernel_launches > 0


def test_speaker_encoder_errors_and_rows_independent():
    enc, sd, wf, a, bands, audio = _speaker()
    with pytest.raises(ValueError):
        enc(torch.zeros(1, 16, 100).cuda())           # not a multiple of the strides
    with pytest.raises(ValueError):
        enc(torch.zeros(1, 8, 1024).cuda())           # wrong band count
    with pytest.raises(ValueError):
        enc.embed_audio(torch.zeros(1, 1, 1000).cuda())
    x = torch.from_numpy(bands).cuda()
    both = enc(x)
    assert torch.equal(enc(x[1:2]), both[1:2])        # rows do not interact (eval-mode batch norm)


def test_set_speaker_from_audio_feeds_encode():
    """model.py:246-247 then :618-620: the embedding of the utterance is what encode() appends."""
    enc, sd, wf, a, bands, audio = _speaker()
    m, meta, ga = _model("v2_c8")
    assert m.cond_channels == 256
    m.load_speaker_encoder(golden_state_dict(load_golden("speaker")[0]))
    emb = m.set_speaker_from_audio(torch.from_numpy(audio).cuda())
    assert rel_l2(emb.cpu().numpy(), a["emb_audio"]) < F32_TOL
    z = m.encode(torch.from_numpy(golden_input(ga)).cuda())
    tail = z[:, -256:, :].cpu().numpy()
    assert np.array_equal(tail, np.broadcast_to(emb.cpu().numpy().reshape(1, 256, 1), tail.shape))


# ----------------------------------------------------------------------------- resampler
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("ratio", [2, 3])
def test_resampler_offline_vs_golden_and_oracle(ratio, dtype):
    """rave/resampler.py:58-66 on the device (ratios of the reference's tests/test_resampler.py)."""
    import os
    from oracle.resampler_oracle import ResamplerOracle
    from oracle.weights_recipe import recipe_input
    from rave_b200.resampler import Resampler
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "resampler.npz"))
    tol = 1e-5 if dtype == torch.float32 else 6e-3
    r = Resampler(ratio * 16000, 16000)
    assert np.array_equal(r.down_weight, d[f"w_down{ratio}"].reshape(-1))
    assert np.array_equal(r.up_weight, d[f"w_up{ratio}"].reshape(ratio, -1))
    x = torch.from_numpy(recipe_input(tuple(int(v) for v in d["x_shape"]), int(d["x_seed"].item()), scale=0.5))
    y = r.to_model_sampling_rate(x.to(dtype).cuda())
    assert y.dtype == dtype and tuple(y.shape) == d[f"down{ratio}"].shape
    assert rel_l2(y.float().cpu().numpy(), d[f"down{ratio}"]) < tol
    z = r.from_model_sampling_rate(torch.from_numpy(d[f"down{ratio}"]).to(dtype).cuda())
    assert tuple(z.shape) == tuple(x.shape)
    assert rel_l2(z.float().cpu().numpy(), d[f"up{ratio}"]) < tol
    # ragged lengths (not a multiple of the CTA tile), more streams: oracle on the same input
    o = ResamplerOracle(ratio)
    for B, T in ((1, 6 * ratio), (5, 1000 * ratio * 3 + 6 * ratio), (3, 2 ** 16 * ratio)):
        x = torch.from_numpy(recipe_input((B, 1, T), T, scale=0.5)).to(dtype)
        y = r.to_model_sampling_rate(x.cuda())
        yo = o.to_model_sampling_rate(x.float())
        assert rel_l2(y.float().cpu().numpy(), yo.numpy()) < tol, (B, T)
        z = r.from_model_sampling_rate(y)
        assert rel_l2(z.float().cpu().numpy(), o.from_model_sampling_rate(y.float().cpu()).numpy()) < tol, (B, T)
    with pytest.raises(ValueError):
        r.to_model_sampling_rate(torch.zeros(1, 1, ratio * 8 + 1).cuda())
    assert r.kernel_launches > 0


def test_resampler_streaming_chunk_by_chunk():
    """cached_conv mode: chunk-by-chunk against the golden stream of the real module and, with ragged
    chunk sizes (shorter than the history) and idle rows, against the oracle."""
    import os
    from oracle.resampler_oracle import ResamplerOracle
    from oracle.weights_recipe import recipe_input
    from rave_b200.resampler import Resampler
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "resampler.npz"))
    x = torch.from_numpy(recipe_input(tuple(int(v) for v in d["x_shape"]), int(d["x_seed"].item()), scale=0.5)).cuda()
    c = int(d["chunk"].item())
    r = Resampler(44100, 22050, streaming=True, max_streams=4)
    ys = [r.to_model_sampling_rate(x[..., i:i + c]) for i in range(0, x.shape[-1], c)]
    zs = [r.from_model_sampling_rate(y) for y in ys]
    assert rel_l2(torch.cat(ys, -1).cpu().numpy(), d["down2_stream"]) < 1e-5
    assert rel_l2(torch.cat(zs, -1).cpu().numpy(), d["up2_stream"]) < 1e-5
    r.reset()
    o = ResamplerOracle(2, streaming=True)
    xs = torch.from_numpy(recipe_input((4, 1, 8192), 3, scale=0.5))
    pos = 0
    for n in (2048, 16, 6, 30, 4096, 1996):
        xc = xs[..., pos:pos + n]
        pos += n
        y = r.to_model_sampling_rate(xc.cuda())
        yo = o.to_model_sampling_rate(xc)
        assert rel_l2(y.cpu().numpy(), yo.numpy()) < 1e-5, n
        z = r.from_model_sampling_rate(y)
        assert rel_l2(z.cpu().numpy(), o.from_model_sampling_rate(yo).numpy()) < 1e-5, n
    # a call with fewer rows must not disturb the idle rows' history
    r.reset()
    o = ResamplerOracle(2, streaming=True)
    a, b = xs[..., :1024], xs[..., 1024:2048]
    r.to_model_sampling_rate(a.cuda())
    o.to_model_sampling_rate(a)
    r.to_model_sampling_rate(xs[:2, :, 4096:4096 + 512].cuda())          # rows 0-1 only
    y = r.to_model_sampling_rate(b.cuda())
    yo_rows23 = o.to_model_sampling_rate(b)[2:]
    assert rel_l2(y[2:].cpu().numpy(), yo_rows23.numpy()) < 1e-5
    with pytest.raises(ValueError):
        Resampler(48000, 16000, streaming=True)


# ----------------------------------------------------------------------------- pitch front-end
def _f0_agreement(got, want, sr=44100):
    """Fraction of frames with the identical f0 value, and the worst period disagreement elsewhere."""
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    same = got == want
    both = (got > 0) & (want > 0) & ~same
    dtau = np.abs(sr / got[both] - sr / want[both]) if both.any() else np.zeros(1)
    flips = int(((got > 0) != (want > 0)).sum())
    return same.mean(), float(dtau.max()), flips


def test_yin_matches_golden_and_oracle():
    """pitch_utils.py:15-106 on the device.  The reference takes the autocorrelation from an fp32 FFT, the
    kernel from direct fp32 sums: periods are integers, so frames either agree exactly or differ where the
    reference's own rounding noise picks the lag; the test bounds how often and by how much."""
    import os
    from oracle import pitch_oracle as P
    from rave_b200 import pitch
    from .util import pitch_signals
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "pitch.npz"))
    x = torch.from_numpy(pitch_signals()).cuda()
    f0 = pitch.get_pitch(x, 1024)
    assert tuple(f0.shape) == d["f0"].shape
    same, dtau, flips = _f0_agreement(f0.cpu().numpy(), d["f0"])
    assert same >= 0.99 and dtau <= 1.0 and flips <= 1, (same, dtau, flips)        # measured: 1.0, 0, 0
    assert (f0[1] == 0).all()                                              # white noise: no period
    fw = pitch.estimate(x, 44100, 50., 800., 0.01, 0.15)
    assert tuple(fw.shape) == d["f0_wide"].shape
    same, dtau, flips = _f0_agreement(fw.cpu().numpy(), d["f0_wide"])
    assert same >= 0.99 and dtau <= 1.0 and flips <= 1, (same, dtau, flips)
    # other shapes: leading dims, a signal shorter than one frame, a low pitch floor (large tau_max)
    xs = torch.from_numpy(pitch_signals(32768))
    for args in ((44100, 30., 1000., 0.02, 0.2), (44100, 20., 2000., 0.05, 0.1), (22050, 70., 400., 0.01, 0.1)):
        ref = P.estimate(xs.reshape(2, 2, -1), *args)
        got = pitch.estimate(xs.reshape(2, 2, -1).cuda(), *args)
        assert got.shape == ref.shape
        same, dtau, flips = _f0_agreement(got.cpu().numpy(), ref.numpy(), args[0])
        # measured: identical except one frame of 52 at pitch_min = 20 Hz (4410-sample frames) that lands one lag away
        assert same >= 0.97 and dtau <= 1.001 and flips <= 1, (args, same, dtau, flips)
    short = xs[:1, :700]
    assert torch.equal(pitch.estimate(short.cuda(), 44100, 70., 400.).cpu() > 0, P.estimate(short, 44100, 70., 400.) > 0)
    with pytest.raises(ValueError):
        pitch.estimate(xs.cuda(), 44100, 70., 400., frame_stride=1e-6)


def test_f0_one_hot_matches_golden_and_oracle():
    """pitch_utils.py:116-135: f0 -> normalised log-f0 -> 257-class one-hot (class 256 = unvoiced)."""
    import os
    from oracle import pitch_oracle as P
    from rave_b200 import pitch
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "pitch.npz"))
    f0 = torch.from_numpy(d["f0"])
    oh, ln = pitch.f0_norm_one_hot(f0.cuda())
    assert tuple(oh.shape) == (4, f0.shape[1], 257) and torch.all(oh.sum(-1) == 1)
    idx = oh.argmax(-1).cpu().numpy()
    edge = np.abs(d["log_f0_norm"] * 256 - np.round(d["log_f0_norm"] * 256)) < 1e-3     # on a bin boundary
    ok = (idx == d["one_hot_idx"]) | edge
    assert ok.all()
    assert np.allclose(np.nan_to_num(ln.cpu().numpy(), nan=-9.), np.nan_to_num(d["log_f0_norm"], nan=-9.), atol=2e-6)
    cf, _ = pitch.f0_norm_one_hot(f0.cuda(), channels_first=True)
    assert torch.equal(cf, oh.permute(0, 2, 1))                            # scripts/export.py:359
    # out-of-range and unvoiced values: below 40 Hz wraps to the last class like upstream's eye() lookup
    f = torch.tensor([[0., 20., 39.9, 40., 41., 399., 2000., 30000.]])
    oh2, _ = pitch.f0_norm_one_hot(f.cuda())
    want, _ = P.f0_norm_one_hot(f)
    assert torch.equal(oh2.cpu(), want)


def test_from_run_loads_config_checkpoint_and_speaker_encoder(tmp_path):
    """scripts/export.py:541-568: run/config.gin + newest *last*.ckpt (state_dict or callbacks.EMA); the
    `speaker_encoder.*` tensors of the checkpoint build the speaker encoder (model.py:164-174)."""
    from rave_b200 import RAVE
    meta, a = load_golden("v2_c8")
    sd = {k: torch.from_numpy(v) for k, v in golden_state_dict(meta).items()}
    spk = {"speaker_encoder." + k: torch.from_numpy(v) for k, v in golden_state_dict(load_golden("speaker")[0]).items()}
    run = tmp_path / "myrun"
    (run / "version_0" / "checkpoints").mkdir(parents=True)
    (run / "config.gin").write_text('include "v2.gin"\nCAPACITY = 8\n')
    ema = {k: (v * 0.5 if k.endswith("weight_g") else v) for k, v in sd.items()       # parameters only (train.py:73-80)
           if k.endswith(("weight_g", "weight_v", "weight", "bias", "alpha"))}
    torch.save({"state_dict": {**sd, **spk}, "callbacks": {"EMA": ema}}, run / "version_0" / "checkpoints" / "epoch9-last.ckpt")
    m = RAVE.from_run(str(run))
    x = torch.from_numpy(golden_input(a)).cuda()
    assert rel_l2(m.encode(x).cpu().numpy(), a["z"]) < F32_TOL
    assert m.speaker_encoder is not None
    emb = m.speaker_encoder(torch.zeros(1, 16, 256).cuda())
    assert tuple(emb.shape) == (1, 256) and torch.isfinite(emb).all()
    m2 = RAVE.from_run(str(run), ema_weights=True)          # halved weight-norm gains: a different model
    with pytest.raises(RuntimeError):                       # EMA holds parameters only: no speaker vector yet
        m2.encode(x)
    m2.set_speaker(sd["speaker"])
    assert rel_l2(m2.encode(x).cpu().numpy()[:, :-256], a["z"][:, :-256]) > 0.1


# ----------------------------------------------------------------------------- latent processing
def test_latent_post_pre_processing_vs_oracle():
    """scripts/export.py:484-499 (PCA projection / crop, inverse with the noise draw) and blocks.py:926-956."""
    from oracle import latent_oracle as L
    from oracle.weights_recipe import recipe_input
    from rave_b200 import latent
    g = torch.Generator().manual_seed(9)
    C, n = 128, 16
    q, _ = torch.linalg.qr(torch.randn(C, C, generator=g))
    mean = 0.3 * torch.randn(C, generator=g)
    z = torch.from_numpy(recipe_input((3, C, 50), 31, scale=1.0))
    noise = torch.from_numpy(recipe_input((3, C - n, 50), 32, scale=1.0))
    pca = latent.LatentPCA(q, mean)
    y = pca.post_process_latent(z.cuda(), n)
    want = L.post_process_latent(z, q, mean, n)
    assert tuple(y.shape) == (3, n, 50) and rel_l2(y.cpu().numpy(), want.numpy()) < 1e-5
    back = pca.pre_process_latent(y, noise.cuda())
    assert rel_l2(back.cpu().numpy(), L.pre_process_latent(want, noise, q, mean).numpy()) < 1e-5
    zero = pca.pre_process_latent(y)
    assert rel_l2(zero.cpu().numpy(), L.pre_process_latent(want, torch.zeros_like(noise), q, mean).numpy()) < 1e-5
    with pytest.raises(ValueError):
        pca.post_process_latent(z[:, :64].cuda(), n)
    # spherical latents
    x = torch.from_numpy(recipe_input((4, 16, 70), 33, scale=1.0))
    x = x / x.norm(dim=1, keepdim=True)
    a = latent.unit_norm_vector_to_angles(x.cuda())
    a_ref = L.unit_norm_vector_to_angles(x.clone())
    assert tuple(a.shape) == (4, 15, 70) and float((a.cpu() - a_ref).abs().max()) < 2e-4   # acos is steep near +-1
    v = latent.angles_to_unit_norm_vector(a_ref.cuda())
    assert rel_l2(v.cpu().numpy(), L.angles_to_unit_norm_vector(a_ref.clone()).numpy()) < 1e-5
    assert rel_l2(latent.angles_to_unit_norm_vector(a).cpu().numpy(), x.numpy()) < 1e-4     # device round trip


@pytest.mark.parametrize("dtype,streaming", [(torch.float32, False), (torch.bfloat16, False), (torch.float32, True),
                                             (torch.bfloat16, True)])
def test_per_stream_conditioning_and_f0_channel_vs_oracle(dtype, streaming):
    """rb_set_conditioning: one speaker row per stream and a pitch channel behind the speaker channels
    (scripts/export.py:381-397 `z = cat(z, target_emb, f0_norm)`, rave/model.py:615-620), offline and chunk
    by chunk, against the oracle's encode(x, speaker, f0)."""
    from oracle.model import OracleRAVE
    from oracle.weights_recipe import build_state_dict, recipe_input
    from rave_b200 import RAVE
    name = "v2causal_c8" if streaming else "v2_c8"
    meta, _ = load_golden(name)
    ns = [(n, (s[0], 321, s[2]) if n == "decoder.net.0.weight_v" else tuple(s)) for n, s in meta["names_shapes"]]
    sd = build_state_dict(ns, meta["seed"])
    files, binds = GIN[name]
    m = RAVE.from_gin(files, binds + ["blocks.GeneratorV2.latent_size = 321"], streaming=streaming, dtype=dtype)
    assert m.spec.f0_channels == 1
    m.load_state_dict(sd)
    o = OracleRAVE(dict(meta["cfg"], dec_latent_size=321), sd, streaming=streaming)
    B, n_chunk, chunks = 3, 2048, 3 if streaming else 2
    T = n_chunk * chunks
    x = torch.from_numpy(recipe_input((B, 1, T), 91))
    spk = torch.from_numpy(recipe_input((B, 256, 1), 92, scale=1.0))
    f0 = torch.from_numpy(recipe_input((B, 1, T // 1024), 93, uniform=True)) * 0.5 + 0.5
    tol = F32_TOL if dtype == torch.float32 else 1.5e-2
    m.set_speaker(spk.cuda())                       # three different rows
    if not streaming:
        y = m.forward(x.cuda().to(dtype), f0=f0.cuda()).float().cpu().numpy()
        yo = o.forward(x, speaker=spk, f0=f0).numpy()
        assert rel_l2(y, yo) < tol
        z = m.encode(x.cuda().to(dtype), f0=f0.cuda()).float().cpu().numpy()
        zo = o.encode(x, speaker=spk, f0=f0).numpy()
        assert z.shape == zo.shape == (B, 321, T // 1024)
        assert rel_l2(z, zo) < tol
        assert np.array_equal(z[:, 64:320, 0], spk[:, :, 0].to(dtype).float().numpy())
        # a single row conditions every stream (the reference's one `speaker` buffer)
        m.set_speaker(spk[1])
        y1 = m.forward(x.cuda().to(dtype), f0=f0.cuda()).float().cpu().numpy()
        yo1 = o.forward(x, speaker=spk[1:2].repeat(B, 1, 1), f0=f0).numpy()
        assert rel_l2(y1, yo1) < tol
        # no pitch channels set: zeros
        m.set_f0(None)
        y0 = m.forward(x.cuda().to(dtype)).float().cpu().numpy()
        assert rel_l2(y0, o.forward(x, speaker=spk[1:2].repeat(B, 1, 1), f0=torch.zeros_like(f0)).numpy()) < tol
        with pytest.raises(ValueError):
            m.set_f0(torch.zeros(B, 2, T // 1024).cuda())
        m.set_speaker(spk[:2].cuda())               # two rows, three streams
        with pytest.raises(ValueError):
            m.forward(x.cuda().to(dtype))
        m.set_f0(f0[:, :, :2].cuda())               # frames of another call
        m.set_speaker(spk.cuda())
        with pytest.raises(ValueError):
            m.forward(x.cuda().to(dtype))
    else:
        for c in range(chunks):
            xs = x[..., c * n_chunk:(c + 1) * n_chunk].contiguous()
            fs = f0[..., c * 2:(c + 1) * 2].contiguous()
            y = m.forward(xs.cuda().to(dtype), f0=fs.cuda()).float().cpu().numpy()
            yo = o.forward(xs, speaker=spk, f0=fs).numpy()
            assert rel_l2(y, yo) < tol, c


@pytest.mark.parametrize("name,dtype", [("v2causal_c8", torch.bfloat16), ("v2causal_c64", torch.bfloat16),
                                        ("v2causal_c8", torch.float32)])
def test_stream_step_is_cuda_graph_capturable(name, dtype):
    """no hidden allocation or synchronisation in the hot call after rb_stream_alloc -- one
    rb_stream_step is captured in a CUDA graph (cudaStreamCaptureModeThreadLocal through torch) and the replays,
    chunk by chunk, return bit for bit what the eager steps return (the state arena advances in place)."""
    m, meta, a = _model(name, streaming=True, dtype=dtype)
    x = torch.from_numpy(golden_input(a)).cuda().to(dtype)
    n = int(a["chunk"].item())
    chunks = [x[..., c * n:(c + 1) * n].contiguous() for c in range(x.shape[-1] // n)]
    m.alloc_streams(64, n)
    eager = [m(c).clone() for c in chunks]
    m.reset_streams()
    xs = torch.zeros_like(chunks[0])
    m(xs)                                            # warm-up outside the capture (function attributes, tensor maps)
    m.reset_streams()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        ys = m(xs)
    m.reset_streams()                                # capture does not execute; start from zero state anyway
    for c, ye in zip(chunks, eager):
        xs.copy_(c)
        g.replay()
        assert torch.equal(ys, ye)
    torch.cuda.synchronize()


def test_discrete_latent_post_pre_processing_vs_oracle():
    """DiscreteScriptedRAVE.post_process_latent / pre_process_latent (scripts/export.py:503-517): float indices out,
    clamp + embedding sum back, against the oracle's RVQ (bit-exact indices, fp32 sums)."""
    from oracle.model import OracleRAVE
    from rave_b200.latent import DiscreteLatent
    m, meta, a = _model("discrete_c8")
    o = OracleRAVE(meta["cfg"], golden_state_dict(meta))
    dl = DiscreteLatent(m)
    z = torch.from_numpy(a["z"][:, :meta["cfg"]["latent_size"]]).contiguous()
    zi = dl.post_process_latent(z.cuda())
    assert zi.dtype == torch.float32 and zi.shape[1] == meta["cfg"]["num_quantizers"]
    assert torch.equal(zi.cpu(), o.rvq_encode(z).float())
    wild = zi.clone()
    wild[:, 0, 0] = -7.0                         # out-of-range values are clamped like the reference's
    wild[:, 1, 1] = 1e6
    zq = dl.pre_process_latent(wild).cpu()
    ref = o.rvq_decode(torch.clamp(wild.cpu(), 0, meta["cfg"]["codebook_size"] - 1).long())
    assert rel_l2(zq.numpy(), ref.numpy()) < 1e-6
    with pytest.raises(ValueError):
        DiscreteLatent(_model("v2_c8")[0])


def test_fake_stereo_forward_vs_oracle():
    """scripts/export.py:399-411: `z = cat([z, z], 0)` before the decoder, `y = cat(y.chunk(2, 0), 1)` after it."""
    from oracle.model import OracleRAVE
    m, meta, a = _model("v2_c8")
    o = OracleRAVE(meta["cfg"], golden_state_dict(meta))
    x = torch.from_numpy(golden_input(a))
    y = m.forward_stereo(x.cuda()).cpu()
    zo = o.encode(x)
    yo = torch.cat(o.decode(torch.cat([zo, zo], 0)).chunk(2, 0), 1)
    assert y.shape == (x.shape[0], 2, x.shape[2])
    assert rel_l2(y.numpy(), yo.numpy()) < 2 * F32_TOL


@pytest.mark.gpu
def test_folded_speaker_channels_vs_oracle():
    """bf16, offline, forward(): the speaker channels never reach the decoder's first conv as data -- they enter
    as a per-stream bias with the first / last row of each stream corrected for the zero padding (tc_engine.cu
    fold; rave/blocks.py:631-638, rave/model.py:615-620).  Per-stream rows, a changed table between calls (the
    bias is cached), stream micro-batches (row offsets), two-frame inputs (only edge rows) and one frame (the
    unfolded path) against the oracle; encode() still returns the conditioned latent."""
    from oracle.model import OracleRAVE
    from rave_b200 import RAVE
    from oracle.weights_recipe import recipe_input
    meta, _ = load_golden("v2_c8")
    sd = golden_state_dict(meta)
    files, binds = GIN["v2_c8"]
    m = RAVE.from_gin(files, binds, dtype=torch.bfloat16)
    m.load_state_dict(sd)
    o = OracleRAVE(meta["cfg"], sd)
    B = 5
    spk = torch.from_numpy(recipe_input((B, 256, 1), 192, scale=1.0))
    for frames in (6, 2, 1):
        T = 1024 * frames
        x = torch.from_numpy(recipe_input((B, 1, T), 191))
        xg = x.cuda().bfloat16()
        m.set_speaker(spk.cuda())
        y = m.forward(xg).float().cpu().numpy()
        assert rel_l2(y, o.forward(x, speaker=spk).numpy()) < 1.5e-2
        m.set_speaker(spk.flip(0).contiguous().cuda())        # same buffer, new contents
        y2 = m.forward(xg).float().cpu().numpy()
        assert rel_l2(y2, o.forward(x, speaker=spk.flip(0)).numpy()) < 1.5e-2
        m.set_microbatch(2)                                    # streams 0-1, 2-3, 4
        y3 = m.forward(xg).float().cpu().numpy()
        m.set_microbatch(0)
        assert np.array_equal(y3, y2)
        z = m.encode(xg).float().cpu().numpy()
        assert np.array_equal(z[:, 64:320, 0], spk.flip(0)[:, :, 0].bfloat16().float().numpy())
