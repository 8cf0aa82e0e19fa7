"""-m gpu: the tcgen05 implicit-GEMM kernel, layer by layer, through the C-ABI testing hook
rb_debug_conv_bf16, against a torch fp32 reference of the same conv on bf16-rounded inputs
and weights (so the only differences are accumulation order and the bf16 rounding of the
stored result: rel-L2 tolerance 4e-3; the end-to-end bf16 bar of BASELINE.json is 1e-2)."""
import ctypes as C

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from rave_b200 import capi

pytestmark = pytest.mark.gpu
TOL = 4e-3


def _bf(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).bfloat16().float()


def _act(x, kind, alpha):
    if kind == 1:
        return F.leaky_relu(x, 0.2)
    if kind == 2:
        a = alpha.view(1, -1, 1)
        return x + (1.0 / (a + 1e-9)) * torch.sin(a * x) ** 2
    return x


def _ref_conv(x, w, b, k, stride, dil, off, ileave, pre_act, alpha):
    xa = _act(x, pre_act, alpha)
    n_out = x.shape[-1] // stride
    need = (n_out - 1) * stride + (k - 1) * dil + 1
    xp = F.pad(xa, (off, max(0, need - off - x.shape[-1])))
    y = F.conv1d(xp, w, None, stride=stride, dilation=dil)[..., :n_out]
    if ileave > 1:
        B, CO, n = y.shape
        y = y.view(B, ileave, CO // ileave, n).permute(0, 2, 3, 1).reshape(B, CO // ileave, n * ileave)
    if b is not None:
        y = y + b.view(1, -1, 1)
    return y


def _ptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


def _run(ops, ws, bs, als, x, res=None, flags=0):
    lib = capi.load_library()
    d = capi.RbDebugConv()
    keep = []
    for i, o in enumerate(ops):
        for name, val in o.items():
            setattr(d.op[i], name, int(val))
        for arr, field in ((ws, d.w), (bs, d.bias), (als, d.alpha)):
            a = None if arr[i] is None else np.ascontiguousarray(arr[i], dtype=np.float32)
            keep.append(a)
            field[i] = _ptr(a)
    d.n_ops, d.B, d.n_in = len(ops), x.shape[0], x.shape[2]
    xc = np.ascontiguousarray(x, dtype=np.float32)
    d.x = _ptr(xc)
    rc_ = None if res is None else np.ascontiguousarray(res, dtype=np.float32)
    d.res = _ptr(rc_)
    o0 = ops[0]
    r = max(o0.get("ileave", 1), 1)
    c_last = ops[-1]["c_out"] // (r if len(ops) == 1 else 1)
    n_out = x.shape[2] // o0.get("stride", 1) * r
    y = np.full((x.shape[0], c_last, n_out), np.nan, dtype=np.float32)
    d.y = _ptr(y)
    fused = C.c_int32(-1)
    d.fused_out = C.pointer(fused)
    d.flags = flags
    capi.check(lib.rb_debug_conv_bf16(C.byref(d)))
    return y, fused.value


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


CASES = [
    # c_in, c_out, k, stride, dil, off, ileave, pre_act, bias, residual, B, n_in
    dict(c_in=64, c_out=64, k=3, stride=1, dil=1, off=1, ileave=1, pre_act=0, bias=1, res=0, B=2, n=512),
    dict(c_in=64, c_out=64, k=3, stride=1, dil=9, off=9, ileave=1, pre_act=1, bias=1, res=1, B=3, n=200),
    dict(c_in=6, c_out=64, k=7, stride=1, dil=1, off=3, ileave=1, pre_act=0, bias=1, res=0, B=2, n=300),
    dict(c_in=64, c_out=32, k=7, stride=1, dil=1, off=3, ileave=1, pre_act=1, bias=1, res=0, B=2, n=1000),
    dict(c_in=64, c_out=128, k=8, stride=4, dil=1, off=3, ileave=1, pre_act=1, bias=1, res=0, B=2, n=1024),
    dict(c_in=64, c_out=128, k=8, stride=4, dil=1, off=7, ileave=1, pre_act=1, bias=1, res=0, B=2, n=1024),
    dict(c_in=256, c_out=512, k=4, stride=2, dil=1, off=1, ileave=1, pre_act=1, bias=1, res=0, B=2, n=768),
    dict(c_in=128, c_out=256, k=3, stride=1, dil=1, off=1, ileave=4, pre_act=1, bias=0, res=0, B=2, n=160),
    dict(c_in=1024, c_out=1024, k=3, stride=1, dil=1, off=1, ileave=2, pre_act=1, bias=0, res=0, B=2, n=96),
    dict(c_in=512, c_out=512, k=3, stride=1, dil=3, off=3, ileave=1, pre_act=1, bias=1, res=0, B=2, n=640),
    dict(c_in=512, c_out=512, k=1, stride=1, dil=1, off=0, ileave=1, pre_act=1, bias=1, res=1, B=2, n=640),
    dict(c_in=320, c_out=1024, k=3, stride=1, dil=1, off=1, ileave=1, pre_act=0, bias=1, res=0, B=3, n=70),
    dict(c_in=1024, c_out=64, k=3, stride=1, dil=1, off=1, ileave=1, pre_act=1, bias=1, res=0, B=2, n=130),
    dict(c_in=128, c_out=128, k=3, stride=1, dil=9, off=18, ileave=1, pre_act=2, bias=1, res=0, B=2, n=400),
    dict(c_in=96, c_out=192, k=8, stride=4, dil=1, off=3, ileave=1, pre_act=1, bias=1, res=0, B=1, n=512),
    dict(c_in=8, c_out=8, k=3, stride=1, dil=3, off=3, ileave=1, pre_act=1, bias=1, res=1, B=2, n=260),
]


@pytest.mark.parametrize("c", CASES, ids=lambda c: "ci{c_in}_co{c_out}_k{k}_s{stride}_d{dil}_il{ileave}_a{pre_act}".format(**c))
def test_single_conv(c):
    g = torch.Generator().manual_seed(c["c_in"] * 131 + c["k"] * 7 + c["dil"])
    x = _bf(torch.randn(c["B"], c["c_in"], c["n"], generator=g).numpy())
    w = _bf((torch.randn(c["c_out"], c["c_in"], c["k"], generator=g) / np.sqrt(c["c_in"] * c["k"])).numpy())
    r = c["ileave"]
    b = torch.randn(c["c_out"] // r, generator=g) * 0.1 if c["bias"] else None
    alpha = 1.0 + 0.1 * torch.randn(c["c_in"], generator=g) if c["pre_act"] == 2 else None
    ref = _ref_conv(x, w, b, c["k"], c["stride"], c["dil"], c["off"], r, c["pre_act"], alpha)
    res = None
    if c["res"]:
        res = _bf(torch.randn(ref.shape, generator=g).numpy())
        ref = ref + res
    op = dict(c_in=c["c_in"], c_out=c["c_out"], k=c["k"], stride=c["stride"], dilation=c["dil"], off=c["off"],
              ileave=r, pre_act=c["pre_act"], has_bias=c["bias"], res_delay=0)
    y, _ = _run([op], [w.numpy()], [None if b is None else b.numpy()], [None if alpha is None else alpha.numpy()],
                x.numpy(), None if res is None else res.numpy())
    assert np.isfinite(y).all()
    assert _rel(y, ref.numpy()) < TOL


@pytest.mark.parametrize("C_,dil,act,n", [(64, 1, 1, 512), (64, 9, 1, 333), (64, 3, 2, 1000), (16, 3, 1, 130),
                                          (32, 9, 1, 4096), (128, 3, 1, 300), (128, 9, 1, 1000), (128, 1, 2, 700),
                                          (256, 3, 1, 300)])
def test_dilated_unit_pair(C_, dil, act, n):
    """Residual(DilatedUnit): x + W1 act(W3 *_d act(x) + b3) + b1 (blocks.py:32-46,84-113)."""
    g = torch.Generator().manual_seed(C_ + dil)
    B = 3
    x = _bf(torch.randn(B, C_, n, generator=g).numpy())
    w3 = _bf((torch.randn(C_, C_, 3, generator=g) / np.sqrt(3 * C_)).numpy())
    w1 = _bf((torch.randn(C_, C_, 1, generator=g) / np.sqrt(C_)).numpy())
    b3, b1 = torch.randn(C_, generator=g) * 0.1, torch.randn(C_, generator=g) * 0.1
    a3 = 1.0 + 0.1 * torch.randn(C_, generator=g) if act == 2 else None
    a1 = 1.0 + 0.1 * torch.randn(C_, generator=g) if act == 2 else None
    h = _ref_conv(x, w3, b3, 3, 1, dil, dil, 1, act, a3)
    hb = h.bfloat16().float() if C_ > 128 else h         # unfused path stores h in bf16
    ref = x + _ref_conv(hb, w1, b1, 1, 1, 1, 0, 1, act, a1)
    ops = [dict(c_in=C_, c_out=C_, k=3, stride=1, dilation=dil, off=dil, ileave=1, pre_act=act, has_bias=1, res_delay=0),
           dict(c_in=C_, c_out=C_, k=1, stride=1, dilation=1, off=0, ileave=1, pre_act=act, has_bias=1, res_delay=0)]
    y, fused = _run(ops, [w3.numpy(), w1.numpy()], [b3.numpy(), b1.numpy()],
                    [None if a3 is None else a3.numpy(), None if a1 is None else a1.numpy()], x.numpy())
    assert fused == (1 if C_ <= 128 else 0)
    assert np.isfinite(y).all()
    assert _rel(y, ref.numpy()) < TOL


def _unleaky(z):
    """What the kernel recovers from a tensor stored as z = bf16(leaky(x)): min(z, 5 z) in fp32."""
    return torch.minimum(z, z * 5.0)


ACT_CASES = [
    # activated input (TMA feeds the MMA directly), activated output, activated residual
    dict(c_in=64, c_out=64, k=3, stride=1, dil=3, off=3, ileave=1, res=1, B=2, n=333, flags=7),
    dict(c_in=512, c_out=512, k=1, stride=1, dil=1, off=0, ileave=1, res=1, B=2, n=640, flags=5),
    dict(c_in=64, c_out=128, k=8, stride=4, dil=1, off=3, ileave=1, res=0, B=2, n=1024, flags=3),
    dict(c_in=256, c_out=512, k=3, stride=1, dil=1, off=1, ileave=4, res=0, B=2, n=160, flags=3),
    dict(c_in=512, c_out=512, k=3, stride=1, dil=9, off=9, ileave=1, res=0, B=2, n=700, flags=1),
    dict(c_in=64, c_out=32, k=7, stride=1, dil=1, off=3, ileave=1, res=0, B=2, n=1000, flags=1),
    # cta_group::2 pairs (streamed weights, 256-wide output tiles, >= 256 rows): odd tile counts, a pair whose
    # second tile lies beyond the tensor, residual, two n-tiles, strided input view, transposed-conv phases
    dict(c_in=256, c_out=256, k=3, stride=1, dil=3, off=3, ileave=1, res=0, B=3, n=1000, flags=3),
    dict(c_in=512, c_out=512, k=1, stride=1, dil=1, off=0, ileave=1, res=1, B=1, n=300, flags=7),
    dict(c_in=256, c_out=256, k=1, stride=1, dil=1, off=0, ileave=1, res=1, B=5, n=2048, flags=7),
    dict(c_in=512, c_out=1024, k=4, stride=2, dil=1, off=1, ileave=1, res=0, B=2, n=1024, flags=3),
    dict(c_in=1024, c_out=1024, k=3, stride=1, dil=1, off=1, ileave=2, res=0, B=2, n=300, flags=3),
    # short tensors: the pair takes two neighbouring streams (odd stream count: one CTA of the last pair idles)
    dict(c_in=512, c_out=512, k=3, stride=1, dil=1, off=1, ileave=1, res=0, B=3, n=260, flags=3),
    dict(c_in=512, c_out=1024, k=4, stride=2, dil=1, off=1, ileave=1, res=0, B=5, n=512, flags=3),
    # split accumulator (MT = 2: the sub-tiles are published / taken back separately), several work items per
    # pair so that every barrier goes through both parities, with and without a residual
    dict(c_in=512, c_out=512, k=3, stride=1, dil=3, off=3, ileave=1, res=1, B=40, n=1024, flags=7),
    dict(c_in=256, c_out=512, k=4, stride=2, dil=1, off=1, ileave=1, res=0, B=64, n=2048, flags=3),
    # single-tap cta_group::2 layers: several items per pair, and fewer work slots than pairs
    dict(c_in=512, c_out=512, k=1, stride=1, dil=1, off=0, ileave=1, res=1, B=48, n=1024, flags=7),
    dict(c_in=512, c_out=1024, k=1, stride=1, dil=1, off=0, ileave=1, res=0, B=3, n=384, flags=3),
    # sector-complete stores of phase outputs (drain_pair), the second one also as a cta_group::2 pair
    dict(c_in=128, c_out=256, k=3, stride=1, dil=1, off=1, ileave=4, res=0, B=2, n=520, flags=3),
    dict(c_in=256, c_out=256, k=3, stride=1, dil=1, off=1, ileave=2, res=0, B=3, n=640, flags=3),
    # ... and with two n-tiles that each hold two of the four phases (the 256->128 x4 transposed conv)
    dict(c_in=256, c_out=512, k=3, stride=1, dil=1, off=1, ileave=4, res=0, B=3, n=520, flags=3),
    dict(c_in=6, c_out=64, k=7, stride=1, dil=1, off=3, ileave=1, res=0, B=2, n=4096, flags=2, pre=0),
]


@pytest.mark.parametrize("c", ACT_CASES, ids=lambda c: "ci{c_in}_co{c_out}_k{k}_s{stride}_il{ileave}_f{flags}_B{B}".format(**c))
def test_single_conv_activated_storage(c):
    """LeakyReLU nets keep tensors as leaky(x): the conv input needs no pass, a residual is
    inverted with min(z, 5z), the output is activated in the epilogue before the bf16 store."""
    g = torch.Generator().manual_seed(c["c_in"] * 17 + c["k"])
    z = _bf(F.leaky_relu(torch.randn(c["B"], c["c_in"], c["n"], generator=g), 0.2).numpy())   # stored input
    w = _bf((torch.randn(c["c_out"], c["c_in"], c["k"], generator=g) / np.sqrt(c["c_in"] * c["k"])).numpy())
    r = c["ileave"]
    b = torch.randn(c["c_out"] // r, generator=g) * 0.1
    ref = _ref_conv(z, w, b, c["k"], c["stride"], c["dil"], c["off"], r, 0, None)
    res = None
    if c["res"]:
        res = _bf(F.leaky_relu(torch.randn(ref.shape, generator=g), 0.2).numpy())
        ref = ref + (_unleaky(res) if c["flags"] & 4 else res)
    if c["flags"] & 2:
        ref = F.leaky_relu(ref, 0.2)
    op = dict(c_in=c["c_in"], c_out=c["c_out"], k=c["k"], stride=c["stride"], dilation=c["dil"], off=c["off"],
              ileave=r, pre_act=c.get("pre", 1), has_bias=1, res_delay=0)
    y, _ = _run([op], [w.numpy()], [b.numpy()], [None], z.numpy(), None if res is None else res.numpy(),
                flags=c["flags"])
    assert np.isfinite(y).all()
    assert _rel(y, ref.numpy()) < TOL


@pytest.mark.parametrize("C_,dil,n", [(64, 9, 777), (32, 1, 256), (128, 3, 300), (128, 9, 2000),
                                      # C = 256: both convs in one cluster-pair launch (tc_du_pair_kernel): several
                                      # tiles per CTA, a pair whose second tile is past the end, one short tile
                                      (256, 3, 2048), (256, 9, 900), (256, 1, 100), (256, 3, 4096),
                                      (128, 1, 4096), (128, 9, 1100)])
def test_dilated_unit_pair_activated_storage(C_, dil, n):
    _dilated_unit_pair_activated(C_, dil, n, 2 if C_ == 256 else 1)


@pytest.mark.parametrize("C_,dil,n", [(128, 1, 4096), (128, 9, 1100), (128, 3, 300), (64, 3, 2000), (64, 9, 130)])
def test_dilated_unit_cluster_pair_two_tmem_stages(C_, dil, n, monkeypatch):
    """tc_du_pair_kernel with two tiles in tensor memory (C <= 128; selected with RB_TC_PD128 / RB_TC_PD64 -- the
    default for those widths is the one-CTA fused kernel, which measures faster): odd tile counts, one-item CTAs."""
    monkeypatch.setenv("RB_TC_PD128", "1")
    monkeypatch.setenv("RB_TC_PD64", "1")
    _dilated_unit_pair_activated(C_, dil, n, 2)


def _dilated_unit_pair_activated(C_, dil, n, want_fused):
    g = torch.Generator().manual_seed(C_ * 3 + dil)
    B = 3 if C_ >= 128 else 2
    z = _bf(F.leaky_relu(torch.randn(B, C_, n, generator=g), 0.2).numpy())
    w3 = _bf((torch.randn(C_, C_, 3, generator=g) / np.sqrt(3 * C_)).numpy())
    w1 = _bf((torch.randn(C_, C_, 1, generator=g) / np.sqrt(C_)).numpy())
    b3, b1 = torch.randn(C_, generator=g) * 0.1, torch.randn(C_, generator=g) * 0.1
    h = _ref_conv(z, w3, b3, 3, 1, dil, dil, 1, 0, None)
    hb = F.leaky_relu(h.bfloat16().float() if C_ > 128 else h, 0.2)  # unfused path stores h in bf16
    ref = F.leaky_relu(_unleaky(z) + _ref_conv(hb, w1, b1, 1, 1, 1, 0, 1, 0, None), 0.2)
    ops = [dict(c_in=C_, c_out=C_, k=3, stride=1, dilation=dil, off=dil, ileave=1, pre_act=1, has_bias=1, res_delay=0),
           dict(c_in=C_, c_out=C_, k=1, stride=1, dilation=1, off=0, ileave=1, pre_act=1, has_bias=1, res_delay=0)]
    y, fused = _run(ops, [w3.numpy(), w1.numpy()], [b3.numpy(), b1.numpy()], [None, None], z.numpy(), flags=3)
    assert fused == want_fused
    assert np.isfinite(y).all()
    assert _rel(y, ref.numpy()) < TOL
