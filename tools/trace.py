"""Pipeline event trace of CTA 0 of one tc_conv launch (through rb_debug_conv_bf16's trace buffer).
usage: python tools/trace.py du C dil | conv c_in c_out k [res]
Prints, per work item, the clock stamps (relative, in cycles) of:
  0 MMA: D_EMPTY ok   1 MMA: A_FULL ok(last chunk)   2 MMA: G1 issued   3 MMA: H_FULL ok
  4 MMA: D2_EMPTY ok  5 MMA: G2 issued   6 E1: D_FULL ok   7 E1: done   8 E2/out: start
  9 E2/out: done     10 A prod: A_EMPTY ok   11 A prod: issued
  12 out: acc barrier ok   13 out: first acc in registers   14 out: math done   15 out: stores issued
"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))  # repo root

import ctypes as C
import os
import sys

import numpy as np

from rave_b200 import capi


def ptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


def main():
    mode = sys.argv[1]
    rng = np.random.default_rng(0)
    lib = capi.load_library()
    d = capi.RbDebugConv()
    if mode == "du":
        Cc, dil = int(sys.argv[2]), int(sys.argv[3])
        B, n = int(os.environ.get("RB_TRACE_B", 48)), 16384 * 64 // Cc
        ops = [dict(c_in=Cc, c_out=Cc, k=3, stride=1, dilation=dil, off=dil, ileave=1, pre_act=1, has_bias=1),
               dict(c_in=Cc, c_out=Cc, k=1, stride=1, dilation=1, off=0, ileave=1, pre_act=1, has_bias=1)]
        ws = [rng.standard_normal((Cc, Cc, 3), dtype=np.float32) / np.sqrt(3 * Cc),
              rng.standard_normal((Cc, Cc, 1), dtype=np.float32) / np.sqrt(Cc)]
        res = None
        c_last = Cc
        flags = 3
    else:
        ci, co, k = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
        B, n = int(os.environ.get("RB_TRACE_B", 48)), 16384 * 64 // ci
        ops = [dict(c_in=ci, c_out=co, k=k, stride=1, dilation=1, off=k // 2, ileave=1, pre_act=1, has_bias=1)]
        ws = [rng.standard_normal((co, ci, k), dtype=np.float32) / np.sqrt(k * ci)]
        res = rng.standard_normal((B, co, n), dtype=np.float32) if len(sys.argv) > 5 else None
        c_last = co
        flags = 3 | (4 if res is not None else 0)
    bs = [np.zeros(o["c_out"], np.float32) for o in ops]
    for i, o in enumerate(ops):
        for name, val in o.items():
            setattr(d.op[i], name, int(val))
        d.w[i] = ptr(ws[i])
        d.bias[i] = ptr(bs[i])
    x = rng.standard_normal((B, ops[0]["c_in"], n), dtype=np.float32)
    y = np.zeros((B, c_last, n), np.float32)
    d.n_ops, d.B, d.n_in = len(ops), B, n
    d.x, d.y, d.res = ptr(x), ptr(y), ptr(res)
    d.flags = flags
    tr = np.zeros((64, 16), np.uint64)
    d.trace = tr.ctypes.data_as(C.POINTER(C.c_uint64))
    capi.check(lib.rb_debug_conv_bf16(C.byref(d)))
    t0 = tr[tr > 0].min()
    print("tile " + " ".join("%7s" % ("e%d" % e) for e in range(16)))
    last = max(t for t in range(64) if tr[t].any())
    for t in range(max(0, last - 14), last + 1):
        print("%4d " % t + " ".join("%7d" % (int(tr[t, e]) - int(t0)) if tr[t, e] else "      -" for e in range(16)))
    for e in (2, 5, 7, 9):
        v = tr[max(2, last - 24):last - 1, e].astype(np.int64)
        if (v > 0).all():
            print("event %d period: %.0f cycles/tile" % (e, np.diff(v).mean()))


if __name__ == "__main__":
    main()
