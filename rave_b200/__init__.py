"""rave_b200 -- B200-native (sm_100a) implementation of RAVE's encode/decode hot path.

Drop-in for `rave.RAVE.encode / decode / forward` of abargum/RAVE behind a C-ABI
(include/rave_b200.h, rave_b200/librave_b200.so).  See DESIGN.md.
"""
from . import gin_lite  # noqa: F401
from .netspec import NetSpec, build_netspec, get_padding  # noqa: F401


def __getattr__(name):
    # torch / ctypes are only imported when the model class is actually used
    if name in ("RAVE", "MAX_BATCH_SIZE"):
        from . import model
        return getattr(model, name)
    raise AttributeError(name)
