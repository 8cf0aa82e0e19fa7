"""ctypes binding of the C-ABI in include/rave_b200.h (librave_b200.so, built in-tree).

There is deliberately no fallback: if the library is missing or no CUDA device is
present, construction raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librave_b200.so")

RB_OK, RB_ERR_INVALID, RB_ERR_CUDA, RB_ERR_STATE, RB_ERR_UNSUPPORTED = 0, -1, -2, -3, -4
RB_F32, RB_BF16 = 0, 1
(RB_W_CONV, RB_W_BIAS, RB_W_ALPHA, RB_W_AFFINE_A, RB_W_AFFINE_B, RB_W_COND, RB_W_CODEBOOK,
 RB_W_ADAIN_STATS) = range(8)
ABI_VERSION = 1


class RbSlot(C.Structure):
    _fields_ = [("channels", C.c_int32), ("per_frame", C.c_int32), ("hist", C.c_int32),
                ("reserved", C.c_int32)]


class RbOp(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "kind", "src", "dst", "res", "src2", "c_in", "c_out", "k", "stride", "dilation", "off",
        "ileave", "pre_act", "res_delay", "has_bias", "n_band", "src_ch_off")] + [("reserved", C.c_int32 * 3)]


class RbNetSpec(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("n_slots", C.c_int32), ("slots", C.POINTER(RbSlot)),
                ("n_enc_ops", C.c_int32), ("n_ops", C.c_int32), ("ops", C.POINTER(RbOp)),
                ("enc_in_slot", C.c_int32), ("enc_out_slot", C.c_int32),
                ("dec_in_slot", C.c_int32), ("dec_out_slot", C.c_int32),
                ("samples_per_frame", C.c_int32), ("streaming", C.c_int32),
                ("cond_channels", C.c_int32), ("noise_slot", C.c_int32),
                ("rvq_quantizers", C.c_int32), ("rvq_codebook_size", C.c_int32), ("rvq_dim", C.c_int32),
                ("reserved", C.c_int32 * 4)]


class RbTensorDesc(C.Structure):
    _fields_ = [("op", C.c_int32), ("role", C.c_int32), ("data", C.POINTER(C.c_float)),
                ("numel", C.c_int64)]


class RbDebugConv(C.Structure):
    _fields_ = [("op", RbOp * 2), ("n_ops", C.c_int32), ("B", C.c_int32), ("n_in", C.c_int32),
                ("w", C.POINTER(C.c_float) * 2), ("bias", C.POINTER(C.c_float) * 2),
                ("alpha", C.POINTER(C.c_float) * 2), ("x", C.POINTER(C.c_float)),
                ("res", C.POINTER(C.c_float)), ("y", C.POINTER(C.c_float)),
                ("fused_out", C.POINTER(C.c_int32)), ("flags", C.c_int32),
                ("trace", C.POINTER(C.c_uint64))]


EXPORTS = {
    "rb_debug_conv_bf16": [C.POINTER(RbDebugConv)],
    "rb_create": [C.POINTER(RbNetSpec), C.c_int, C.POINTER(C.c_void_p)],
    "rb_destroy": [C.c_void_p],
    "rb_load_weights": [C.c_void_p, C.POINTER(RbTensorDesc), C.c_int],
    "rb_encode": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_decode": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_forward": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_forward_host": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p],
    "rb_forward_host_async": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int],
    "rb_host_join": [C.c_void_p, C.c_void_p],
    "rb_host_sync": [C.c_void_p],
    "rb_set_noise": [C.c_void_p, C.c_void_p, C.c_int64],
    "rb_set_conditioning": [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p],
    "rb_rvq_encode": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_rvq_decode": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_stream_alloc": [C.c_void_p, C.c_int, C.c_int, C.c_int],
    "rb_stream_reset": [C.c_void_p, C.POINTER(C.c_int32), C.c_int],
    "rb_stream_encode": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_stream_decode": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_stream_step": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_stream_state_elements": [C.c_void_p, C.POINTER(C.c_int64)],
    "rb_workspace_bytes": [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)],
    "rb_kernel_launches": [C.c_void_p, C.POINTER(C.c_int64)],
    "rb_set_microbatch": [C.c_void_p, C.c_int],
    "rb_adain_update": [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int],
    "rb_adain_get_stats": [C.c_void_p, C.c_int, C.POINTER(C.c_float), C.c_int64, C.POINTER(C.c_int),
                           C.POINTER(C.c_int)],
    "rb_speaker_create": [C.c_int, C.POINTER(C.c_void_p)],
    "rb_speaker_destroy": [C.c_void_p],
    "rb_speaker_load": [C.c_void_p, C.c_int, C.POINTER(C.c_float), C.c_int64, C.POINTER(C.c_float), C.c_int64],
    "rb_speaker_load_bn": [C.c_void_p, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float),
                           C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_int, C.c_float],
    "rb_speaker_load_pqmf": [C.c_void_p, C.POINTER(C.c_float), C.c_int, C.c_int],
    "rb_speaker_embed": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_speaker_embed_audio": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_speaker_launches": [C.c_void_p, C.POINTER(C.c_int64)],
    "rb_speaker_embed_noalloc": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_resampler_create": [C.c_int, C.c_int, C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_float), C.c_int,
                            C.c_int, C.c_int, C.POINTER(C.c_void_p)],
    "rb_resampler_destroy": [C.c_void_p],
    "rb_resampler_reset": [C.c_void_p, C.c_void_p],
    "rb_resampler_to_model": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_resampler_from_model": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_resampler_launches": [C.c_void_p, C.POINTER(C.c_int64)],
    "rb_yin_frames": [C.c_int, C.c_int, C.c_double, C.c_double, C.POINTER(C.c_int)],
    "rb_yin_estimate": [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double,
                        C.c_void_p, C.c_int, C.c_void_p],
    "rb_f0_one_hot": [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p],
    "rb_latent_create": [C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_void_p)],
    "rb_latent_destroy": [C.c_void_p],
    "rb_latent_post": [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_latent_pre": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_unit_norm_to_angles": [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_angles_to_unit_norm": [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p],
    "rb_profile_ops": [C.c_void_p, C.c_int],
    "rb_op_times": [C.c_void_p, C.POINTER(C.c_float), C.c_int],
    "rb_last_error": [],
    "rb_abi_version": [],
}

_lib = None


def load_library():
    """dlopen the in-tree library and declare every entry point of the header."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m rave_b200.build` "
            "(rave_b200 has no CPU or PyTorch fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in EXPORTS.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = C.c_char_p if name == "rb_last_error" else C.c_int
    if lib.rb_abi_version() != ABI_VERSION:
        raise RuntimeError("librave_b200.so ABI version mismatch; rebuild")
    _lib = lib
    return lib


def check(rc):
    """C status -> the exception the reference's Python would raise."""
    if rc == RB_OK:
        return
    msg = (load_library().rb_last_error() or b"").decode(errors="replace")
    if rc == RB_ERR_INVALID:
        raise ValueError(msg)
    if rc == RB_ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise RuntimeError(msg)
