"""ctypes binding of libtmdiff_b200.so (include/tmdiff_b200.h).  There is no CPU fallback:
if the shared library is missing or no CUDA device is present, calls raise."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# TMDIFF_B200_LIB: load another build of the same library (A/B runs of kernel variants on one box, profiles/_variant.py)
LIB_PATH = os.environ.get("TMDIFF_B200_LIB") or os.path.join(_HERE, "lib", "libtmdiff_b200.so")

F32, BF16 = 0, 1
SENSOR_IDS = {"QB": 0, "WV3": 1, "GF2": 2, "WV2": 3, "WV4": 4}
NUM_SLOTS = 8


class Config(C.Structure):
    _fields_ = [("channels", C.c_int32 * 4), ("embed_dim", C.c_int32), ("inter_dim", C.c_int32),
                ("max_batch", C.c_int32), ("max_bands", C.c_int32), ("height", C.c_int32), ("width", C.c_int32),
                ("dtype", C.c_int32)]


class Eval(C.Structure):
    _fields_ = [("src", C.c_int32), ("dst", C.c_int32), ("t_input", C.c_float), ("alpha_t", C.c_float),
                ("sigma_t", C.c_float)]


class Update(C.Structure):
    _fields_ = [("dst", C.c_int32), ("x", C.c_int32), ("m", C.c_int32), ("m1", C.c_int32),
                ("c1", C.c_float), ("c2", C.c_float), ("c3", C.c_float), ("form", C.c_int32), ("m2", C.c_int32),
                ("k", C.c_float * 5)]


class Step(C.Structure):
    _fields_ = [("kind", C.c_int32), ("eval", Eval), ("update", Update)]


class Plan(C.Structure):
    _fields_ = [("steps", C.POINTER(Step)), ("n_steps", C.c_int32), ("result_slot", C.c_int32),
                ("q_weight", C.c_float), ("q_lo", C.c_int32), ("max_val", C.c_float), ("model_type", C.c_int32),
                ("cfg", C.c_int32), ("guidance_scale", C.c_float)]


MODEL_X_START, MODEL_NOISE = 0, 1


KIND_NAMES = ("conv3x3x3", "conv1x1x1", "elementwise", "solver", "conditioning")


class ProfileKind(C.Structure):
    _fields_ = [("ms", C.c_double), ("flops", C.c_double), ("bytes", C.c_double), ("launches", C.c_int64),
                ("flops_executed", C.c_double)]


class Profile(C.Structure):
    _fields_ = [("kind", ProfileKind * len(KIND_NAMES))]


# name -> (restype, argtypes); every symbol declared in include/tmdiff_b200.h
_P = C.c_void_p
SIGNATURES = {
    "tmd_last_error": (C.c_char_p, []),
    "tmd_version": (C.c_int, []),
    "tmd_create": (C.c_int, [C.POINTER(Config), C.POINTER(_P)]),
    "tmd_destroy": (C.c_int, [_P]),
    "tmd_load_weight": (C.c_int, [_P, C.c_char_p, _P, C.POINTER(C.c_int64), C.c_int32]),
    "tmd_set_time_frequencies": (C.c_int, [_P, _P, C.c_int32]),
    "tmd_set_text_embedding": (C.c_int, [_P, C.c_int32, _P]),
    "tmd_finalize_weights": (C.c_int, [_P]),
    "tmd_precompute_condition": (C.c_int, [_P, _P, _P, C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_int32,
                                           C.c_int32, _P]),
    "tmd_denoise": (C.c_int, [_P, _P, _P, _P, _P]),
    "tmd_sample": (C.c_int, [_P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(Plan), _P, _P, _P, _P]),
    "tmd_dwt2d": (C.c_int, [_P, _P, _P, _P, _P, C.c_int64, C.c_int32, C.c_int32, _P]),
    "tmd_idwt2d": (C.c_int, [_P, _P, _P, _P, _P, C.c_int64, C.c_int32, C.c_int32, _P]),
    "tmd_conv3d_mod": (C.c_int, [_P, _P, _P, _P, _P] + [C.c_int32] * 8 + [_P]),
    "tmd_linear": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P]),
    "tmd_dynamic_threshold": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_float, _P]),
    "tmd_dynamic_threshold_static": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_float, _P]),
    "tmd_groupnorm": (C.c_int, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_float, C.c_int32, _P]),
    "tmd_attention": (C.c_int, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_float, _P]),
    "tmd_attention_bf16": (C.c_int, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_float, _P]),
    "tmd_attention_bf16_strided": (C.c_int, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int64,
                                             C.c_int32, C.c_int64, C.c_int32, C.c_float, _P]),
    "tmd_attention_bf16_tokens": (C.c_int, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_int64,
                                            C.c_int64, C.c_int64, C.c_int32, C.c_float, _P]),
    "tmd_dynamic_conv3d": (C.c_int, [_P] * 6 + [C.c_int32] * 9 + [C.c_float, C.c_int32, _P]),
    "tmd_quality_metrics": (C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_float, _P]),
    "tmd_ancestral_update": (C.c_int, [_P, _P, _P, _P] + [C.c_float] * 5 + [C.c_int32, C.c_uint64, C.c_uint64, C.c_int64, _P]),
    "tmd_dynamic_conv": (C.c_int, [_P] * 7 + [C.c_int32] * 10 + [C.c_float, C.c_int32, _P]),
    "tmd_dynamic_conv_workspace_bytes": (C.c_int64, [C.c_int32] * 9),
    "tmd_dynamic_conv_ws": (C.c_int, [_P] * 7 + [C.c_int32] * 10 + [C.c_float, C.c_int32, _P, C.c_int64, _P]),
    "tmd_layernorm": (C.c_int, [_P, _P, _P, _P, C.c_int32, C.c_int32, C.c_float, _P]),
    "tmd_geglu": (C.c_int, [_P, _P, C.c_int64, C.c_int32, _P]),
    "tmd_band_moments": (C.c_int, [_P, _P, _P] + [C.c_int32] * 6 + [_P]),
    "tmd_psf_downsample": (C.c_int, [_P, _P, _P, _P] + [C.c_int32] * 7 + [_P]),
    "tmd_scene_to_tiles": (C.c_int, [_P, _P, _P, _P] + [C.c_int32] * 7 + [C.c_float, _P]),
    "tmd_tiles_to_scene": (C.c_int, [_P, _P] + [C.c_int32] * 6 + [C.c_float, C.c_int32, _P]),
    "tmd_debug_check_redzones": (C.c_int, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "tmd_launch_count": (C.c_int64, [_P]),
    "tmd_profile_begin": (C.c_int, [_P]),
    "tmd_profile_end": (C.c_int, [_P, C.POINTER(Profile), C.c_char_p]),
}

_lib = None


def load():
    """Load the shared library (no device work).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libtmdiff_b200.so is not built (run `python -m tmdiff_b200.build`); "
                               "tmdiff_b200 has no CPU or PyTorch fallback")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc):
    if rc != 0:
        raise RuntimeError("tmdiff_b200: " + load().tmd_last_error().decode())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    """Device/host pointer of a contiguous tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())
