"""ctypes binding of libstofnet_b200.so (the C ABI declared in include/stofnet_b200.h).

There is no CPU fallback: if the shared library is missing this module raises at import, and
every compute entry point returns STOF_E_CUDA without a CUDA device.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# STOFNET_B200_LIB: development override for A/B runs of two builds on one GPU box (tools/); the default is the in-tree library
LIB_PATH = os.environ.get("STOFNET_B200_LIB") or os.path.join(_HERE, "libstofnet_b200.so")

MAX_MID = 24
KIND_STOFNET, KIND_ESPCN, KIND_EDSR = 0, 1, 2
PATH_SIMT, PATH_TC, PATH_TC_FP16 = 0, 1, 2

E_INVALID, E_LENGTH, E_LENGTH_ODD, E_WORKSPACE, E_CUDA, E_NCCL, E_UNSUPPORTED = -1, -2, -3, -4, -5, -6, -7


class Weights(C.Structure):
    _fields_ = [
        ("kind", C.c_int),
        ("conv1_w", C.c_void_p), ("conv1_b", C.c_void_p),
        ("n_mid", C.c_int),
        ("mid_w", C.c_void_p * MAX_MID), ("mid_b", C.c_void_p * MAX_MID),
        ("last_w", C.c_void_p), ("last_b", C.c_void_p),
        ("contract_w", C.c_void_p), ("contract_b", C.c_void_p),
        ("expand_w", C.c_void_p), ("expand_b", C.c_void_p),
        ("num_features", C.c_int), ("upsample_factor", C.c_int),
        ("k_first", C.c_int), ("k_mid", C.c_int), ("k_last", C.c_int),
        ("sgb_scale", C.c_int), ("sgb_feat", C.c_int), ("sgb_k", C.c_int),
    ]


class Grads(C.Structure):
    _fields_ = [
        ("conv1_w", C.c_void_p), ("conv1_b", C.c_void_p),
        ("mid_w", C.c_void_p * MAX_MID), ("mid_b", C.c_void_p * MAX_MID),
        ("last_w", C.c_void_p), ("last_b", C.c_void_p),
        ("contract_w", C.c_void_p), ("contract_b", C.c_void_p),
        ("expand_w", C.c_void_p), ("expand_b", C.c_void_p),
    ]


if not os.path.exists(LIB_PATH):
    raise ImportError(
        "stofnet_b200: %s not found - build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "or `make -C stofnet_b200/csrc` (there is no CPU fallback)" % LIB_PATH)

lib = C.CDLL(LIB_PATH)

_vp, _i, _f, _sz, _d = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_double
_SIGS = {
    "stof_version": (C.c_char_p, []),
    "stof_last_error": (C.c_char_p, []),
    "stof_device_count": (_i, []),
    "stof_launch_count": (C.c_uint64, []),
    "stof_model_create": (_i, [C.POINTER(Weights), _i, C.POINTER(_vp)]),
    "stof_model_set_weights": (_i, [_vp, C.POINTER(Weights)]),
    "stof_model_set_weights_device": (_i, [_vp, C.POINTER(Weights), _vp]),
    "stof_model_destroy": (None, [_vp]),
    "stof_forward_workspace_bytes": (_i, [_vp, _i, _i, _i, C.POINTER(_sz)]),
    "stof_forward": (_i, [_vp, _vp, _i, _i, _vp, _vp, _sz, _i, _vp]),
    "stof_peaks": (_i, [_vp, _i, _i, _i, _i, _f, _i, _vp, _vp, _i, _vp, _vp, _vp]),
    "stof_peaks_indices": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "stof_nms_1d": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "stof_infer_host": (_i, [_vp, _vp, _i, _i, _i, _i, _f, _vp, _i, _vp, _vp, C.POINTER(_i)]),
    "stof_model_set_host_chunk_rows": (_i, [_vp, _i]),
    "stof_sample_shuffle": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
    "stof_model_set_path": (_i, [_vp, _i, _i]),
    "stof_model_set_chain_kernel": (_i, [_vp, _i]),
    "stof_chain_segment": (_i, [_vp, _i, _i, _vp, _vp, _vp, _i, _i, _vp]),
    "stof_debug_set_timeline": (_i, [_vp, _i]),
    "stof_sgb_expand": (_i, [_vp, _vp, _i, _i, _vp, _i, _vp]),
    "stof_profile_enable": (_i, [_vp, _i]),
    "stof_profile_collect": (_i, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    # operators either side of the path
    "stof_toa_rmse": (_i, [_vp, _i, _vp, _i, _i, _f, _vp, _vp]),
    "stof_coords2mask": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "stof_reduce_echoes": (_i, [_vp, _i, _i, _vp, _i, _i, _i, _f, _vp, _vp]),
    "stof_loss_scratch_bytes": (_sz, [_i, _i]),
    "stof_mask_loss": (_i, [_vp, _vp, _i, _i, _i, _vp, _i, _f, _f, _i, _vp, _vp, _vp, _vp, _vp]),
    "stof_mask_loss_static": (_i, [_vp, _vp, _i, _i, _i, _vp, _i, _f, _f, _i, _vp, _vp, _vp, _vp]),
    "stof_threshold_scratch_bytes": (_sz, [_i, _i]),
    "stof_find_threshold": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "stof_rf_upscale": (_i, [_vp, _vp, _i, _i, _d, _d, _d, _vp, _vp]),
    "stof_train_workspace_bytes": (_i, [_vp, _i, _i, C.POINTER(_sz)]),
    "stof_forward_train": (_i, [_vp, _vp, _i, _i, _vp, _vp, _sz, _vp]),
    "stof_backward": (_i, [_vp, _vp, _vp, _i, _i, _vp, _sz, C.POINTER(Grads), _vp]),
    "stof_adamw_step": (_i, [_vp, _vp, _vp, _vp, _sz, _i, _f, _f, _f, _f, _f, _vp]),
    "stof_adamw_step_dev": (_i, [_vp, _vp, _vp, _vp, _sz, _vp, _vp]),
    "stof_model_mark_weights_changed": (_i, [_vp]),
    "stof_thresholding": (_i, [_vp, _i, _i, _i, _f, _vp, _vp]),
    "stof_get_amplitudes": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "stof_reduce_echoes_pairs": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
    "stof_sgb_workspace_bytes": (_i, [_vp, _i, _i, C.POINTER(_sz)]),
    "stof_semi_global_block": (_i, [_vp, _vp, _i, _i, _vp, _vp, _sz, _vp]),
    "stof_attention_scratch_bytes": (_sz, [_i, _i, _i]),
    "stof_attention_block": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _f, C.c_uint64, _vp, _vp, _vp]),
}
for _name, (_res, _args) in _SIGS.items():
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args

EXPORTS = tuple(_SIGS)
PROF_CLASSES = 8
PROF_NAMES = ("conv_first", "contract_pool", "expand", "sgb_add", "conv_mid", "conv_last", "peaks", "other")


def last_error():
    return lib.stof_last_error().decode("utf-8", "replace")


def check(rc):
    """Map a C status to the exception the reference's torch path would raise."""
    if rc == 0:
        return
    msg = "stofnet_b200: %s (code %d)" % (last_error(), rc)
    if rc in (E_INVALID, E_UNSUPPORTED):
        raise ValueError(msg)
    raise RuntimeError(msg)
