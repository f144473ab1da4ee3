"""ctypes binding of include/sbm_capi.h (libsbm_b200.so).  No fallback: a missing library is an error."""
import ctypes as C
import os

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SBM_LIB_PATH") or os.path.join(PKG, "libsbm_b200.so")  # SBM_LIB_PATH: variant builds (tools/build_variant.py)

TEMPLATE_DESC = np.dtype([("width", "<i4"), ("height", "<i4"), ("tl_x", "<i4"), ("tl_y", "<i4"), ("num_features", "<i4")])
FEATURE = np.dtype([("x", "<i4"), ("y", "<i4"), ("label", "<i4")])
MATCH_REC = np.dtype([("x", "<i4"), ("y", "<i4"), ("similarity", "<f4"), ("class_idx", "<i4"), ("template_id", "<i4")])
DEVICE_REC = np.dtype([("x", "<i4"), ("y", "<i4"), ("similarity", "<f4"), ("template_gid", "<u4"), ("order_key", "<u4")])

# every symbol include/sbm_capi.h declares (tests/test_capi_symbols.py checks the header against this list)
SYMBOLS = [
    "sbm_last_error", "sbm_build_info", "sbm_create", "sbm_destroy", "sbm_set_class_templates", "sbm_clear_templates",
    "sbm_num_templates", "sbm_match", "sbm_match_candidates", "sbm_match_device", "sbm_device_results",
    "sbm_finalize_records", "sbm_set_debug", "sbm_get_intermediate", "sbm_similarity_map", "sbm_similarity_local",
    "sbm_last_stats", "sbm_last_timing", "sbm_set_timing", "sbm_match_batch", "sbm_match_batch_device", "sbm_time_similarity", "sbm_time_kernels", "sbm_quantized_orientations", "sbm_merge_shard_records", "sbm_time_l2_read",
    "sbm_set_candidate_capacity", "sbm_group_unique_id", "sbm_group_create", "sbm_group_destroy", "sbm_group_info", "sbm_group_match",
    "sbm_group_match_batch", "sbm_group_block_bytes", "sbm_group_match_batch_device", "sbm_group_finalize", "sbm_phase_degrees", "sbm_quantized_level1_static",
]

GROUP_TEMPLATES, GROUP_TEMPLATES_CONTIGUOUS, GROUP_FRAMES = 0, 1, 2
GROUP_ID_BYTES = 128

SBM_QUANTIZED, SBM_SPREAD, SBM_RESPONSE, SBM_LINEAR, SBM_PYRAMID_SRC = range(5)

_lib = None


class SbmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("sbm error %d: %s" % (code, msg))
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "%s is missing: build it with `python -m shape_based_matching_b200.build` "
                "(there is no CPU or PyTorch fallback for the match path)" % LIB_PATH
            )
        L = C.CDLL(LIB_PATH)
        L.sbm_last_error.restype = C.c_char_p
        L.sbm_build_info.restype = C.c_char_p
        vp, ci, cf = C.c_void_p, C.c_int, C.c_float
        L.sbm_create.argtypes = [C.POINTER(ci), ci, cf, cf, ci, ci, C.POINTER(vp)]
        L.sbm_destroy.argtypes = [vp]
        L.sbm_set_class_templates.argtypes = [vp, ci, ci, vp, vp]
        L.sbm_clear_templates.argtypes = [vp]
        L.sbm_num_templates.argtypes = [vp, ci]
        margs = [vp, vp, ci, ci, ci, C.c_size_t, vp, cf, vp, ci, vp, ci, C.POINTER(ci)]
        L.sbm_match.argtypes = margs
        L.sbm_match_candidates.argtypes = margs
        L.sbm_match_device.argtypes = [vp, vp, ci, ci, ci, vp, cf, vp, ci, vp, ci, vp, vp]
        L.sbm_match_batch.argtypes = [vp, vp, vp, ci, ci, ci, ci, C.c_size_t, cf, vp, ci, vp, ci, vp]
        L.sbm_match_batch_device.argtypes = [vp, vp, C.c_size_t, ci, ci, ci, ci, cf, vp, ci, vp, ci, vp, vp]
        L.sbm_time_similarity.argtypes = [vp, cf, ci, C.POINTER(cf)]
        L.sbm_time_kernels.argtypes = [vp, cf, ci, vp, ci]
        L.sbm_quantized_orientations.argtypes = [vp, vp, ci, ci, ci, C.c_size_t, ci, vp, vp, vp, vp, C.POINTER(ci), C.POINTER(ci)]
        L.sbm_device_results.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(ci)]
        L.sbm_finalize_records.argtypes = [vp, vp, ci, vp, ci, ci, vp, ci, C.POINTER(ci)]
        L.sbm_time_l2_read.argtypes = [vp, C.c_size_t, ci, C.POINTER(cf)]
        L.sbm_merge_shard_records.argtypes = [vp, vp, ci, vp, ci, ci, vp, ci, ci, vp, ci, C.POINTER(ci)]
        L.sbm_set_candidate_capacity.argtypes = [vp, ci]
        L.sbm_group_unique_id.argtypes = [vp]
        L.sbm_group_create.argtypes = [vp, ci, ci, vp, ci]
        L.sbm_group_destroy.argtypes = [vp]
        L.sbm_group_info.argtypes = [vp, C.POINTER(ci), C.POINTER(ci), C.POINTER(ci)]
        L.sbm_group_match.argtypes = margs
        L.sbm_group_match_batch.argtypes = [vp, vp, vp, ci, ci, ci, ci, C.c_size_t, cf, vp, ci, vp, ci, vp]
        L.sbm_group_block_bytes.argtypes = [ci]
        L.sbm_group_block_bytes.restype = C.c_size_t
        L.sbm_group_match_batch_device.argtypes = [vp, vp, C.c_size_t, ci, ci, ci, ci, cf, vp, ci, vp, ci, vp]
        L.sbm_group_finalize.argtypes = [vp, vp, ci, ci, vp, ci, ci, vp, ci, vp]
        L.sbm_phase_degrees.argtypes = [ci, vp, vp, ci, vp, vp]
        L.sbm_quantized_level1_static.argtypes = [vp, ci, ci, ci, C.c_float, vp, vp]
        L.sbm_set_debug.argtypes = [vp, ci]
        L.sbm_set_timing.argtypes = [vp, ci]
        L.sbm_get_intermediate.argtypes = [vp, ci, ci, vp, C.c_size_t, C.POINTER(C.c_size_t)]
        L.sbm_similarity_map.argtypes = [vp, ci, ci, vp, C.c_size_t]
        L.sbm_similarity_local.argtypes = [vp, ci, ci, ci, ci, ci, vp]
        L.sbm_last_stats.argtypes = [vp, vp]
        L.sbm_last_timing.argtypes = [vp, vp]
        _lib = L
    return _lib


def prefer_bundled_nccl():
    """Before the first sbm_group_* call of a Python process: point the library's run-time NCCL binding (SBM_NCCL_LIB)
    at the libnccl.so.2 this Python environment ships (the nvidia-nccl wheel torch links against), if there is one and the
    caller has not chosen.  The C library by itself takes the copy already mapped into the process, else the system's --
    and a system NCCL older than torch's, loaded first under the same soname, would break a later `import torch`."""
    if os.environ.get("SBM_NCCL_LIB"):
        return
    try:
        import importlib.util

        spec = importlib.util.find_spec("nvidia.nccl")
        for base in (spec.submodule_search_locations if spec else []):
            cand = os.path.join(base, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                os.environ["SBM_NCCL_LIB"] = cand
                return
    except Exception:
        pass


def check(rc):
    if rc < 0:
        raise SbmError(rc, lib().sbm_last_error().decode())
    return rc
