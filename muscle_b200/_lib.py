"""ctypes binding of libmuscle_b200.so (the C ABI declared in include/muscle_b200.h).

There is no CPU or eager-PyTorch fallback: if the library is missing, or the device is not an
sm_100 GPU, every op raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_double, c_float, c_int, c_int64, c_size_t, c_uint32, c_void_p, POINTER

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libmuscle_b200.so")

MUSCLE_OK = 0
LAYOUT_PIXEL_MAJOR = 0
LAYOUT_NCHW = 1
MAX_CLASSES = 31

# name -> (restype, argtypes); mirrors include/muscle_b200.h one to one
_SIGNATURES = {
    "muscle_last_error": (c_char_p, []),
    "muscle_version": (c_int, []),
    "muscle_pad_rows": (c_int64, [c_int64]),
    "muscle_dense_infonce_workspace_bytes": (c_size_t, [c_int64, c_int, c_int]),
    "muscle_dense_infonce_fwd_bwd": (c_int, [c_void_p, c_int, c_int64, c_int64, c_int, c_void_p, c_int, c_int,
                                             c_float, c_float, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "muscle_dense_infonce_fwd": (c_int, [c_void_p, c_int, c_int64, c_int64, c_int, c_void_p, c_int, c_int,
                                         c_float, c_float, c_void_p, c_void_p, c_size_t, c_void_p]),
    "muscle_dense_infonce_bwd": (c_int, [c_int, c_int64, c_int64, c_int, c_int, c_float, c_void_p, c_void_p,
                                         c_void_p, c_size_t, c_void_p]),
    "muscle_dense_infonce_proto_fwd": (c_int, [c_void_p, c_int, c_int64, c_int64, c_int, c_void_p, c_int, c_int,
                                               c_float, c_float, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "muscle_dense_infonce_proto_bwd": (c_int, [c_int, c_int64, c_int64, c_int, c_int, c_float, c_void_p, c_int, c_void_p,
                                               c_void_p, c_void_p, c_size_t, c_void_p]),
    "muscle_infonce_prepare_scratch_bytes": (c_size_t, [c_int64]),
    "muscle_infonce_prepare": (c_int, [c_void_p, c_int, c_int64, c_int64, c_int, c_void_p, c_int, c_int, c_float,
                                       c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_void_p, c_size_t, c_void_p]),
    "muscle_infonce_plan": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_void_p, c_void_p]),
    "muscle_infonce_rowsum": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64,
                                      c_void_p, c_int, c_float, c_void_p, c_void_p]),
    "muscle_infonce_plan_sym": (c_int, [c_void_p, c_int64, c_void_p, c_void_p]),
    "muscle_infonce_rowsum_sym": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int, c_float, c_void_p,
                                          c_void_p]),
    "muscle_infonce_plan_sym_ring_scratch_bytes": (c_size_t, [c_int64]),
    "muscle_infonce_plan_sym_ring": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "muscle_infonce_rowsum_sym_ring": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64,
                                               c_void_p, c_int, c_int, c_int, c_float, c_void_p, c_void_p]),
    "muscle_infonce_plan_sym_ring_static_max_rows": (c_int, []),
    "muscle_infonce_plan_sym_ring_static": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p]),
    "muscle_infonce_rowsum_sym_share": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int, c_int, c_int, c_float,
                                                c_void_p, c_void_p]),
    "muscle_infonce_mid": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                   c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "muscle_infonce_part_slots": (c_int, []),
    "muscle_infonce_grad": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p,
                                    c_void_p, c_int64, c_void_p, c_int, c_float, c_void_p, c_void_p]),
    "muscle_infonce_final": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64,
                                     c_int64, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_float, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "muscle_pool_prototypes_fwd": (c_int, [c_void_p, c_int, c_void_p, c_int64, c_int, c_int, c_float, c_void_p, c_void_p,
                                           c_void_p, c_void_p]),
    "muscle_pool_prototypes_fwd_scratch_bytes": (c_size_t, [c_int, c_int]),
    "muscle_pool_prototypes_fwd_ws": (c_int, [c_void_p, c_int, c_void_p, c_int64, c_int, c_int, c_float, c_void_p, c_void_p,
                                              c_void_p, c_void_p, c_size_t, c_void_p]),
    "muscle_pool_prototypes_bwd": (c_int, [c_void_p, c_int, c_void_p, c_int64, c_int, c_int, c_float, c_void_p, c_void_p,
                                           c_void_p, c_void_p, c_void_p, c_void_p]),
    "muscle_imc_workspace_bytes": (c_size_t, [c_int, c_int]),
    "muscle_imc_fwd_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_size_t, c_void_p]),
    "muscle_cls_losses_fwd_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_float, c_void_p,
                                          c_void_p, c_void_p, c_void_p]),
    "muscle_focal_loss_fwd_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_float, c_float, c_void_p, c_void_p,
                                          c_void_p, c_void_p]),
    "muscle_lsep_fwd_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "muscle_upsample_bilinear_fwd": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "muscle_upsample_bilinear_bwd_scratch_bytes": (c_size_t, [c_int64, c_int, c_int]),
    "muscle_upsample_bilinear_bwd": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_size_t,
                                             c_void_p]),
    "muscle_er_loss_state_bytes": (c_size_t, [c_int]),
    "muscle_er_loss_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_size_t,
                                   c_void_p]),
    "muscle_er_loss_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p,
                                   c_void_p, c_void_p]),
    "muscle_pixpro_fwd_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                      c_void_p, c_void_p, c_void_p, c_void_p]),
    "muscle_cam_head_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "muscle_cam_head_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                    c_void_p]),
    "muscle_gap_fwd": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "muscle_linear_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "muscle_linear_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "muscle_cam_softmaxnorm_fwd": (c_int, [c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p]),
    "muscle_cam_softmaxnorm_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p]),
    "muscle_cam_maxnorm_fwd": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "muscle_cam_maxnorm_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "muscle_pcm_state_bytes": (c_size_t, [c_int, c_int, c_int64]),
    "muscle_pcm_bwd_scratch_bytes": (c_size_t, [c_int, c_int, c_int64]),
    "muscle_pcm_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_void_p,
                               c_void_p, c_size_t, c_void_p]),
    "muscle_pcm_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_void_p, c_size_t,
                               c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "muscle_field_edges": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_int, c_void_p, c_void_p,
                                   c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "muscle_field_edges_sobel": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_int, c_int, c_void_p,
                                         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "muscle_field_compact": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p,
                                     c_void_p]),
    "muscle_field_loss_scratch_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "muscle_field_loss": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "muscle_field_loss_decoder": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                          c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_size_t,
                                          c_void_p]),
    "muscle_conv1x1_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_void_p, c_void_p]),
    "muscle_irn_transition_init": (c_int, [c_void_p, c_int, c_float, c_void_p, c_void_p, c_void_p]),
    "muscle_irn_mask_edge": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "muscle_irn_split3_bytes": (c_size_t, [c_int, c_int]),
    "muscle_irn_split3": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "muscle_irn_matmul_split3": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "muscle_crops_fwd": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int64, c_void_p,
                                 c_void_p]),
    "muscle_crops_bwd": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int64, c_void_p,
                                 c_void_p]),
    "muscle_emd_scratch_bytes": (c_size_t, [c_int, c_int, c_int64, c_int]),
    "muscle_emd_dynamic_fwd_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_int64,
                                           c_int64, c_float, c_int, c_void_p, c_void_p, c_int64, c_void_p, c_size_t,
                                           c_void_p]),
    "muscle_emd_static_scratch_bytes": (c_size_t, [c_int, c_int]),
    "muscle_emd_static_fwd_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                          c_int64, c_void_p, c_size_t, c_void_p]),
    "muscle_info_nce_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int]),
    "muscle_info_nce_fwd_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_float, c_int,
                                        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                        c_void_p]),
    "muscle_info_nce_fwd_bwd_ex": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_float, c_int,
                                           c_float, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                           c_void_p, c_size_t, c_void_p]),
}

_lib = None


def exported_symbols():
    return sorted(_SIGNATURES)


def load():
    """Loads the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "muscle_b200: %s not found. Build it with `python -m muscle_b200.build` "
            "(needs nvcc); there is no CPU fallback." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


DIAG_LIB_PATH = os.path.join(_HERE, "lib", "libmuscle_b200_diag.so")
_diag = None


def load_diag():
    """The test fixture library (include/muscle_b200_diag.h): one-tile checks of the tcgen05 building blocks."""
    global _diag
    if _diag is None:
        if not os.path.exists(DIAG_LIB_PATH):
            raise RuntimeError("muscle_b200: %s not found. Build it with `python -m muscle_b200.build`." % DIAG_LIB_PATH)
        lib = ctypes.CDLL(DIAG_LIB_PATH)
        lib.muscle_debug_tile.restype = c_int
        lib.muscle_debug_tile.argtypes = [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]
        _diag = lib
    return _diag


def check(rc: int, what: str):
    if rc != MUSCLE_OK:
        msg = load().muscle_last_error()
        raise RuntimeError("muscle_b200 %s failed (code %d): %s" % (what, rc, msg.decode() if msg else "?"))
