"""ctypes binding of libgnncca.so (include/gnncca.h).  There is no CPU or eager fallback: if the library
cannot be loaded, importing the kernels fails loudly."""
from __future__ import annotations

import ctypes as C
import os
import threading

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GNNCCA_LIB") or os.path.join(PKG, "libgnncca.so")     # GNNCCA_LIB: an experimental build (tools/variant.sh)

OK, ERR_BAD_ARG, ERR_UNSUPPORTED, ERR_LAUNCH, ERR_WORKSPACE = 0, -1, -2, -3, -4
TOTALS_MAGIC = 0x5eed5eed5eed          # GNNCCA_TOTALS_MAGIC
AGG = {"sum": 0, "mean": 1, "max": 2}
MATH_FP32, MATH_TF32 = 0, 1
ATTR_MODES = {"full": 0, "only_appearance": 1, "only_dist": 2}
CSR_NFLAGS = 4

c_f32p = C.c_void_p   # all device pointers travel as plain integers
i64, i32, vp = C.c_int64, C.c_int32, C.c_void_p


class MpCfg(C.Structure):
    _fields_ = [("f_node", i32), ("f_edge", i32), ("cls_hidden", i32), ("agg", i32), ("reattach_nodes", i32),
                ("reattach_edges", i32), ("math", i32), ("edge_phase", i32), ("node_begin", i32), ("node_end", i32)]


class MpWeights(C.Structure):
    _fields_ = [("w_edge", vp), ("b_edge", vp), ("w_node", vp), ("b_node", vp), ("w_c1", vp), ("b_c1", vp),
                ("bn_weight", vp), ("bn_bias", vp), ("bn_mean", vp), ("bn_var", vp), ("w_c2", vp), ("b_c2", vp),
                ("bn_eps", C.c_float), ("pad_", i32)]


class FinSeg(C.Structure):
    _fields_ = [("partial", vp), ("dst", vp), ("n_blocks", i32), ("width", i32), ("src_off", i32), ("src_stride", i32),
                ("rows", i32), ("cols", i32), ("dst_stride", i32), ("accumulate", i32)]


class FrameBwdStep(C.Structure):      # gnncca_frame_bwd_step
    _fields_ = [(n, vp) for n in ("R", "e", "e_prev", "h_out", "h_prev", "g_logits", "aff", "pa_agg", "pa_edge", "pa_proj", "pa_cls", "gh_prev")]


FRAMES_BWD_MAX_STEPS = 8
FRAMES_BWD_SMEM_LIMIT = 190 * 1024
MAX_FIN_SEGS = 24
CLS_ACC = 41
BLOCKS_PER_EDGE, BLOCKS_PER_NODE_WARP, BLOCKS_NODE_PROJECT, BLOCKS_ROW_CHUNKS = 0, 1, 2, 3

# name -> (restype, argtypes); must list every symbol include/gnncca.h declares (tests check this)
SIGNATURES = {
    "gnncca_abi_version": (C.c_int, []),
    "gnncca_status_string": (C.c_char_p, [C.c_int]),
    "gnncca_last_cuda_error": (C.c_char_p, []),
    "gnncca_launch_count": (i64, []),
    "gnncca_csr_from_edge_index": (C.c_int, [vp, i64, i64, vp, vp, vp, vp, vp]),
    "gnncca_csr_from_edge_rows": (C.c_int, [vp, vp, i64, i64, i64, vp, vp, vp, vp, vp]),
    "gnncca_segment_agg": (C.c_int, [vp, vp, i64, i32, i32, vp, vp]),
    "gnncca_csr_reverse_edges": (C.c_int, [vp, vp, vp, i64, i64, vp, vp, vp]),
    "gnncca_csr_sort_workspace_bytes": (i64, [i64, i64]),
    "gnncca_csr_sort": (C.c_int, [i64, i64, vp, vp, vp, vp, vp, i64, vp]),
    "gnncca_permute_rows": (C.c_int, [vp, vp, vp, i64, i32, i32, vp]),
    "gnncca_linear": (C.c_int, [vp, vp, vp, vp, i64, i32, i32, i32, vp, vp, vp, vp, C.c_float, i32, vp]),
    "gnncca_node_encoder": (C.c_int, [vp, vp, vp, vp, vp, vp, i64, i32, i32, i32, i32, vp]),
    "gnncca_node_encoder_train": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, i32, vp]),
    "gnncca_node_encoder_splits": (i32, [i64, i32]),
    "gnncca_node_encoder_scratch_bytes": (i64, [i64, i32, i32]),
    "gnncca_node_encoder_ws": (C.c_int, [vp, vp, vp, vp, vp, vp, vp, i64, i32, i32, i32, vp, i64, vp]),
    "gnncca_node_project": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, vp, i64, vp, vp, vp, vp]),
    "gnncca_mp_step_tile": (i32, []),
    "gnncca_mp_step_workspace_bytes": (i64, [C.POINTER(MpCfg), i64, i64]),
    "gnncca_mp_step_fwd": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, vp, vp, i64, i64, vp, vp, vp, vp, vp,
                                     vp, vp, vp, vp, vp, i64, vp]),
    "gnncca_mp_step_fwd_attr": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, vp, vp, i64, i64, vp, vp, vp, vp, i32, vp, vp, vp, vp,
                                          vp, vp, vp, vp, i64, vp]),
    "gnncca_mp_step_fwd_next": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, vp, vp, i64, i64, vp, vp, vp, vp, i32, vp, vp, vp, vp, vp, vp, vp,
                                          vp, vp, vp, vp, vp, vp, i64, vp]),
    "gnncca_mp_fused_fwd": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, vp, i32, vp, i64, vp, vp, vp, vp, vp, i64, i64,
                                      vp, vp, i32, i32, i32, i32, vp, vp, vp, vp, vp]),
    "gnncca_mpn_forward_frames": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, vp, vp, vp, i32, i32, vp, vp, i32, vp, vp, vp, vp, vp,
                                            i64, i64, i64, i32, i32, i32, i32, vp, vp, vp, vp, i64, vp]),
    "gnncca_train_blocks": (i32, [i64, i32]),
    "gnncca_node_agg_bwd": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, i64, vp, vp, vp, vp, vp, vp, vp, vp]),
    "gnncca_edge_bwd": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "gnncca_cls_moments": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, i64, vp, vp, vp]),
    "gnncca_cls_bn_prepare": (C.c_int, [i32, vp, i64, vp, vp, C.c_float, C.c_float, vp, vp, vp, vp]),
    "gnncca_cls_fwd": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, i64, vp, vp, vp]),
    "gnncca_cls_bwd": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, i64, vp, i32, vp, vp, vp, vp, vp]),
    "gnncca_pq_reduce": (C.c_int, [vp, vp, vp, vp, i64, vp, vp, vp]),
    "gnncca_node_project_bwd": (C.c_int, [C.POINTER(MpCfg), C.POINTER(MpWeights), vp, vp, vp, vp, vp, i64, vp, vp, vp, vp]),
    "gnncca_mp_frames_bwd_smem_bytes": (i64, [i32, i32]),
    "gnncca_mp_frames_bwd": (C.c_int, [vp, vp, vp, vp, i64, vp, vp, vp, i64, i64, i32, i32, vp, i32, vp, vp, vp, vp, vp, vp, vp]),
    "gnncca_edge_enc_bwd": (C.c_int, [vp, i32, vp, vp, i64, vp, vp]),
    "gnncca_linear_bwd": (C.c_int, [vp, vp, vp, vp, i32, i64, i32, i32, vp, vp, vp, vp, i32, i32, vp]),
    "gnncca_finalize_grads": (C.c_int, [C.POINTER(FinSeg), i32, vp]),
    "gnncca_loss_blocks": (i32, [i64]),
    "gnncca_bce_stats_fwd": (C.c_int, [C.POINTER(vp), i32, vp, i64, C.c_float, i32, C.c_float, C.c_float, vp, vp, vp, vp, vp]),
    "gnncca_bce_stats_bwd": (C.c_int, [C.POINTER(vp), C.POINTER(vp), i32, vp, i64, C.c_float, vp, vp, vp]),
    "gnncca_frame_clusters": (C.c_int, [vp, i32, vp, i64, vp, vp, i64, i64, i32, vp, vp, vp, vp, vp, vp]),
    "gnncca_foot_to_world": (C.c_int, [vp, vp, vp, vp, vp, i32, i64, vp, vp, vp, vp]),
    "gnncca_frame_valid": (C.c_int, [vp, vp, vp, vp, i64, vp, vp]),
    "gnncca_edge_mlp": (C.c_int, [vp, vp, i64, i32, C.POINTER(i32), C.POINTER(vp), C.POINTER(vp), C.POINTER(i32), vp]),
    "gnncca_colnorm_scratch_bytes": (i64, [i64, i32]),
    "gnncca_colnorm": (C.c_int, [vp, vp, i64, i32, vp, vp]),
    "gnncca_colnorm_static": (C.c_int, [vp, vp, i64, i32, vp]),
    "gnncca_colnorm_norms": (C.c_int, [vp, i64, i32, vp, vp, vp]),
    "gnncca_colnorm_apply": (C.c_int, [vp, vp, i64, i32, vp, vp]),
    "gnncca_scale_columns": (C.c_int, [vp, vp, i64, i32, vp, vp]),
    "gnncca_graph_prepare_scratch_bytes": (i64, [i64, i64]),
    "gnncca_graph_edge_counts": (C.c_int, [vp, vp, i64, i64, vp, vp, i64, vp]),
    "gnncca_graph_tile_total_ptr": (vp, [vp, i64, i64]),
    "gnncca_build_graphs": (C.c_int, [vp, vp, i64, i64, i64, i64, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp, vp, vp, vp, vp]),
    "gnncca_graph_prepare_all": (C.c_int, [vp, vp, i64, i64, vp, vp, i64, vp, i32, vp, vp, vp, i32, vp, vp, vp, vp]),
    "gnncca_graph_join": (C.c_int, [vp]),
}

_lock = threading.Lock()
_lib = None


class GnnccaError(RuntimeError):
    pass


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load libgnncca.so (building it with nvcc when absent).  Raises if that is impossible."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            if not build_if_missing:
                raise GnnccaError("libgnncca.so is missing (%s); run `python __graft_entry__.py build`" % LIB_PATH)
            from .build import build
            build()
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)   # AttributeError if the library does not export a declared symbol
            fn.restype, fn.argtypes = res, args
        if lib.gnncca_abi_version() != 1:
            raise GnnccaError("libgnncca.so ABI version mismatch")
        _lib = lib
        return lib


def check(status: int, what: str) -> None:
    if status == OK:
        return
    lib = load()
    msg = "%s: %s" % (what, lib.gnncca_status_string(status).decode())
    if status == ERR_LAUNCH:
        msg += " [%s]" % lib.gnncca_last_cuda_error().decode()
    if status == ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    if status == ERR_BAD_ARG:
        raise ValueError(msg)
    raise GnnccaError(msg)


def ptr(t) -> int:
    """Device pointer of a tensor (None -> NULL)."""
    return 0 if t is None else t.data_ptr()
