"""ctypes binding of libvqacx.so (the C ABI declared in include/vqacx.h).

There is deliberately no fallback: if the CUDA library is missing the import fails, and every
entry point raises on a non-zero status. PyTorch is used only for device memory and streams.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvqacx.so")

VQACX_KNN_MAX_K = 64
VQACX_KNN_FORCE_EXACT = 1
VQACX_KNN_NO_SYMMETRIC = 2
VQACX_KNN_NSTATS = 8
VQACX_KNN_NPHASES = 4
VQACX_KNN_PACK_INTS = 130
VQACX_EINVAL, VQACX_EALIGN, VQACX_EWORKSPACE, VQACX_ETMA, VQACX_ERANGE, VQACX_ESTATIC = -1, -2, -3, -4, -5, -6


class VqacxError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU or PyTorch fallback for this path."
        )
    return ctypes.CDLL(LIB_PATH)


lib = _load()

_p = ctypes.c_void_p
_i64 = ctypes.c_int64
_i32 = ctypes.c_int
_u32 = ctypes.c_uint32
_sz = ctypes.c_size_t
_f32 = ctypes.c_float

lib.vqacx_version.restype = ctypes.c_char_p
lib.vqacx_error_string.restype = ctypes.c_char_p
lib.vqacx_error_string.argtypes = [_i32]
lib.vqacx_launch_count.restype = _i64
lib.vqacx_launch_count.argtypes = []

_SIGS = {
    "vqacx_knn_prep": (_i32, [_p, _i64, _i32, _p, _p, _p, _p, _p]),
    "vqacx_knn_workspace_bytes": (_sz, [_i64, _i64, _i32]),
    "vqacx_knn_search": (
        _i32,
        [_p, _p, _p, _p, _i64, _i64, _p, _p, _p, _i64, _i32, _i32, _p, _p, _p, _sz, _u32, _p, _p, _p],
    ),
    "vqacx_knn_search_phase": (
        _i32,
        [_p, _p, _p, _p, _i64, _i64, _p, _p, _p, _i64, _i32, _i32, _p, _p, _p, _sz, _u32, _p, _p, _p, _i32, _p],
    ),
    "vqacx_knn_kth_bound": (_i32, [_p, _i32, _i64, _i32, _p, _p]),
    "vqacx_knn_selfjoin_workspace_bytes": (_sz, [_i64, _i32, _i32]),
    "vqacx_knn_selfjoin_host": (_i32, [_p, _i64, _i32, _i32, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p, _p]),
    "vqacx_topk_merge": (_i32, [_p, _p, _i32, _i64, _i32, _p, _p, _p]),
    "vqacx_knn_selfjoin_shard_pack": (_i32, [_i64, _i32, _i32, _i64, _p, _p, _sz, _p]),
    "vqacx_knn_selfjoin_shard_owner": (_i32, [_p, _p, _p, _i64, _i32, _i32, _i32, _p, _i64, _i64, _i64, _p, _p, _p, _sz,
                                              _p, _p, _p]),
    "vqacx_topk_merge_checked": (_i32, [_p, _p, _i32, _i64, _i32, _p, _p, _p, _p, _p]),
    "vqacx_knn_selfjoin_shard_supported": (_i32, [_i64, _i32, _i32]),
    "vqacx_knn_selfjoin_shard_workspace_bytes": (_sz, [_i64, _i32, _i32]),
    "vqacx_knn_selfjoin_shard": (_i32, [_p, _p, _p, _p, _i64, _i32, _i32, _i32, _i32, _i32, _p, _p, _p, _p, _p, _sz, _p,
                                        _p, _p]),
    "vqacx_gemm_tf32_splitk": (_i32, [_i32, _i32, _i32, _p, _i64, _i32, _p, _i64, _i32, _p, _i64, _p, _i32, _i32, _i32,
                                      _p, _sz, _p]),
    "vqacx_gemm_tf32_splitk_static": (_i32, [_i32, _i32, _i32, _p, _i64, _i32, _p, _i64, _i32, _p, _i64, _p, _i32,
                                             _i32, _i32, _p]),
    "vqacx_gemm_tf32": (_i32, [_i32, _i32, _i32, _p, _i64, _i32, _p, _i64, _i32, _p, _i64, _p, _i32, _i32, _i32,
                               _i32, _p]),
}


class NcxDims(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in
                ("dim_v", "dim_q", "dim_hv", "dim_hq", "dim_mm", "R", "n_ans", "dim_a", "dim_h", "n_layers", "knn")]


class NcxFrozen(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in ("wv", "bv", "wq", "bq", "whv", "bhv", "whq", "bhq", "wc", "bc",
                                               "whv_f16")]


class NcxLayout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int64) for n in
                ("total", "off_emb", "off_w1", "off_b1", "off_w2", "off_b2", "off_w3", "off_b3", "off_wout", "off_bout",
                 "ld_w1", "ld_h", "col_vother", "col_vmult", "col_zother", "col_dist", "col_vorig", "col_q",
                 "col_zorig", "col_agt", "col_rank", "col_aother", "kc", "ks")]


class NcxBatch(ctypes.Structure):
    _fields_ = [("batch", ctypes.c_int64), ("features", ctypes.c_void_p), ("ld_features", ctypes.c_int64),
                ("img_idx", ctypes.c_void_p), ("q_emb", ctypes.c_void_p), ("ld_q", ctypes.c_int64),
                ("answer_aid", ctypes.c_void_p), ("drop_p", ctypes.c_float), ("drop_seed", ctypes.c_uint64),
                ("drop_mask", ctypes.c_void_p), ("drop_seed_dev", ctypes.c_void_p), ("lesion", ctypes.c_uint32),
                ("v_rank", ctypes.c_void_p), ("hv_table", ctypes.c_void_p)]


class NcxHeadLayout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int64) for n in
                ("total", "off_wv", "off_bv", "off_wq", "off_bq", "off_whv", "off_bhv", "off_whq", "off_bhq", "off_wc",
                 "off_bc")]


class NcxHeadTrain(ctypes.Structure):
    _fields_ = [("p_v", ctypes.c_float), ("p_q", ctypes.c_float), ("p_c", ctypes.c_float), ("seed", ctypes.c_uint64),
                ("mask_v", ctypes.c_void_p), ("mask_q", ctypes.c_void_p), ("mask_c", ctypes.c_void_p)]


VQACX_LESION_V_MULT, VQACX_LESION_V_DIST, VQACX_LESION_V_RANK = 1, 2, 4


_SIGS.update({
    "vqacx_ncx_layout": (_i32, [ctypes.POINTER(NcxDims), ctypes.POINTER(NcxLayout)]),
    "vqacx_ncx_workspace_bytes": (_sz, [ctypes.POINTER(NcxDims), _i64, _i32]),
    "vqacx_ncx_forward": (_i32, [ctypes.POINTER(NcxDims), ctypes.POINTER(NcxFrozen), _p, ctypes.POINTER(NcxBatch), _p,
                                 _sz, _i32, _p, _p]),
    "vqacx_ce_recall": (_i32, [_p, _p, _i64, _i32, _p, _p, _f32, _p, _p]),
    "vqacx_ncx_backward": (_i32, [ctypes.POINTER(NcxDims), _p, ctypes.POINTER(NcxBatch), _p, _sz, _p, _p, _p]),
    "vqacx_ncx_hidden": (_i32, [ctypes.POINTER(NcxDims), _i64, _i32, _p, _sz, _i32, _p, _p]),
    "vqacx_adam_step": (_i32, [_p, _p, _p, _p, _i64, _f32, _f32, _f32, _f32, _i32, _f32, _p]),
    "vqacx_adam_step_dev": (_i32, [_p, _p, _p, _p, _i64, _f32, _f32, _f32, _f32, _p, _f32, _p]),
    "vqacx_step_tick": (_i32, [_p, _p, ctypes.c_uint64, _p]),
    "vqacx_launch_count_add": (None, [_i64]),
    "vqacx_gru_workspace_bytes": (_sz, [_i64, _i32, _i32, _i32]),
    "vqacx_gru_encode": (_i32, [_p, _i64, _i32, _p, _i64, _i32, _p, _p, _p, _i32, _i32, _p, _p, _sz, _p]),
    "vqacx_gru_train_workspace_bytes": (_sz, [_i64, _i32, _i32, _i32]),
    "vqacx_gru_forward_train": (_i32, [_p, _i64, _i32, _p, _i64, _i32, _p, _p, _p, _i32, _f32, ctypes.c_uint64, _p, _p, _p,
                                       _p, _sz, _p]),
    "vqacx_gru_backward": (_i32, [_i64, _i32, _p, _i64, _i32, _p, _p, _i32, _p, _i64, _p, _sz, _p, _p, _p, _p, _p]),
    "vqacx_pool_features": (_i32, [_p, _i64, _i32, _i32, _p, _p, _p, _p, _p, _p]),
    "vqacx_cx_build_examples": (_i32, [_p, _i64, _i32, _p, _p, _i64, _p, _p, _p]),
    "vqacx_cx_vqa_forward": (_i32, [ctypes.POINTER(NcxDims), ctypes.POINTER(NcxFrozen), ctypes.POINTER(NcxBatch), _p, _sz,
                                    _p, _p, _p, _p]),
    "vqacx_cx_head_scores": (_i32, [ctypes.POINTER(NcxDims), ctypes.POINTER(NcxFrozen), ctypes.POINTER(NcxBatch), _p, _sz,
                                    _i32, _p, _f32, _p, _p]),
    "vqacx_cx_linear_context_forward": (_i32, [ctypes.POINTER(NcxDims), ctypes.POINTER(NcxFrozen),
                                               ctypes.POINTER(NcxBatch), _p, _sz, _p, _p, _p, _p, _p]),
    "vqacx_cx_linear_context_backward": (_i32, [ctypes.POINTER(NcxDims), _i64, _p, _sz, _p, _p, _p, _p, _p]),
    "vqacx_cosine_similarity": (_i32, [_p, _i32, _i32, _p, _p]),
    "vqacx_ncx_fast_tables_bytes": (_sz, [ctypes.POINTER(NcxDims), _i64]),
    "vqacx_ncx_fast_twins_bytes": (_sz, [ctypes.POINTER(NcxDims)]),
    "vqacx_ncx_fast_workspace_bytes": (_sz, [ctypes.POINTER(NcxDims), _i64]),
    "vqacx_ncx_fast_build_tables": (_i32, [ctypes.POINTER(NcxDims), ctypes.POINTER(NcxFrozen), _p, _i64, _i64, _p, _sz, _p,
                                           _sz, _p]),
    "vqacx_ncx_fast_build_twins": (_i32, [ctypes.POINTER(NcxDims), ctypes.POINTER(NcxFrozen), _p, _p, _sz, _p, _sz, _p]),
    "vqacx_ncx_fast_forward": (_i32, [ctypes.POINTER(NcxDims), ctypes.POINTER(NcxFrozen), _p, ctypes.POINTER(NcxBatch), _p,
                                      _i64, _p, _p, _p, _sz, _p, _p]),
    "vqacx_ncx_fast_p1_bytes": (_sz, [ctypes.POINTER(NcxDims), _i64]),
    "vqacx_ncx_fast_build_p1": (_i32, [ctypes.POINTER(NcxDims), _p, _p, _i64, _i64, _p, _sz, _p, _sz, _p]),
    "vqacx_gemm_f16": (_i32, [_i32, _i32, _i32, _p, _i64, _p, _i64, _p, _i64, _p, _i64, _p, _i32, _p]),
    "vqacx_ncx_head_layout": (_i32, [ctypes.POINTER(NcxDims), ctypes.POINTER(NcxHeadLayout)]),
    "vqacx_ncx_tv_workspace_bytes": (_sz, [ctypes.POINTER(NcxDims), _i64]),
    "vqacx_ncx_forward_tv": (_i32, [ctypes.POINTER(NcxDims), _p, _p, ctypes.POINTER(NcxBatch),
                                    ctypes.POINTER(NcxHeadTrain), _p, _sz, _p, _sz, _p, _p]),
    "vqacx_ncx_backward_tv": (_i32, [ctypes.POINTER(NcxDims), _p, _p, ctypes.POINTER(NcxBatch),
                                     ctypes.POINTER(NcxHeadTrain), _p, _sz, _p, _sz, _p, _p, _p, _p, _i64, _p]),
})

(OPT_KNN_LOCKSTEP, OPT_KNN_SYM_GROUPS, OPT_KNN_SYM_MIRROR_CAP, OPT_KNN_SYM_LOG_CAP, OPT_KNN_NO_SYMMETRIC, OPT_KNN_NO_QSORT,
 OPT_NCX_SERIAL, OPT_NCX_NO_MUTAN_FUSE, OPT_KNN_SINGLE_EXCHANGE, OPT_KNN_SHARD_SYM_GROUPS, OPT_KNN_SYM_LOCKSTEP,
 OPT_NCX_NO_PDL, OPT_GEMM_NO_PAIR, OPT_KNN_MIRROR_WARPS) = range(1, 15)
OPT_KNN_SINGLE_CTA, OPT_KNN_OVERLAP, OPT_KNN_WAVE_TRACE, OPT_KNN_DEBUG, OPT_GEMM_DEBUG = range(16, 21)   # debug builds only
_SIGS.update({"vqacx_get_option": (_i32, [_i32]), "vqacx_set_option": (_i32, [_i32, _i32])})


class option:
    """`with option(OPT_KNN_LOCKSTEP, 0): ...` -- sets a library option for the block (tests, A/B tools)."""

    def __init__(self, opt, value):
        self.opt, self.value = opt, value

    def __enter__(self):
        self.old = lib.vqacx_set_option(self.opt, int(self.value))
        if self.old < 0:
            raise VqacxError(f"option {self.opt} does not exist in this build (debug-only options need -DVQACX_DEBUG)")
        return self

    def __exit__(self, *exc):
        lib.vqacx_set_option(self.opt, self.old)
        return False


def is_debug_build():
    return b"debug build" in lib.vqacx_version()


VQACX_HEAD_BLACKBOX = 0
VQACX_HEAD_SEMANTIC = 1


def _bind(sigs):
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


_bind(_SIGS)


def check(status, what):
    if status != 0:
        msg = lib.vqacx_error_string(int(status)).decode()
        raise VqacxError(f"{what} failed: {msg} (status {status})")


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def current_stream():
    import torch

    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
