"""CPU: the C-ABI library loads without a GPU and exports every function include/vqacx.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "vqacx.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vqacx_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported():
    from vqa_counterexamples_b200 import _lib

    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = declared_functions()
    assert len(names) >= 6
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/vqacx.h but not exported"


def test_version_and_error_strings():
    from vqa_counterexamples_b200._lib import lib

    assert b"sm_100a" in lib.vqacx_version()
    assert lib.vqacx_error_string(0) == b"ok"
    assert b"workspace" in lib.vqacx_error_string(-3)


def test_workspace_query_needs_no_gpu():
    from vqa_counterexamples_b200._lib import lib

    small = lib.vqacx_knn_workspace_bytes(10, 1000, 2048)
    big = lib.vqacx_knn_workspace_bytes(123287, 123287, 2048)
    # the self-join size includes the symmetric sweep's slots, mirror buffers and append logs (~2.7 GB of 180 GB)
    assert 0 < small < big < (4 << 30)
    assert lib.vqacx_knn_workspace_bytes(123287, 123286, 2048) < (1 << 30)   # not a self-join: plain sweep only


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    import importlib

    from vqa_counterexamples_b200 import _lib

    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    try:
        _lib._load()
    except ImportError as e:
        assert "no CPU or PyTorch fallback" in str(e)
    else:
        raise AssertionError("loading a missing library must raise")
    importlib.reload(_lib)


def test_options_are_explicit_api_not_per_call_environment(monkeypatch):
    """Tuning knobs are a table read from the environment ONCE and changed only through vqacx_set_option; the
    measurement options that can invalidate results do not exist in the default build."""
    from vqa_counterexamples_b200 import _lib

    lib = _lib.lib
    assert lib.vqacx_get_option(_lib.OPT_KNN_SYM_GROUPS) in (3, 5, 7, 9, 11, 13, 15)
    old = lib.vqacx_get_option(_lib.OPT_KNN_LOCKSTEP)
    monkeypatch.setenv("VQACX_KNN_LOCKSTEP", "3")              # too late: the environment was read at first use
    assert lib.vqacx_get_option(_lib.OPT_KNN_LOCKSTEP) == old
    with _lib.option(_lib.OPT_KNN_LOCKSTEP, 2):
        assert lib.vqacx_get_option(_lib.OPT_KNN_LOCKSTEP) == 2
    assert lib.vqacx_get_option(_lib.OPT_KNN_LOCKSTEP) == old
    assert lib.vqacx_set_option(999, 1) == _lib.VQACX_EINVAL
    if not _lib.is_debug_build():
        for o in (_lib.OPT_KNN_SINGLE_CTA, _lib.OPT_KNN_OVERLAP, _lib.OPT_KNN_DEBUG, _lib.OPT_GEMM_DEBUG):
            assert lib.vqacx_set_option(o, 1) == _lib.VQACX_EINVAL
            assert lib.vqacx_get_option(o) == 0


def test_trainable_vqa_layout_and_workspace_queries_need_no_gpu():
    """vqacx_ncx_head_layout / *_tv_workspace_bytes / vqacx_gru_train_workspace_bytes are pure host arithmetic: offsets are
    16-byte aligned, ordered as the header says, and sized by the reference's tensor shapes (fusion.py:62-76, noatt.py:19)."""
    import ctypes

    from vqa_counterexamples_b200._lib import NcxDims, NcxHeadLayout, lib

    d = NcxDims(2048, 2400, 360, 360, 360, 10, 2000, 2400, 300, 2, 24)
    hl = NcxHeadLayout()
    assert lib.vqacx_ncx_head_layout(ctypes.byref(d), ctypes.byref(hl)) == 0
    offs = [hl.off_wv, hl.off_bv, hl.off_wq, hl.off_bq, hl.off_whv, hl.off_bhv, hl.off_whq, hl.off_bhq, hl.off_wc, hl.off_bc]
    sizes = [360 * 2048, 360, 360 * 2400, 360, 3600 * 360, 3600, 3600 * 360, 3600, 2000 * 360, 2000]
    assert offs[0] == 0 and all(o % 4 == 0 for o in offs)
    for i in range(9):
        assert offs[i] + sizes[i] <= offs[i + 1]
    assert offs[9] + sizes[9] <= hl.total < offs[9] + sizes[9] + 64
    small, big = lib.vqacx_ncx_tv_workspace_bytes(ctypes.byref(d), 8), lib.vqacx_ncx_tv_workspace_bytes(ctypes.byref(d), 64)
    assert 0 < small < big < (1 << 30)
    bad = NcxDims(2048, 2400, 360, 360, 362, 10, 2000, 2400, 300, 2, 24)          # dim_mm not a multiple of 4
    assert lib.vqacx_ncx_head_layout(ctypes.byref(bad), ctypes.byref(hl)) < 0
    assert lib.vqacx_ncx_tv_workspace_bytes(ctypes.byref(bad), 8) == 0
    g1, g2 = lib.vqacx_gru_train_workspace_bytes(8, 26, 620, 2400), lib.vqacx_gru_train_workspace_bytes(64, 26, 620, 2400)
    assert 0 < g1 < g2 < (2 << 30)
    assert lib.vqacx_gru_train_workspace_bytes(8, 26, 621, 2400) == 0             # E must be a multiple of 4
    # the no-allocation GEMM entry point rejects before touching the device
    assert lib.vqacx_gemm_tf32_splitk_static(4096, 512, 256, None, 256, 0, None, 256, 0, None, 512, None, 0, 0, 0, None) < 0
