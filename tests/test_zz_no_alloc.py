"""vqacx_gemm_tf32_splitk_static: the split-K product with its scratch in a static device array (no allocation).

Collected last on purpose. The entry point has no counterpart in the reference; the
oracle here is the existing vqacx_gemm_tf32_splitk with a caller-provided scratch of the same size, which chooses the
same tile / split and sums the K-slices in the same order, so the two must agree bit for bit.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_BYTES = 8 << 20          # VQACX_STATIC_SCRATCH_BYTES in include/vqacx.h
ESTATIC = -6                    # VQACX_ESTATIC

# (m, n, k, a_mn, b_mn, bias, act, split_k, bn): two different shapes back to back, so the second call finds the first
# call's partial tiles in the shared array
CASES = [(64, 2400, 2048, 0, 0, True, 1, 8, 0),
         (300, 310, 1000, 0, 1, True, 2, 0, 0),
         (64, 2400, 2048, 0, 0, True, 1, 8, 0)]


def _operands(m, n, k, a_mn, b_mn, use_bias):
    g = torch.Generator(device="cuda").manual_seed(m * 5 + n * 11 + k)
    pad = lambda x: (x + 3) // 4 * 4
    a = torch.randn((k, pad(m)) if a_mn else (m, pad(k)), generator=g, device="cuda")
    b = torch.randn((k, pad(n)) if b_mn else (n, pad(k)), generator=g, device="cuda")
    a = a[:, :m] if a_mn else a[:, :k]
    b = b[:, :n] if b_mn else b[:, :k]
    bias = torch.randn(n, generator=g, device="cuda") if use_bias else None
    return a, b, bias


def test_static_scratch_splitk_matches_caller_scratch_entry_point():
    from vqa_counterexamples_b200._lib import lib, check, ptr, current_stream

    scratch = torch.full((STATIC_BYTES // 4,), float("nan"), device="cuda")
    work = []
    for m, n, k, a_mn, b_mn, use_bias, act, split_k, bn in CASES:      # every torch allocation happens up front
        a, b, bias = _operands(m, n, k, a_mn, b_mn, use_bias)
        work.append((a, b, bias, torch.full((m, n), float("nan"), device="cuda"),
                     torch.full((m, n), float("nan"), device="cuda")))
    torch.cuda.synchronize()
    for i, ((m, n, k, a_mn, b_mn, use_bias, act, split_k, bn), (a, b, bias, c_new, c_old)) in enumerate(zip(CASES, work)):
        free_before = torch.cuda.mem_get_info()[0]
        reserved_before = torch.cuda.memory_reserved()
        check(lib.vqacx_gemm_tf32_splitk_static(m, n, k, ptr(a), a.stride(0), a_mn, ptr(b), b.stride(0), b_mn,
                                                ptr(c_new), c_new.stride(0), ptr(bias), act, split_k, bn,
                                                current_stream()), "vqacx_gemm_tf32_splitk_static")
        torch.cuda.synchronize()
        if i > 0:        # the first call may load the module lazily; after that device memory must not move at all
            assert torch.cuda.mem_get_info()[0] == free_before
        assert torch.cuda.memory_reserved() == reserved_before
        check(lib.vqacx_gemm_tf32_splitk(m, n, k, ptr(a), a.stride(0), a_mn, ptr(b), b.stride(0), b_mn, ptr(c_old),
                                         c_old.stride(0), ptr(bias), act, split_k, bn, ptr(scratch),
                                         scratch.numel() * 4, current_stream()), "vqacx_gemm_tf32_splitk")
        torch.cuda.synchronize()
        assert bool(torch.isfinite(c_new).all())
        assert torch.equal(c_new, c_old), f"case {i}: max diff {(c_new - c_old).abs().max().item():.3e}"
        ref = (a.double().t() if a_mn else a.double()) @ (b.double() if b_mn else b.double().t())
        mag = (a.double().abs().t() if a_mn else a.double().abs()) @ (b.double().abs() if b_mn else b.double().abs().t())
        if bias is not None:
            ref, mag = ref + bias.double(), mag + bias.double().abs()
        ref = torch.tanh(ref) if act == 1 else (torch.relu(ref) if act == 2 else ref)
        tol = 2.0 ** -10 * mag + 1e-5 * mag + 1e-6          # the TF32 bound tests/test_gemm_gpu.py uses
        assert bool(((c_new.double() - ref).abs() <= tol).all())


def test_static_scratch_rejects_oversized_problem():
    from vqa_counterexamples_b200._lib import lib, ptr, current_stream

    m, n, k = 4096, 512, 256                       # one slice = 8 MiB: two cannot fit
    a = torch.zeros((m, k), device="cuda")
    b = torch.zeros((n, k), device="cuda")
    c = torch.full((m, n), 7.0, device="cuda")
    rc = lib.vqacx_gemm_tf32_splitk_static(m, n, k, ptr(a), k, 0, ptr(b), k, 0, ptr(c), n, None, 0, 0, 0,
                                           current_stream())
    assert rc == ESTATIC
    assert "static scratch" in lib.vqacx_error_string(rc).decode()
    rc = lib.vqacx_gemm_tf32_splitk_static(64, 2400, k, ptr(a), k, 0, ptr(b), k, 0, ptr(c), 2400, None, 0, 32, 0,
                                           current_stream())          # 32 slices of 64 x 2400 floats > 8 MiB
    assert rc == ESTATIC
    torch.cuda.synchronize()
    assert bool((c == 7.0).all())                  # nothing was launched
