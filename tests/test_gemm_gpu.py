"""GPU: the TF32 tcgen05 GEMM (vqacx_gemm_tf32, the one dense product of stage 2) against a float64 matmul.

TF32 keeps 10 mantissa bits of each operand (products are exact, accumulation is fp32), so the element-wise
tolerance is  2^-10 * sum_k |a_k b_k|  (both operands truncated) plus fp32 accumulation noise."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def gemm(a, b, m, n, k, a_mn, b_mn, bias=None, act=0, split_k=1, bn=0, c=None, accumulate=0):
    from vqa_counterexamples_b200._lib import lib, check, ptr, current_stream

    if c is None:
        c = torch.zeros((m, n), dtype=torch.float32, device=a.device)
    check(lib.vqacx_gemm_tf32(m, n, k, ptr(a), a.stride(0), a_mn, ptr(b), b.stride(0), b_mn, ptr(c), c.stride(0),
                              ptr(bias), act, split_k, bn, accumulate, current_stream()), "vqacx_gemm_tf32")
    return c


def reference(a, b, a_mn, b_mn, bias, act):
    A = a.double().t() if a_mn else a.double()
    B = b.double().t() if b_mn else b.double()
    ref = A @ B.t()
    mag = A.abs() @ B.abs().t()
    if bias is not None:
        ref = ref + bias.double()
    if act == 1:
        ref = torch.tanh(ref)
    elif act == 2:
        ref = torch.relu(ref)
    return ref, mag


CASES = [
    # m, n, k, a_mn, b_mn, bias, act, split_k, bn
    (128, 128, 64, 0, 0, False, 0, 1, 128),
    (128, 64, 32, 1, 0, False, 0, 1, 64),
    (128, 64, 32, 0, 1, False, 0, 1, 64),
    (128, 128, 96, 1, 1, False, 0, 1, 128),
    (1536, 360, 2048, 0, 0, True, 1, 1, 0),      # MUTAN linear_v shape
    (1600, 300, 4457, 0, 0, True, 2, 1, 0),      # layer-1 candidate part, ragged K and N
    (64, 320, 7232, 0, 0, False, 0, 8, 0),       # shared part, split-K
    (300, 4512, 1536, 1, 1, False, 0, 1, 0),     # dW1 = dpre^T . X
    (2000, 320, 1536, 1, 1, False, 0, 1, 0),     # dM = p^T . dpre
    (1536, 320, 320, 0, 1, False, 0, 1, 0),      # dh1 = dpre2 . W2
    (2000, 2400, 300, 0, 1, False, 0, 1, 256),   # dE = dM . W1a
    (37, 50, 100, 0, 0, True, 0, 1, 32),
    (1, 33, 36, 1, 1, False, 0, 3, 64),
]


@pytest.mark.parametrize("m,n,k,a_mn,b_mn,use_bias,act,split_k,bn", CASES)
def test_gemm_matches_fp64(m, n, k, a_mn, b_mn, use_bias, act, split_k, bn):
    g = torch.Generator(device="cuda").manual_seed(m * 7 + n * 3 + k)
    pad = lambda x: (x + 3) // 4 * 4
    a = torch.randn((k, pad(m)) if a_mn else (m, pad(k)), generator=g, device="cuda")
    b = torch.randn((k, pad(n)) if b_mn else (n, pad(k)), generator=g, device="cuda")
    a = a[:, :m] if a_mn else a[:, :k]   # views with padded leading dimension
    b = b[:, :n] if b_mn else b[:, :k]
    bias = torch.randn(n, generator=g, device="cuda") if use_bias else None
    out = gemm(a, b, m, n, k, a_mn, b_mn, bias=bias, act=act, split_k=split_k, bn=bn)
    ref, mag = reference(a, b, a_mn, b_mn, bias, act)
    tol = 2.0 ** -10 * mag + 1e-5 * mag + 1e-6
    err = (out.double() - ref).abs()
    assert bool((err <= tol).all()), f"max err {err.max().item():.3e} (tol at that point {tol.flatten()[err.argmax()].item():.3e})"
    # and it is not accidentally exact-zero / untouched
    assert out.abs().sum().item() > 0


def test_gemm_accumulate_adds():
    g = torch.Generator(device="cuda").manual_seed(1)
    a = torch.randn((256, 64), generator=g, device="cuda")
    b = torch.randn((96, 64), generator=g, device="cuda")
    c0 = torch.randn((256, 96), generator=g, device="cuda")
    c = c0.clone()
    gemm(a, b, 256, 96, 64, 0, 0, c=c, accumulate=1)
    ref = c0.double() + a.double() @ b.double().t()
    assert torch.allclose(c.double(), ref, atol=5e-2, rtol=1e-2)
    assert not torch.allclose(c, c0)


def test_gemm_rejects_bad_alignment():
    from vqa_counterexamples_b200._lib import VqacxError

    a = torch.randn((128, 66), device="cuda")
    b = torch.randn((64, 66), device="cuda")
    with pytest.raises(VqacxError):
        gemm(a, b, 128, 64, 66, 0, 0)


SPLITK_CASES = [
    # m, n, k, a_mn, b_mn, bias, act, split_k (0 = cost model), bn (0 = cost model)
    (1536, 300, 6480, 0, 0, True, 2, 0, 0),      # layer 1 at B = 64: the model must pick a wide tile and split K
    (1536, 300, 6480, 0, 0, True, 2, 6, 160),
    (64, 360, 2400, 0, 0, True, 1, 0, 0),        # MUTAN linear_q at B = 64 (one row tile)
    (2000, 300, 2400, 0, 0, False, 0, 0, 0),     # M = E . W1a^T
    (300, 4480, 1536, 1, 1, False, 0, 0, 0),     # dW1 = dpre^T . X
    (2000, 300, 1536, 1, 1, False, 0, 4, 0),     # dM = p^T . dpre
    (300, 300, 1536, 1, 1, False, 0, 0, 0),      # dW2
    (37, 50, 100, 0, 0, True, 1, 2, 32),         # ragged everything
    (1600, 360, 2048, 0, 0, True, 1, 0, 0),
]


@pytest.mark.parametrize("m,n,k,a_mn,b_mn,use_bias,act,split_k,bn", SPLITK_CASES)
def test_gemm_deterministic_splitk_matches_fp64(m, n, k, a_mn, b_mn, use_bias, act, split_k, bn):
    from vqa_counterexamples_b200._lib import lib, check, ptr, current_stream

    g = torch.Generator(device="cuda").manual_seed(m * 5 + n * 11 + k)
    pad = lambda x: (x + 3) // 4 * 4
    a = torch.randn((k, pad(m)) if a_mn else (m, pad(k)), generator=g, device="cuda")
    b = torch.randn((k, pad(n)) if b_mn else (n, pad(k)), generator=g, device="cuda")
    a = a[:, :m] if a_mn else a[:, :k]
    b = b[:, :n] if b_mn else b[:, :k]
    bias = torch.randn(n, generator=g, device="cuda") if use_bias else None
    scratch = torch.full((32 * m * ((n + 31) // 32 * 32),), float("nan"), device="cuda")
    outs = []
    for _ in range(2):
        c = torch.full((m, n), float("nan"), device="cuda")
        check(lib.vqacx_gemm_tf32_splitk(m, n, k, ptr(a), a.stride(0), a_mn, ptr(b), b.stride(0), b_mn, ptr(c),
                                         c.stride(0), ptr(bias), act, split_k, bn, ptr(scratch),
                                         scratch.numel() * 4, current_stream()), "vqacx_gemm_tf32_splitk")
        outs.append(c)
    ref, mag = reference(a, b, a_mn, b_mn, bias, act)
    tol = 2.0 ** -10 * mag + 1e-5 * mag + 1e-6
    err = (outs[0].double() - ref).abs()
    assert bool((err <= tol).all()), f"max err {err.max().item():.3e}"
    assert torch.equal(outs[0], outs[1])          # fixed summation order: bit-identical run to run


@pytest.mark.parametrize("m,n,k,ld_extra,act", [(128, 64, 64, 0, 0), (300, 300, 360, 24, 2), (1000, 2000, 384, 0, 0),
                                                 (70, 37, 200, 8, 1), (4096, 300, 4096, 0, 2), (64, 320, 2048, 0, 0),
                                                 # >= 296 m-tiles: CTA-pair mode (cta_group::2, shared weight tile); odd tile
                                                 # counts leave the last pair's second CTA without rows
                                                 (38000, 300, 520, 8, 2), (37900, 2000, 384, 0, 0), (38100, 96, 200, 0, 1)])
def test_gemm_f16_operands_match_fp32_reference(m, n, k, ld_extra, act):
    """kind::f16 product of the fast evaluation path (vqacx_gemm_f16): fp16 operands, fp32 accumulation, ragged K (TMA
    zero fill past the k extent, garbage in the row padding must not leak in), bias + activation, fp32 and fp16 outputs."""
    from vqa_counterexamples_b200._lib import lib, check, ptr, current_stream

    g = torch.Generator(device="cuda").manual_seed(m * 7 + n)
    lda = (k + ld_extra + 7) // 8 * 8
    a = torch.full((m, lda), 1e4, dtype=torch.float16, device="cuda")      # padding holds large values on purpose
    b = torch.full((n, lda), -1e4, dtype=torch.float16, device="cuda")
    a[:, :k] = (torch.randn((m, k), generator=g, device="cuda") * 0.5).half()
    b[:, :k] = (torch.randn((n, k), generator=g, device="cuda") * 0.1).half()
    bias = torch.randn(n, generator=g, device="cuda")
    c = torch.empty((m, n), dtype=torch.float32, device="cuda")
    ldh = (n + 7) // 8 * 8
    ch = torch.zeros((m, ldh), dtype=torch.float16, device="cuda")
    check(lib.vqacx_gemm_f16(m, n, k, ptr(a), lda, ptr(b), lda, ptr(c), n, ptr(ch), ldh, ptr(bias), act, current_stream()),
          "vqacx_gemm_f16")
    ref = a[:, :k].double() @ b[:, :k].double().t() + bias.double()
    ref = torch.tanh(ref) if act == 1 else (torch.relu(ref) if act == 2 else ref)
    scale = max(1.0, float(ref.abs().max()))
    assert float((c.double() - ref).abs().max()) <= 2e-4 * scale          # fp32 accumulation of exact fp16 products
    assert float((ch[:, :n].double() - ref).abs().max()) <= 2e-3 * scale  # + one fp16 rounding of the output
    # fp16-only output (no fp32 C)
    ch2 = torch.zeros_like(ch)
    check(lib.vqacx_gemm_f16(m, n, k, ptr(a), lda, ptr(b), lda, None, 0, ptr(ch2), ldh, ptr(bias), act, current_stream()),
          "vqacx_gemm_f16")
    assert torch.equal(ch2[:, :n], ch[:, :n])
