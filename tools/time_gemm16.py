"""Warp-role cycle counters and timing experiments of one vqacx_gemm_f16 launch (needs a -DVQACX_DEBUG build:
VQACX_DEBUG_BUILD=1 python -m vqa_counterexamples_b200.build): python tools/time_gemm16.py M N K [half_out_only]"""
import ctypes, sys
import torch
sys.path.insert(0, ".")
from vqa_counterexamples_b200 import _lib
from vqa_counterexamples_b200._lib import lib, check, ptr, current_stream


def run(M, N, K, half_only=1, reps=5):
    g = torch.Generator(device="cuda").manual_seed(0)
    ka = (K + 63) // 64 * 64
    a = (torch.randn((M, ka), generator=g, device="cuda") * 0.3).half()
    b = (torch.randn((N, ka), generator=g, device="cuda") * 0.1).half()
    ldh = (N + 63) // 64 * 64
    c = None if half_only else torch.zeros((M, ldh), device="cuda")
    ch = torch.zeros((M, ldh), dtype=torch.float16, device="cuda")
    bv = torch.randn(N, generator=g, device="cuda")
    prof = torch.zeros(8 * 4096, dtype=torch.int64, device="cuda")
    lib.vqacx_debug_set_knn_profile.argtypes = [ctypes.c_void_p]
    lib.vqacx_debug_set_knn_profile.restype = None

    def call():
        check(lib.vqacx_gemm_f16(M, N, K, ptr(a), ka, ptr(b), ka, ptr(c), ldh, ptr(ch), ldh, ptr(bv), 0, current_stream()), "gemm16")

    def timeit():
        for _ in range(3):
            call()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            call()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps * 1e3

    us = timeit()
    print(f"M={M} N={N} K={K} half_only={half_only}: {us:.1f} us  {2.0 * M * N * K / us / 1e6:.1f} TFLOP/s")
    if _lib.is_debug_build():
        for bits, what in ((1, "no fp32 stores"), (2, "no tmem ld"), (4, "no addend loads"), (8, "no fp16 stores"),
                           (16, "no phase B"), (18, "no phase B, no tmem ld")):
            with _lib.option(_lib.OPT_GEMM_DEBUG, bits):
                print(f"   debug {bits} ({what}): {timeit():.1f} us")
        lib.vqacx_debug_set_knn_profile(ctypes.c_void_p(prof.data_ptr()))
        call(); torch.cuda.synchronize()
        lib.vqacx_debug_set_knn_profile(None)
        p = prof.view(-1, 8).double().cpu()
        p = p[p[:, 6] > 0]
        names = ["prod wait-empty", "prod loop", "mma wait-full", "mma loop", "epi wait-tfull", "epi work", "cta total", "prologue"]
        print("   " + "  ".join(f"{nm} {p[:, j].mean():.0f}" for j, nm in enumerate(names)))


if __name__ == "__main__":
    if len(sys.argv) > 3:
        run(*[int(x) for x in sys.argv[1:]])
    else:
        run(98304, 2000, 384)        # classifier, fp16 logits only
        run(98304, 300, 320, 0)      # layer 2, fp32 out
        run(98304, 300, 6528)        # layer 1 (single segment of the same K)
