"""Warp-role cycle counters of one vqacx_gemm_tf32 launch: python tools/time_gemm.py M N K [bn] [split] [a_mn] [b_mn]"""
import ctypes, sys
import torch
sys.path.insert(0, ".")
from vqa_counterexamples_b200._lib import lib, check, ptr, current_stream

def run(M, N, K, bn=0, split=1, a_mn=0, b_mn=0, bias=True, reps=5):
    g = torch.Generator(device="cuda").manual_seed(0)
    ka = (K + 31) // 32 * 32
    a = torch.randn((K, (M + 3) // 4 * 4) if a_mn else (M, ka), generator=g, device="cuda")
    b = torch.randn((K, (N + 3) // 4 * 4) if b_mn else (N, ka), generator=g, device="cuda")
    c = torch.zeros((M, (N + 31) // 32 * 32), device="cuda")
    bv = torch.randn(N, generator=g, device="cuda") if bias else None
    prof = torch.zeros(8 * 4096, dtype=torch.int64, device="cuda")
    lib.vqacx_debug_set_knn_profile.argtypes = [ctypes.c_void_p]
    lib.vqacx_debug_set_knn_profile.restype = None
    def call():
        check(lib.vqacx_gemm_tf32(M, N, K, ptr(a), a.stride(0), a_mn, ptr(b), b.stride(0), b_mn, ptr(c), c.stride(0),
                                  ptr(bv), 0, split, bn, 1 if split > 1 else 0, current_stream()), "gemm")
    for _ in range(3): call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): call()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    lib.vqacx_debug_set_knn_profile(ctypes.c_void_p(prof.data_ptr()))
    call(); torch.cuda.synchronize()
    lib.vqacx_debug_set_knn_profile(None)
    p = prof.view(-1, 8).double().cpu()
    p = p[p[:, 6] > 0]
    names = ["prod wait-empty", "prod loop", "mma wait-full", "mma loop", "epi wait-tfull", "epi work", "cta total", "prologue"]
    print(f"M={M} N={N} K={K} bn={bn} split={split} a_mn={a_mn} b_mn={b_mn}: {us:.1f} us  {2.0*M*N*K/us/1e6:.1f} TFLOP/s  ctas={len(p)}")
    print("   " + "  ".join(f"{nm} {p[:, j].mean():.0f}" for j, nm in enumerate(names)))

if __name__ == "__main__":
    if len(sys.argv) > 3:
        run(*[int(x) for x in sys.argv[1:]])
    else:
        run(1536, 3600, 360)          # hvc
        run(64, 360, 2048)            # xvs
        run(1536, 360, 2048)          # xvc
        run(1536, 2000, 360)          # classif
        run(2000, 300, 2400)          # M = E.W1a^T
        run(1536, 300, 4480)          # layer 1 (one segment)
        run(1536, 300, 300)           # layer 2
        run(300, 4480, 1536, a_mn=1, b_mn=1)   # dW1c
        run(2000, 2400, 300, b_mn=1)  # dE
        run(4096, 4096, 4096)         # big square
        run(4096 * 24, 3600, 360)     # hvc at eval batch 4096
