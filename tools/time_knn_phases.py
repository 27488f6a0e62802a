"""Phase timing of the kNN self-join at bench size: python tools/time_knn_phases.py [N] [reps]"""
import sys
import torch
sys.path.insert(0, ".")
from vqa_counterexamples_b200 import synth
from vqa_counterexamples_b200.knn import NearestNeighbors

n = int(sys.argv[1]) if len(sys.argv) > 1 else 123287
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
x = synth.clustered_features_cuda(n, 2048, seed=1234)
nn_ = NearestNeighbors(n_neighbors=25).fit(x)
nn_.profile = True
for r in range(reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    d, i = nn_.kneighbors(x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    ph = nn_.last_phase_ms
    print(f"n={n} rep={r} total {ms:.2f} ms  {n / ms * 1e3:.0f} q/s | gemm {ph['candidates']:.2f} ms = "
          f"{2.0 * n * n * 2048 / ph['candidates'] / 1e9:.0f} TFLOP/s  finalize {ph['finalize']:.2f}  rerank {ph['rerank']:.2f}  "
          f"fallback {ph['fallback']:.2f}  stats={nn_.last_stats}")

# warp-role cycle counters of the pair kernel (vqacx_debug_set_knn_profile)
import ctypes
from vqa_counterexamples_b200._lib import lib
prof = torch.zeros(2 * 148 * 8, dtype=torch.int64, device="cuda")
lib.vqacx_debug_set_knn_profile.argtypes = [ctypes.c_void_p]
lib.vqacx_debug_set_knn_profile.restype = None
lib.vqacx_debug_set_knn_profile(ctypes.c_void_p(prof.data_ptr()))
nn_.kneighbors(x)
torch.cuda.synchronize()
lib.vqacx_debug_set_knn_profile(None)
p2 = prof.view(2, 148, 8)[1].double().cpu()
p = prof.view(2, 148, 8)[0].double().cpu()
names = ["prod wait-empty / (odd CTAs) epi push+compact", "prod total", "mma wait-full", "mma wait-tempty", "mma total", "epi wait-tfull", "epi total", "epi tmem ld+wait"]
print("odd-CTA slot 0 = epilogue push+compact cycles:", float(p[1::2, 0].mean()) / 1e6, "Mcyc; even-CTA slot 0 = producer wait-empty:", float(p[0::2, 0].mean()) / 1e6)
for j, nm in enumerate(names):
    col = p[:, j][p[:, j] > 0]
    if len(col):
        print(f"{nm:16s} mean {col.mean() / 1e6:8.2f} Mcyc  min {col.min() / 1e6:8.2f}  max {col.max() / 1e6:8.2f}  (n={len(col)})")
names2 = ["prod lock-step wait", "prod TMA issue", "mirror wait-tfull", "mirror scan", "mirror total"]
for j, nm in enumerate(names2):
    col = p2[:, j][p2[:, j] > 0]
    if len(col):
        print(f"{nm:20s} mean {col.mean() / 1e6:8.2f} Mcyc  min {col.min() / 1e6:8.2f}  max {col.max() / 1e6:8.2f}  (n={len(col)})")
