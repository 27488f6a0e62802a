"""Where the host-in / host-out kNN call spends its time: python tools/time_e2e.py [N]"""
import sys
import time

import torch

sys.path.insert(0, ".")
from vqa_counterexamples_b200 import synth
from vqa_counterexamples_b200.knn import NearestNeighbors

n = int(sys.argv[1]) if len(sys.argv) > 1 else 123287
x = synth.clustered_features_cuda(n, 2048, seed=1234)
h = torch.empty((n, 2048), dtype=torch.float32, pin_memory=True)
h.copy_(x)
torch.cuda.synchronize()
x_np = h.numpy()


def timed(fn, reps=5):
    out = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize()
        out.append((time.perf_counter() - t0) * 1e3)
    return sorted(out)[len(out) // 2], r


ms, xd = timed(lambda: torch.from_numpy(x_np).to("cuda"))
print(f"H2D {x_np.nbytes / 1e9:.2f} GB: {ms:.2f} ms = {x_np.nbytes / ms / 1e6:.1f} GB/s")
ms, nn_ = timed(lambda: NearestNeighbors(n_neighbors=25).fit(x_np))
print(f"fit(host): {ms:.2f} ms")
ms, (d, i) = timed(lambda: nn_.kneighbors(x))
print(f"kneighbors(device tensor): {ms:.2f} ms")
ms, _ = timed(lambda: (d.cpu().numpy(), i.cpu().numpy()))
print(f"D2H pageable: {ms:.2f} ms")
from vqa_counterexamples_b200.knn import _to_host
ms, _ = timed(lambda: _to_host(d, i))
print(f"D2H pinned staging: {ms:.2f} ms")
ms, _ = timed(lambda: NearestNeighbors(n_neighbors=25).fit(x_np).kneighbors(x_np))
print(f"e2e fit(host) + kneighbors(host): {ms:.2f} ms = {n / ms * 1e3:.0f} q/s")
for C in (1, 2, 3, 4, 6, 8):
    ms, (dd, ii) = timed(lambda: NearestNeighbors(n_neighbors=25).fit_kneighbors(x_np, n_chunks=C))
    print(f"fit_kneighbors(host, n_chunks={C}): {ms:.2f} ms = {n / ms * 1e3:.0f} q/s")
assert (ii == i.cpu().numpy()).all() and (dd == d.cpu().numpy()).all(), "pipelined self-join differs from fit + kneighbors"
print("pipelined results identical to fit + kneighbors")
