"""Decomposes the mirrored launch of the symmetric self-join (123 287 x 2048, k = 25) with the timing experiments of a
-DVQACX_DEBUG build (VQACX_DEBUG_BUILD=1 python -m vqa_counterexamples_b200.build): OPT_KNN_DEBUG bit 1 = no row-side
epilogue, 8 = no mirror scan, 16 = mirror hits found but not parked in the ring, 32 = mirror compares only, 64 = per-lane hit
queues never overflow (no warp-wide flush path at all), 512 = ring flushed but not written to the log, 128 / 256 = log
written with an L2 evict_first hint / st.global.cs. Every variant except 0, 128 and 256 INVALIDATES results; only the
candidate-phase time is read."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vqa_counterexamples_b200 import _lib, synth
from vqa_counterexamples_b200.knn import NearestNeighbors

assert _lib.is_debug_build(), "needs the -DVQACX_DEBUG build"
n = int(sys.argv[1]) if len(sys.argv) > 1 else 123287
variants = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0, 64, 32, 8, 9]
x = synth.clustered_features_cuda(n, 2048, seed=1234)


def run_once():
    nn_ = NearestNeighbors(n_neighbors=25)
    nn_.profile = True
    nn_.fit(x)
    try:
        nn_.kneighbors(x)
    except Exception as exc:          # invalid results may trip the proof / fallbacks: the phase times are still set
        print("   (search raised:", type(exc).__name__, ")")
    torch.cuda.synchronize()
    return round(nn_.last_phase_ms["candidates"], 2)


# interleaved rounds: the chip drifts into its power limit over the first seconds, so variants are compared round by round
rounds = int(sys.argv[3]) if len(sys.argv) > 3 else 5
res = {v: [] for v in variants}
for _ in range(rounds):
    for v in variants:
        with _lib.option(_lib.OPT_KNN_DEBUG, v):
            res[v].append(run_once())
for v in variants:
    print(f"debug={v:3d}: candidates ms {res[v]}  median {sorted(res[v])[len(res[v]) // 2]}", flush=True)
