"""Decomposes the mirrored launch of the symmetric self-join (123 287 x 2048, k = 25) with the timing experiments of a
-DVQACX_DEBUG build (VQACX_DEBUG_BUILD=1 python -m vqa_counterexamples_b200.build): OPT_KNN_DEBUG bit 1 = no row-side
epilogue, 8 = no mirror scan, 16 = mirror hits found but not stored, 32 = mirror compares only. Every variant except 0
INVALIDATES results; only the candidate-phase time is read."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vqa_counterexamples_b200 import _lib, synth
from vqa_counterexamples_b200.knn import NearestNeighbors

assert _lib.is_debug_build(), "needs the -DVQACX_DEBUG build"
n = int(sys.argv[1]) if len(sys.argv) > 1 else 123287
variants = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0, 32, 16, 8, 1, 9, 0]
x = synth.clustered_features_cuda(n, 2048, seed=1234)


def run(reps=3):
    out = []
    for _ in range(reps):
        nn_ = NearestNeighbors(n_neighbors=25)
        nn_.profile = True
        nn_.fit(x)
        try:
            nn_.kneighbors(x)
        except Exception as exc:          # invalid results may trip the proof / fallbacks: the phase times are still set
            print("   (search raised:", type(exc).__name__, ")")
        torch.cuda.synchronize()
        out.append(round(nn_.last_phase_ms["candidates"], 2))
    return out


for v in variants:
    with _lib.option(_lib.OPT_KNN_DEBUG, v):
        print(f"debug={v:2d}: candidates ms {run()}", flush=True)
