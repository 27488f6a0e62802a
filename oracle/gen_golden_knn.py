"""Generates tests/golden/knn_*.npz by running the reference's own kNN call sequence
(knn.py:41-56 -> scikit-learn NearestNeighbors) on seeded synthetic features.

Run in the build container:  python oracle/gen_golden_knn.py
Inputs are NOT stored: they are regenerated from (n, d, seed, dup_frac) by
vqa_counterexamples_b200.synth.clustered_features_np (numpy Philox, deterministic)."""
import importlib.util
import os
import sys

import numpy as np
import sklearn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


synth = _load("synth", os.path.join(ROOT, "vqa_counterexamples_b200", "synth.py"))
from oracle.knn_oracle import knn_reference_sklearn, knn_sklearn_onecall  # noqa: E402

CASES = [
    # name, n, d, seed, dup_frac, k, n_queries (None = self-join through the knn.py batch loop)
    ("small_d128", 1500, 128, 11, 0.0, 25, None),
    ("dups_d64", 1000, 64, 12, 0.02, 25, None),
    ("coco_d2048", 4096, 2048, 13, 0.0, 25, 256),
]


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, n, d, seed, dup, k, nq in CASES:
        x = synth.clustered_features_np(n, d, seed=seed, dup_frac=dup)
        if nq is None:
            dist, idx = knn_reference_sklearn(x, k=k, batch_size=10)
        else:
            dist, idx = knn_sklearn_onecall(x, x[:nq], k=k)
        np.savez_compressed(
            os.path.join(out_dir, f"knn_{name}.npz"),
            distances=dist, indices=idx.astype(np.int32),
            meta=np.array([n, d, seed, k, -1 if nq is None else nq], dtype=np.int64), dup_frac=np.float64(dup),
            sklearn_version=np.array(sklearn.__version__), generator_version=np.int64(synth.GENERATOR_VERSION),
        )
        print(name, dist.shape, "checksum", float(dist.sum()), int(idx.sum()))


if __name__ == "__main__":
    main()
