"""CPU: pin the kNN oracle (oracle/knn_oracle.py) against the golden vectors produced by the reference's own
call sequence on scikit-learn (oracle/gen_golden_knn.py), and exercise the tie-aware comparator."""
import glob
import os

import numpy as np
import pytest

from oracle.knn_oracle import compare_knn, knn_restatement_f64, knn_reference_sklearn
from vqa_counterexamples_b200 import synth

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def load_case(path):
    z = np.load(path)
    n, d, seed, k, nq = [int(v) for v in z["meta"]]
    x = synth.clustered_features_np(n, d, seed=seed, dup_frac=float(z["dup_frac"]))
    q = x if nq < 0 else x[:nq]
    return x, q, k, z["distances"], z["indices"].astype(np.int64)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "knn_*.npz"))))
@pytest.mark.parametrize("direct", [False, True])
def test_restatement_matches_sklearn_golden(path, direct):
    x, q, k, gd, gi = load_case(path)
    d, i = knn_restatement_f64(x, q, k=k, direct=direct)
    rep = compare_knn(d, i, gd, gi)
    assert rep["ok"], rep
    # self is the nearest neighbour at ~0 (sklearn's cancellation noise is < 2e-6)
    assert np.all(gd[:, 0] < 2e-6)
    assert np.all(np.diff(d, axis=1) >= 0)


def test_reference_sequence_small_batches_equals_one_call():
    x = synth.clustered_features_np(300, 64, seed=5)
    d1, i1 = knn_reference_sklearn(x, k=10, batch_size=10)
    d2, i2 = knn_reference_sklearn(x, k=10, batch_size=300)
    assert compare_knn(d1, i1, d2, i2)["ok"]


def test_comparator_rejects_wrong_neighbour():
    x = synth.clustered_features_np(200, 64, seed=6)
    d, i = knn_restatement_f64(x, x, k=5)
    bad_i = i.copy()
    bad_i[3, 2] = (bad_i[3, 2] + 1) % 200
    assert not compare_knn(d, bad_i, d, i)["ok"]
    bad_d = d.copy()
    bad_d[7, 4] *= 1.01
    assert not compare_knn(bad_d, i, d, i)["ok"]


def test_comparator_accepts_tie_swap():
    d = np.array([[0.0, 1.0, 1.0, 2.0]])
    i = np.array([[0, 5, 9, 3]])
    j = np.array([[0, 9, 5, 3]])
    assert compare_knn(d, j, d, i)["ok"]
