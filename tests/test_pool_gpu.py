"""GPU: the pooled-feature producer against extract.py:123-124 evaluated in float64, and the fused
fit (fp16 copy + norms emitted by the pooling pass) against the regular NearestNeighbors.fit on the pooled rows."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _maps(seed, b, c=2048, h=14, w=14):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.relu(torch.randn((b, c, h, w), generator=g, device="cuda") * 0.7 + 0.2)   # post-ReLU activations


@pytest.mark.parametrize("shape", [(5, 2048, 14, 14), (3, 64, 7, 7), (2, 16, 3, 5)])
def test_pool_matches_extract_line(shape):
    from vqa_counterexamples_b200.extract import pool_noatt

    att = _maps(1, *shape)
    got = pool_noatt(att)
    b, c, h, w = shape
    ref = att.double().sum(3).sum(2).div(h * w).view(-1, c)          # extract.py:123-124 in float64
    assert got.shape == (b, c)
    assert torch.allclose(got.double(), ref, rtol=2e-6, atol=1e-7)   # fp32 summation order differs from torch's


def test_pooled_index_equals_fit_on_pooled_rows():
    from vqa_counterexamples_b200.extract import PooledIndex, pool_noatt
    from vqa_counterexamples_b200.knn import NearestNeighbors

    n, c = 700, 256
    batches = [_maps(10 + i, 100, c, 7, 7) for i in range(7)]
    index = PooledIndex(n, dim=c)
    rows = [index.add(a) for a in batches]
    pooled = torch.cat([pool_noatt(a) for a in batches])
    assert torch.equal(torch.cat(rows), pooled)
    nn_a = index.neighbors(10)
    nn_b = NearestNeighbors(n_neighbors=10).fit(pooled)
    assert torch.equal(nn_a._db.lp, nn_b._db.lp)
    assert torch.allclose(nn_a._db.norm_sq[:n], nn_b._db.norm_sq[:n], rtol=1e-6)
    assert abs(float(nn_a._db.max_bits.view(torch.float32)) / float(nn_b._db.max_bits.view(torch.float32)) - 1) < 1e-6
    da, ia = nn_a.kneighbors(pooled)
    db, ib = nn_b.kneighbors(pooled)
    assert torch.equal(ia, ib) and torch.allclose(da, db)
    assert nn_a.last_stats["bound_violations"] == 0
