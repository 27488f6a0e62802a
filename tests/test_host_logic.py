"""CPU tests of the host-side logic either side of the CUDA path (no GPU, no compute calls into libvqacx.so):
option merging, batchify order, sharding, and the world_size-2 protocols (gloo) of the row-sharded kNN and the
data-parallel NeuralCX reductions. The CUDA kernels behind these protocols are covered by the -m gpu tests."""
import os
import random
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_update_values_cli_over_yaml():
    from vqa_counterexamples_b200.counterexamples import update_values

    cli = {"optim": {"lr": None, "batch_size": 32, "epochs": None}, "cx_model": {"pretrained_vqa": True, "trainable_vqa": False}}
    yml = {"optim": {"lr": 1e-4, "batch_size": 64, "epochs": 100}, "cx_model": {"name": "n", "dim_h": 300, "trainable_vqa": True},
           "model": {"arch": "MutanNoAtt"}}
    out = update_values(cli, yml)
    assert out is yml
    assert out["optim"] == {"lr": 1e-4, "batch_size": 32, "epochs": 100}          # None leaves do not overwrite
    assert out["cx_model"]["trainable_vqa"] is False and out["cx_model"]["pretrained_vqa"] is True
    assert out["cx_model"]["dim_h"] == 300 and out["model"]["arch"] == "MutanNoAtt"
    with pytest.raises(KeyError):                                                  # reference behaviour (utils.py:9)
        update_values({"missing": {"a": 1}}, {})


def test_load_options_reads_reference_yaml_keys(tmp_path):
    from vqa_counterexamples_b200.counterexamples import load_options, parser

    p = tmp_path / "opt.yaml"
    p.write_text("optim:\n  lr: 0.0001\n  batch_size: 64\n  epochs: 100\ncx_model:\n  name: neuralcx_300_2_all\n"
                 "  dim_h: 300\n  n_layers: 2\n  drop_p: 0.25\n  pretrained_emb: true\n  v_emb: true\n"
                 "model:\n  arch: MutanNoAtt\n  fusion:\n    dim_v: 2048\n")
    args = parser.parse_args(["--cx_model", "NeuralModel", "--path_opt", str(p), "-b", "16", "--epochs", "2"])
    opt = load_options(args)
    assert opt["optim"] == {"lr": 0.0001, "batch_size": 16, "epochs": 2}
    assert opt["cx_model"]["dim_h"] == 300 and opt["cx_model"]["pretrained_vqa"] is True
    assert opt["vgenome"] is None


def test_batchify_matches_reference_order():
    """Probe of the reference: seed(42); batchify(range(10), 4) -> [[7,3,2,8],[5,6,9,4],[0,1]]."""
    from vqa_counterexamples_b200.counterexamples import batchify

    random.seed(42)
    ex = list(range(10))
    assert batchify(ex, 4) == [[7, 3, 2, 8], [5, 6, 9, 4], [0, 1]]
    assert ex == [7, 3, 2, 8, 5, 6, 9, 4, 0, 1]                                  # shuffled in place
    assert batchify(list(range(5)), 2, shuffle=False) == [[0, 1], [2, 3], [4]]
    assert batchify([], 3) == []


def test_shard_ranges_cover_exactly_once():
    from vqa_counterexamples_b200.counterexamples import shard_range
    from vqa_counterexamples_b200.knn import shard_rows

    for fn in (shard_range, shard_rows):
        for n in (0, 1, 7, 64, 123287, 1000000):
            for world in (1, 2, 3, 4, 8):
                spans = [fn(n, r, world) for r in range(world)]
                assert spans[0][0] == 0 and spans[-1][1] == n
                assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
                assert all(0 <= e - b <= (n + world - 1) // world for b, e in spans)


# ------------------------------------------------------------------------------------------------
# world_size-2 gloo workers
# ------------------------------------------------------------------------------------------------
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _init(rank, world, port):
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)


def _cpu_local_topk(rows, row0, queries, k):
    """CPU stand-in for the CUDA shard search: the float64 oracle on this rank's rows, global ids."""
    from oracle.knn_oracle import knn_restatement_f64

    kk = min(k, rows.shape[0])
    if kk == 0:
        n = queries.shape[0]
        return torch.empty((n, 0), dtype=torch.float64), torch.empty((n, 0), dtype=torch.int64), None
    d, i = knn_restatement_f64(rows.numpy(), queries.numpy(), k=kk, direct=True)
    return torch.from_numpy(d), torch.from_numpy(i + row0), None


def _cpu_merge(d_all, i_all):
    """CPU stand-in for vqacx_topk_merge: global top-k by (distance, index)."""
    G, n, k = i_all.shape
    d = d_all.permute(1, 0, 2).reshape(n, G * k).numpy()
    i = i_all.permute(1, 0, 2).reshape(n, G * k).numpy()
    order = np.lexsort((i, d), axis=1)[:, :k]
    return torch.from_numpy(np.take_along_axis(d, order, 1)), torch.from_numpy(np.take_along_axis(i, order, 1))


def _knn_worker(rank, world, port, n_db, ret):
    _init(rank, world, port)
    from oracle.knn_oracle import knn_restatement_f64
    from vqa_counterexamples_b200 import synth
    from vqa_counterexamples_b200.knn import shard_rows, sharded_kneighbors

    x = torch.from_numpy(synth.clustered_features_np(n_db, 64, seed=5))
    q = x[:17]
    r0, r1 = shard_rows(n_db, rank, world)
    k = 6
    d, i, _ = sharded_kneighbors(x[r0:r1], r0, q, k, local_topk=_cpu_local_topk, merge=_cpu_merge)
    ref_d, ref_i = knn_restatement_f64(x.numpy(), q.numpy(), k=min(k, n_db), direct=True)
    kk = ref_i.shape[1]
    ok = bool(np.array_equal(i.numpy()[:, :kk], ref_i) and np.allclose(d.numpy()[:, :kk], ref_d, rtol=0, atol=1e-12))
    if kk < k:
        ok = ok and bool((i.numpy()[:, kk:] == -1).all() and np.isinf(d.numpy()[:, kk:]).all())
    ret[rank] = ok
    dist.destroy_process_group()


@pytest.mark.parametrize("n_db", [200, 7, 1])
def test_sharded_knn_protocol_world2_gloo(n_db):
    """Row-sharded search: per-shard top-k with global ids -> all_gather -> merge equals the single-process answer,
    including shards shorter than k (padding with +inf / -1) and an empty shard (n_db = 1 over 2 ranks)."""
    port = _free_port()
    ret = mp.Manager().dict()
    mp.spawn(_knn_worker, args=(2, port, n_db, ret), nprocs=2, join=True)
    assert ret[0] and ret[1]


class _CpuShard:
    """CPU stand-in for knn._CudaShard: phase 1 publishes this shard's k smallest exact distances inflated by 1e-3
    (a valid upper bound, like the CUDA error model's hi), phase 2 returns only the candidates that can beat the agreed
    bound (+inf / -1 elsewhere), as the pruned CUDA re-rank does."""

    def __init__(self, rows, row0, queries, k):
        from oracle.knn_oracle import knn_restatement_f64

        self.k, self.n = k, queries.shape[0]
        kk = min(k, rows.shape[0])
        if kk == 0:
            self.d = torch.empty((self.n, 0), dtype=torch.float64)
            self.i = torch.empty((self.n, 0), dtype=torch.int64)
        else:
            d, i = knn_restatement_f64(rows.numpy(), queries.numpy(), k=kk, direct=True)
            self.d, self.i = torch.from_numpy(d), torch.from_numpy(i + row0)

    def _pad(self, t, v):
        pad = torch.full((self.n, self.k - t.shape[1]), v, dtype=t.dtype)
        return torch.cat([t, pad], 1)

    def phase1(self):
        return self._pad((self.d * (1 + 1e-3) + 1e-6).float(), float("inf"))

    def phase2(self, bound):
        keep = self.d <= bound.double().unsqueeze(1)
        d = torch.where(keep, self.d, torch.full_like(self.d, float("inf")))
        i = torch.where(keep, self.i, torch.full_like(self.i, -1))
        return self._pad(d, float("inf")), self._pad(i, -1)


def _cpu_kth(hi_all):
    G, n, k = hi_all.shape
    return hi_all.permute(1, 0, 2).reshape(n, G * k).sort(dim=1).values[:, k - 1].contiguous()


def _knn_two_phase_worker(rank, world, port, n_db, ret):
    _init(rank, world, port)
    from oracle.knn_oracle import knn_restatement_f64
    from vqa_counterexamples_b200 import synth
    from vqa_counterexamples_b200.knn import shard_rows, sharded_kneighbors

    x = torch.from_numpy(synth.clustered_features_np(n_db, 64, seed=5))
    q = x[:17]
    r0, r1 = shard_rows(n_db, rank, world)
    k = 6
    d, i, _ = sharded_kneighbors(x[r0:r1], r0, q, k, shard_factory=_CpuShard, kth=_cpu_kth, merge=_cpu_merge)
    ref_d, ref_i = knn_restatement_f64(x.numpy(), q.numpy(), k=min(k, n_db), direct=True)
    kk = ref_i.shape[1]
    ok = bool(np.array_equal(i.numpy()[:, :kk], ref_i) and np.allclose(d.numpy()[:, :kk], ref_d, rtol=0, atol=1e-12))
    if kk < k:
        ok = ok and bool((i.numpy()[:, kk:] == -1).all() and np.isinf(d.numpy()[:, kk:]).all())
    # result="sliced": the answer stays distributed by query, rank g holds rows shard_rows(n_q, g, world) (17 queries
    # over 2 ranks: 9 + 8, so the exchange pads the last slice)
    ds, is_, _ = sharded_kneighbors(x[r0:r1], r0, q, k, shard_factory=_CpuShard, kth=_cpu_kth, merge=_cpu_merge,
                                    result="sliced")
    q0, q1 = shard_rows(q.shape[0], rank, world)
    ok = ok and tuple(is_.shape) == (q1 - q0, k) and bool(np.array_equal(is_.numpy()[:, :kk], ref_i[q0:q1]))
    ok = ok and bool(np.allclose(ds.numpy()[:, :kk], ref_d[q0:q1], rtol=0, atol=1e-12))
    ret[rank] = ok
    dist.destroy_process_group()


@pytest.mark.parametrize("n_db", [200, 7, 1])
def test_sharded_knn_two_phase_protocol_world2_gloo(n_db):
    """Two-exchange protocol: phase-1 bounds -> all_gather -> global k-th bound -> pruned phase-2 results ->
    all_gather -> merge equals the single-process answer (short and empty shards included)."""
    port = _free_port()
    ret = mp.Manager().dict()
    mp.spawn(_knn_two_phase_worker, args=(2, port, n_db, ret), nprocs=2, join=True)
    assert ret[0] and ret[1]


def _dp_worker(rank, world, port, ret):
    _init(rank, world, port)
    from vqa_counterexamples_b200.counterexamples import DataParallelCX, shard_batch

    dp = DataParallelCX()
    assert dp.on and dp.world == world and dp.rank == rank
    # gradient protocol: every rank back-propagates sum-CE(slice) / global batch -> SUM all-reduce == full-batch gradient
    g = torch.Generator().manual_seed(0)
    per_example = torch.randn(10, 33, generator=g, dtype=torch.float64)        # one "gradient" per example
    ids = list(range(10))
    mine = shard_batch(ids, rank, world)
    local = per_example[mine].sum(0) / len(ids) if len(mine) else torch.zeros(33, dtype=torch.float64)
    total = dp.allreduce_grads(local.clone())
    ok = torch.allclose(total, per_example.sum(0) / len(ids), atol=1e-14)
    # metric protocol
    loss_sum = torch.tensor(1.5 + rank, dtype=torch.float32)
    hits = torch.tensor([3 + rank, 7 + 2 * rank], dtype=torch.int32)
    ls, h1, h5, n = dp.reduce_metrics(loss_sum, hits, 5 + rank)
    ok = ok and abs(ls - 4.0) < 1e-9 and (h1, h5, n) == (7, 16, 11)
    # an uneven last batch: 3 examples over 2 ranks -> 2 + 1; 1 example -> 1 + 0 (the empty rank still reduces)
    ok = ok and [len(shard_batch(list(range(3)), r, 2)) for r in range(2)] == [2, 1]
    ok = ok and [len(shard_batch(list(range(1)), r, 2)) for r in range(2)] == [1, 0]
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_data_parallel_reductions_world2_gloo():
    port = _free_port()
    ret = mp.Manager().dict()
    mp.spawn(_dp_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret[0] and ret[1]


def test_bench_reference_arm_prints_one_json_line_and_uses_all_threads():
    """`bench.py --impl reference` (task statement (4)): stdout is exactly one JSON line with the contract's keys, also
    when torchrun has exported OMP_NUM_THREADS=1 to the worker (the arm must still use every host thread)."""
    import json
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--n-db", "3000",
                          "--steps", "1", "--warmup", "0", "--ref-queries", "20"], env=env, capture_output=True,
                         text=True, timeout=600, check=True).stdout
    lines = [ln for ln in out.split("\n") if ln.strip()]
    assert len(lines) == 1, out
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["metric"] == "knn_queries_per_sec" and j["unit"] == "queries/s"
    assert j["e2e"] == {"value": j["value"], "unit": j["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = j["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["value"] == j["value"]
    n_host = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    assert cb["cores"] >= n_host
