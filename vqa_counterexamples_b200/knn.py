"""Stage 1 of the VQA-CX hot path: exact k-nearest-neighbour retrieval over pooled image features.

Mirrors the reference entry point `knn.py` (CLI flags knn.py:11-21, flow knn.py:23-58) and the
sklearn-shaped pair it relies on: `NearestNeighbors(n_neighbors=k).fit(X)` / `.kneighbors(Q)`
(knn.py:41-42,49-51). All arithmetic runs in libvqacx.so (tcgen05 GEMM + fused top-k' + fp64
re-rank, see csrc/knn_gemm.cu); this file is host plumbing: device buffers, H2D/D2H, sharding.
"""
import argparse
import ctypes
import os

import numpy as np
import torch

from . import _lib
from ._lib import lib, check, ptr, current_stream


class _Prepped:
    """Device-resident float32 rows plus the side arrays the tensor path needs."""

    def __init__(self, x_f32, max_bits=None):
        assert x_f32.is_cuda and x_f32.dtype == torch.float32 and x_f32.dim() == 2 and x_f32.is_contiguous()
        n, d = x_f32.shape
        self.f32 = x_f32
        self.n, self.d = n, d
        dev = x_f32.device
        self.lp = torch.empty((n, d), dtype=torch.float16, device=dev)
        self.norm_sq = torch.empty((max(n, 1),), dtype=torch.float32, device=dev)
        self.max_bits = max_bits if max_bits is not None else torch.zeros(1, dtype=torch.int32, device=dev)
        self.bad = torch.zeros(1, dtype=torch.int32, device=dev)
        if n > 0:
            check(
                lib.vqacx_knn_prep(ptr(x_f32), n, d, ptr(self.lp), ptr(self.norm_sq), ptr(self.max_bits),
                                   ptr(self.bad), current_stream()),
                "vqacx_knn_prep",
            )


def _as_device_f32(x, device):
    if isinstance(x, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
        return t.to(device, non_blocking=False)
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=torch.float32).contiguous()
    raise TypeError(f"expected numpy array or torch tensor, got {type(x)}")


class NearestNeighbors:
    """Drop-in for `sklearn.neighbors.NearestNeighbors` as used by the reference (knn.py:41-51).

    Only what that call site exercises is supported: brute-force Euclidean search,
    `fit(X)` then `kneighbors(Q) -> (distances float64 [n,k], indices int64 [n,k])`, rows sorted by
    increasing distance (ties by index). numpy in -> numpy out; CUDA tensors in -> CUDA tensors out.
    """

    def __init__(self, n_neighbors=5, *, device=None, force_exact=False, row_offset=0):
        self.n_neighbors = int(n_neighbors)
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.force_exact = bool(force_exact)
        self.row_offset = int(row_offset)  # global index of local row 0 (row-sharded databases)
        self._db = None
        self._fit_src = None
        self._ws = None
        self.last_stats = None
        self.profile = False  # True: record CUDA events around each phase -> last_phase_ms
        self.last_phase_ms = None

    # -- sklearn API -------------------------------------------------------------------------
    def fit(self, X, y=None):
        with torch.cuda.device(self.device):
            self._db = _Prepped(_as_device_f32(X, self.device))
            self._fit_src = X if isinstance(X, np.ndarray) else None  # knn.py queries slices of the fitted array
            if int(self._db.bad.item()) != 0:
                raise _lib.VqacxError("features are not representable in fp16 (|x| > 65504 or non-finite)")
        return self

    def kneighbors(self, X=None, n_neighbors=None, return_distance=True):
        if self._db is None:
            raise RuntimeError("This NearestNeighbors instance is not fitted yet.")
        k = self.n_neighbors if n_neighbors is None else int(n_neighbors)
        want_numpy = X is None or isinstance(X, np.ndarray)
        with torch.cuda.device(self.device):
            if X is None:
                # sklearn semantics: the query set is the training set and each point is not its own neighbour
                dist, idx = self._search(self._db, k + 1)
                dist, idx = _drop_self(dist, idx, self.row_offset)
            else:
                if X is self._fit_src:
                    q = self._db.f32  # same host array that was fitted: already resident, no second upload
                else:
                    q = _as_device_f32(X, self.device)
                if q.dim() != 2 or q.shape[1] != self._db.d:
                    raise ValueError(f"query has {q.shape[1]} features, index has {self._db.d}")
                same = q.data_ptr() == self._db.f32.data_ptr() and q.shape == self._db.f32.shape
                qp = self._db if same else _Prepped(q, max_bits=self._db.max_bits.clone())
                dist, idx = self._search(qp, k)
        if want_numpy:
            dist, idx = dist.cpu().numpy(), idx.cpu().numpy()
        return (dist, idx) if return_distance else idx

    # -- device-side search -------------------------------------------------------------------
    def _search(self, qp, k):
        db = self._db
        if k > _lib.VQACX_KNN_MAX_K:
            raise ValueError(f"n_neighbors={k} exceeds the supported maximum {_lib.VQACX_KNN_MAX_K}")
        if k > db.n:
            raise ValueError(f"Expected n_neighbors <= n_samples_fit, but n_neighbors = {k}, n_samples_fit = {db.n}")
        dev = db.f32.device
        out_dist = torch.empty((qp.n, k), dtype=torch.float64, device=dev)
        out_idx = torch.empty((qp.n, k), dtype=torch.int64, device=dev)
        need = lib.vqacx_knn_workspace_bytes(qp.n, db.n, db.d)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        stats = (ctypes.c_int32 * _lib.VQACX_KNN_NSTATS)()
        # the max norm must cover database AND query rows (error bound uses |x| + |y|max with y in the db only,
        # |x| is recomputed exactly per query) -> database max is what matters; qp.max_bits is a superset.
        flags = _lib.VQACX_KNN_FORCE_EXACT if self.force_exact else 0
        phases = (ctypes.c_float * _lib.VQACX_KNN_NPHASES)() if self.profile else None
        check(
            lib.vqacx_knn_search(ptr(db.f32), ptr(db.lp), ptr(db.norm_sq), ptr(db.max_bits), db.n, self.row_offset,
                                 ptr(qp.f32), ptr(qp.lp), ptr(qp.norm_sq), qp.n, db.d, k, ptr(out_dist), ptr(out_idx),
                                 ptr(self._ws), self._ws.numel(), flags, stats, phases, current_stream()),
            "vqacx_knn_search",
        )
        self.last_stats = {
            "flagged_rows": stats[0], "unresolved_rows": stats[1], "bound_violations": stats[2],
            "tensor_path": bool(stats[3]),
            # candidates (of 64) whose float64 distance was evaluated, average over a 1-in-16 sample of the queries
            "rerank_evaluated_avg": round(stats[4] / max(1, (int(qp.n) + 15) // 16), 2),
        }
        if phases is not None:
            self.last_phase_ms = {"candidates": phases[0], "finalize": phases[1], "rerank": phases[2],
                                  "fallback": phases[3]}
        return out_dist, out_idx


def _drop_self(dist, idx, row_offset):
    """Remove each row's own index (or, if duplicates displaced it, the last column) -> k columns."""
    n, k1 = idx.shape
    own = torch.arange(n, device=idx.device, dtype=idx.dtype).unsqueeze(1) + row_offset
    is_self = idx == own
    has_self = is_self.any(dim=1, keepdim=True)
    drop = torch.where(has_self, is_self, torch.zeros_like(is_self))
    drop[:, -1] |= ~has_self.squeeze(1)
    keep = ~drop
    return dist[keep].view(n, k1 - 1), idx[keep].view(n, k1 - 1)


def merge_topk(dist_g, idx_g):
    """[G, n, k] per-shard results -> global [n, k] by (distance, index). Device tensors."""
    G, n, k = idx_g.shape
    out_d = torch.empty((n, k), dtype=torch.float64, device=idx_g.device)
    out_i = torch.empty((n, k), dtype=torch.int64, device=idx_g.device)
    check(lib.vqacx_topk_merge(ptr(dist_g.contiguous()), ptr(idx_g.contiguous()), G, n, k, ptr(out_d), ptr(out_i),
                               current_stream()), "vqacx_topk_merge")
    return out_d, out_i


def shard_rows(n, rank, world):
    """Contiguous database rows [row0, row1) owned by `rank` (row0 = rank * ceil(n / world))."""
    per = (n + world - 1) // world
    return min(n, rank * per), min(n, (rank + 1) * per)


def _local_topk_cuda(local_rows, row0, queries, k, force_exact=False, profile=False):
    nn_ = NearestNeighbors(n_neighbors=k, device=local_rows.device, force_exact=force_exact, row_offset=row0)
    nn_.profile = profile
    if local_rows.shape[0] == 0:  # more ranks than rows: this shard contributes only padding
        nq = queries.shape[0]
        return (torch.empty((nq, 0), dtype=torch.float64, device=local_rows.device),
                torch.empty((nq, 0), dtype=torch.int64, device=local_rows.device), nn_)
    nn_.fit(local_rows)
    kk = min(k, nn_._db.n)
    d_loc, i_loc = nn_.kneighbors(queries, n_neighbors=kk)
    return d_loc, i_loc, nn_


def sharded_kneighbors(local_rows, row0, queries, k, group=None, force_exact=False, profile=False,
                       return_index=False, local_topk=None, merge=None):
    """Row-sharded exact kNN: every rank holds `local_rows` = global rows [row0, row0+n_local),
    all ranks hold the same `queries`; per-shard top-k (global row ids) -> all_gather -> on-device merge by
    (distance, index). `local_topk` / `merge` default to the CUDA paths; the world_size-2 gloo tests substitute CPU
    stand-ins to exercise this host protocol (padding of short shards, gather layout, offsets) without a GPU."""
    import torch.distributed as dist

    local_topk = local_topk or (lambda rows, r0, q, kk: _local_topk_cuda(rows, r0, q, kk, force_exact, profile))
    merge = merge or merge_topk
    d_loc, i_loc, nn_ = local_topk(local_rows, row0, queries, k)
    last = (lambda: nn_) if return_index else (lambda: getattr(nn_, "last_stats", None))
    kk = d_loc.shape[1]
    if kk < k:  # tiny (or empty) shard: pad to k so the gather is rectangular
        pad_d = torch.full((d_loc.shape[0], k - kk), float("inf"), dtype=d_loc.dtype, device=d_loc.device)
        pad_i = torch.full((i_loc.shape[0], k - kk), -1, dtype=i_loc.dtype, device=i_loc.device)
        d_loc, i_loc = torch.cat([d_loc, pad_d], 1), torch.cat([i_loc, pad_i], 1)
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return d_loc, i_loc, last()
    d_all = torch.empty((world,) + tuple(d_loc.shape), dtype=d_loc.dtype, device=d_loc.device)
    i_all = torch.empty((world,) + tuple(i_loc.shape), dtype=i_loc.dtype, device=i_loc.device)
    if d_loc.is_cuda:
        dist.all_gather_into_tensor(d_all, d_loc.contiguous(), group=group)
        dist.all_gather_into_tensor(i_all, i_loc.contiguous(), group=group)
    else:  # gloo has no all_gather_into_tensor
        dist.all_gather(list(d_all.unbind(0)), d_loc.contiguous(), group=group)
        dist.all_gather(list(i_all.unbind(0)), i_loc.contiguous(), group=group)
    d, i = merge(d_all, i_all)
    return d, i, last()


# ------------------------------------------------------------------------------------------------
# CLI: same flags and output file as the reference's knn.py
# ------------------------------------------------------------------------------------------------
parser = argparse.ArgumentParser(description="Exact kNN over pooled image features (B200-native knn.py)")
parser.add_argument("base_dir", type=str, help="path to dir containing extracted hdf5 features")
parser.add_argument("--hdf5_file", type=str, default="trainset.hdf5")
parser.add_argument("--save_dir", type=str, default="data/coco/extract")
parser.add_argument("--save_file", type=str, default="knn_results_trainset.npy")
parser.add_argument("-k", "--k", default=25, type=int)
parser.add_argument("-b", "--batch_size", default=10, type=int,
                    help="kept for CLI compatibility; the GPU path searches all rows in one call")


def load_features(path):
    """`noatt` dataset of an extract.py HDF5 file (extract.py:105-109), or a .npy array of the same shape."""
    if path.endswith(".npy"):
        return np.load(path)
    try:
        import h5py
    except ImportError as e:  # pragma: no cover - h5py is not in this image
        raise ImportError("reading .hdf5 feature files needs h5py; pass a .npy file instead") from e
    with h5py.File(path, "r") as f:
        return np.array(f.get("noatt"))


def main(argv=None):
    args = parser.parse_args(argv)
    load_path = os.path.join(args.base_dir, args.hdf5_file)
    assert os.path.exists(load_path)
    assert os.path.isdir(args.save_dir)
    save_path = os.path.join(args.save_dir, args.save_file)
    if os.path.exists(save_path):
        print("Warning: {} already exists and will be overwritten.".format(save_path))
    print("Saving results to {}".format(save_path))

    features = load_features(load_path)
    print("Loaded features as array of size {}".format(features.shape))

    print("Fitting KNN model for k={}".format(args.k))
    nbrs = NearestNeighbors(n_neighbors=args.k)
    nbrs.fit(features)

    print("Starting KNN computations...")
    all_distances, all_indices = nbrs.kneighbors(nbrs._db.f32)
    all_distances, all_indices = all_distances.cpu().numpy(), all_indices.cpu().numpy()
    print("kNN stats: {}".format(nbrs.last_stats))

    np.save(save_path, {"indices": all_indices, "distances": all_distances})
    return all_distances, all_indices


if __name__ == "__main__":
    main()
