"""The pooled-feature producer of extract.py (the step before stage 1), B200-native.

extract.py:121-124 turns the CNN's activation maps `output_att` [b, 2048, 14, 14] into the `noatt` rows the kNN stage
indexes: `output_att.sum(3).sum(2).div(nb_regions).view(-1, 2048)`. `pool_noatt` is that line as one CUDA pass;
`PooledIndex` additionally collects, in the SAME pass over the maps, what `NearestNeighbors.fit` would otherwise compute
from the pooled rows (fp16 copy, norms), so the index is ready the moment the last batch has been pooled. The CNN itself
(FBResNet-152, convnets.py) and the HDF5 / .txt file contract stay outside the hot path."""
import torch

from ._lib import lib, check, ptr, current_stream


def _maps(output_att):
    if not (output_att.is_cuda and output_att.dim() == 4):
        raise ValueError("expected a CUDA tensor [b, C, H, W] (there is no CPU fallback for this path)")
    a = output_att.to(torch.float32).contiguous()
    b, c, h, w = a.shape
    return a, b, c, h * w


def pool_noatt(output_att):
    """extract.py:123-124: [b, C, H, W] -> [b, C] spatial mean (float32)."""
    a, b, c, hw = _maps(output_att)
    out = torch.empty((b, c), dtype=torch.float32, device=a.device)
    check(lib.vqacx_pool_features(ptr(a), b, c, hw, ptr(out), None, None, None, None, current_stream()),
          "vqacx_pool_features")
    return out


class PooledIndex:
    """Accumulates pooled rows batch by batch (the loop at extract.py:117-130, `hdf5_noatt[idx:idx+batch_size] = ...`)
    into device-resident arrays and hands them to the kNN stage without a second pass: `index.neighbors(k)` returns a
    fitted `knn.NearestNeighbors`."""

    def __init__(self, n_images, dim=2048, device=None):
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.n, self.dim, self.idx = int(n_images), int(dim), 0
        self.f32 = torch.empty((self.n, self.dim), dtype=torch.float32, device=dev)
        self.lp = torch.empty((self.n, self.dim), dtype=torch.float16, device=dev)
        self.norm_sq = torch.empty((max(self.n, 1),), dtype=torch.float32, device=dev)
        self.max_bits = torch.zeros(1, dtype=torch.int32, device=dev)
        self.bad = torch.zeros(1, dtype=torch.int32, device=dev)

    def add(self, output_att):
        a, b, c, hw = _maps(output_att)
        if c != self.dim or self.idx + b > self.n:
            raise ValueError(f"batch of {b} x {c} does not fit the index ({self.idx}/{self.n} rows of {self.dim})")
        i = self.idx
        check(lib.vqacx_pool_features(ptr(a), b, c, hw, ptr(self.f32[i:]), ptr(self.lp[i:]), ptr(self.norm_sq[i:]),
                                      ptr(self.max_bits), ptr(self.bad), current_stream()), "vqacx_pool_features")
        self.idx += b
        return self.f32[i:i + b]

    def neighbors(self, n_neighbors=25):
        from .knn import NearestNeighbors

        if self.idx != self.n:
            raise RuntimeError(f"only {self.idx} of {self.n} rows have been pooled")
        return NearestNeighbors(n_neighbors=n_neighbors, device=self.f32.device).fit_prepared(
            self.f32, self.lp, self.norm_sq, self.max_bits, self.bad)
