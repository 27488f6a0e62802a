"""Seeded synthetic inputs of the COCO / VQA-CX shapes.

Pooled FBResNet-152 features are spatial means of post-ReLU maps (extract.py:121-124): non-negative,
about half zeros, float32 [N, 2048]. i.i.d. Gaussians in 2048-d are all nearly equidistant, which would
turn kNN parity into a tie lottery, so rows are drawn around cluster centres.
"""
import numpy as np
import torch

GENERATOR_VERSION = 1


def clustered_features_np(n, d=2048, seed=1234, n_centres=None, dup_frac=0.0):
    """numpy float32 [n, d]; deterministic for (n, d, seed, n_centres, dup_frac, GENERATOR_VERSION)."""
    rng = np.random.Generator(np.random.Philox(seed))
    c = n_centres if n_centres is not None else max(1, n // 60)
    centres = np.maximum(rng.normal(0.35, 0.5, size=(c, d)), 0.0).astype(np.float32)
    assign = rng.integers(0, c, size=n)
    x = np.empty((n, d), dtype=np.float32)
    step = 8192
    for i in range(0, n, step):
        j = min(n, i + step)
        noise = rng.normal(0.0, 0.15, size=(j - i, d)).astype(np.float32)
        x[i:j] = np.maximum(centres[assign[i:j]] + noise, 0.0)
    if dup_frac > 0.0 and n > 1:
        m = max(1, int(n * dup_frac))
        dst = rng.choice(n, size=m, replace=False)
        src = rng.integers(0, n, size=m)
        x[dst] = x[src]
    return x


def clustered_features_cuda(n, d=2048, seed=1234, n_centres=None, device="cuda"):
    """Same law generated on the device (different stream than the numpy version; used at bench sizes)."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    c = n_centres if n_centres is not None else max(1, n // 60)
    centres = torch.clamp_min(torch.randn((c, d), generator=g, device=device) * 0.5 + 0.35, 0.0)
    assign = torch.randint(0, c, (n,), generator=g, device=device)
    x = torch.empty((n, d), dtype=torch.float32, device=device)
    step = 16384
    for i in range(0, n, step):
        j = min(n, i + step)
        noise = torch.randn((j - i, d), generator=g, device=device) * 0.15
        x[i:j] = torch.clamp_min(centres[assign[i:j]] + noise, 0.0)
    return x


def cluster_centres_cuda(n_centres, d=2048, seed=1234, device="cuda"):
    """The centres of `clustered_rows_cuda` ([n_centres, d] float32), a pure function of (n_centres, d, seed)."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return torch.clamp_min(torch.randn((n_centres, d), generator=g, device=device) * 0.5 + 0.35, 0.0)


def clustered_rows_cuda(row0, row1, centres, seed=1234, device="cuda", block=16384):
    """Rows [row0, row1) of a virtual clustered matrix of the same law as `clustered_features_cuda`, generated
    block-wise with one counter-based seed per 16 384-row block: any rank can generate exactly its shard of a large
    database (BASELINE config 4: 1 000 000 x 2048 over G GPUs) and the union over ranks is the same matrix a single
    process generates. Queries drawn around the SAME centres (another seed) have real neighbours in that database."""
    d = centres.shape[1]
    x = torch.empty((row1 - row0, d), dtype=torch.float32, device=device)
    g = torch.Generator(device=device)
    b = row0 // block
    while b * block < row1:
        lo, hi = max(row0, b * block), min(row1, (b + 1) * block)
        g.manual_seed((seed * 1000003 + b * 7919 + 12345) & 0x7FFFFFFFFFFFFFFF)
        assign = torch.randint(0, centres.shape[0], (block,), generator=g, device=device)
        noise = torch.randn((block, d), generator=g, device=device)
        o0, o1 = lo - b * block, hi - b * block
        x[lo - row0:hi - row0] = torch.clamp_min(centres[assign[o0:o1]] + 0.15 * noise[o0:o1], 0.0)
        b += 1
    return x


def cx_examples_np(n_examples, n_images, knn_idx, n_questions, seed=4321, n_answers=2000, knn_size=24):
    """Index-form VQA-CX examples (what getDataFromBatch, counterexamples.py:519-547, derives from the
    example dicts): image row, its 24 neighbour rows, question id, answer id, ground-truth candidate.

    `knn_idx` is an int array [n_images, >= knn_size + 1] with column 0 = self (stage-1 output).
    comp index law: P(idx < 1) ~ 0.115, P(idx < 5) ~ 0.445 (DistanceBaseline recalls, README.md:39).
    """
    rng = np.random.Generator(np.random.Philox(seed))
    img = rng.integers(0, n_images, size=n_examples)
    nbrs = knn_idx[img, 1:knn_size + 1]
    # truncated geometric-ish law over 0..23 fitted to the two README quantiles
    p = 0.115 * (1.0 - 0.115) ** np.arange(knn_size)
    p = p / p.sum()
    comp = rng.choice(knn_size, size=n_examples, p=p)
    zipf = 1.0 / np.arange(1, n_answers + 1)
    zipf /= zipf.sum()
    aid = rng.choice(n_answers, size=n_examples, p=zipf)
    qid = rng.integers(0, n_questions, size=n_examples)
    img_idx = np.concatenate([img[:, None], nbrs], axis=1).astype(np.int32)
    return {
        "img_idx": img_idx,
        "q_idx": qid.astype(np.int64),
        "answer_aid": aid.astype(np.int64),
        "comp_idx": comp.astype(np.int64),
    }
