"""Generates tests/golden/builder_{val,train}.npz by EXECUTING the reference notebooks' own cell code
(/root/reference/vqacx_valset_builder.ipynb, vqacx_trainset_builder.ipynb: the cells that build `comp_q`, `knns`,
`coco_name_to_num`, `intersection` and the filter loop with `knns[i1].index(i2)`), extracted from the notebook JSON and
exec'd on seeded synthetic `*_examples_list` / complementary pairs / neighbour lists.

Run in the build container:  python oracle/gen_golden_builder.py

The notebooks read three files; the only shims are the readers: `open` returns the path, `pickle.load` / `json.load`
dispatch on it and hand back the synthetic structures (json.load gives the neighbour dict with STRING keys, as a real
JSON file would). No line of the notebooks' logic is restated here. TEST INFRASTRUCTURE.
"""
import json
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
K1 = 25

# (notebook, name of the examples list, name of the filtered list, COCO split of the image names)
NOTEBOOKS = {
    "val": ("vqacx_valset_builder.ipynb", "val_examples_list", "val_examples_filtered", "val"),
    "train": ("vqacx_trainset_builder.ipynb", "train_examples_list", "train_examples_filtered", "train"),
}
# a cell is executed iff it contains one of these markers (the data-reading cells and the join itself); everything
# else in the notebooks is imports, display and printing
MARKERS = ("_examples_list = pickle.load", "comp_pairs = json.load", "knns = json.load", "def coco_name_to_num",
           "q_ids_alpha = []", "q_ids_beta_set =", "intersection = q_ids_alpha_set", "_examples_filtered = []")


def synthetic(seed, n_img, n_q, split):
    """Seeded inputs in the notebooks' formats. Images get sparse COCO-like numbers; the stage-1 style neighbour table
    `knn` [n_img, 25] (row = image position, column 0 usually self) is the device join's input, the notebook gets the
    24-neighbour lists derived from it (self removed, or the last entry when duplicates displaced it)."""
    rng = np.random.default_rng(seed)
    coco = np.sort(rng.choice(np.arange(9, 600_000), size=n_img, replace=False)).astype(np.int64)
    knn = np.empty((n_img, K1), dtype=np.int64)
    for i in range(n_img):
        knn[i] = np.concatenate([[i], rng.permutation(np.delete(np.arange(n_img), i))[:K1 - 1]])
    displaced = rng.choice(n_img, size=max(1, n_img // 50), replace=False)
    for i in displaced:
        knn[i] = rng.permutation(np.delete(np.arange(n_img), i))[:K1]
    swap = rng.choice(np.setdiff1d(np.arange(n_img), displaced), size=max(1, n_img // 50), replace=False)
    knn[swap, 0], knn[swap, 1] = knn[swap, 1].copy(), knn[swap, 0].copy()
    q_img = rng.integers(n_img, size=n_q)
    qids = 1000 + rng.permutation(4 * n_q)[:n_q].astype(np.int64) * 3        # sparse, unordered question ids
    order = rng.permutation(n_q)
    half = n_q // 2
    pairs = [(int(qids[a]), int(qids[b])) for a, b in zip(order[0:half:2], order[1:half:2])]
    pairs.append((int(qids[order[-1]]), 999_999_999))        # complement outside the example set
    pairs.append((888_888_888, int(qids[order[-2]])))        # first id outside the example set
    pos = {int(q): i for i, q in enumerate(qids)}
    for a, b in pairs[:len(pairs) // 2]:                     # make the complement's image a neighbour for half of them
        if a in pos and b in pos:
            q_img[pos[b]] = knn[q_img[pos[a]], 1 + rng.integers(K1 - 1)]
    name = "COCO_{}2014_{{:012d}}.jpg".format(split)
    examples = [{"question_id": int(qids[i]), "image_name": name.format(int(coco[q_img[i]]))} for i in range(n_q)]
    nn_lists = {}
    for i in range(n_img):
        row = knn[i].tolist()
        if i in row:
            row.remove(i)
        else:
            row = row[:-1]
        nn_lists[str(int(coco[i]))] = [int(coco[j]) for j in row]
    return coco, knn, q_img.astype(np.int32), qids, np.asarray(pairs, dtype=np.int64), examples, nn_lists


def run_notebook(nb_file, list_name, filtered_name, examples, pairs, nn_lists):
    with open(os.path.join(REF, nb_file)) as f:
        nb = json.load(f)
    cells = ["".join(c["source"]) for c in nb["cells"] if c["cell_type"] == "code"]
    run = [src for src in cells if any(m in src for m in MARKERS)]
    assert len(run) == 8, [s[:40] for s in run]

    def fake_load(path):
        if "complementary_pairs" in path:
            return [list(p) for p in pairs.tolist()]
        if "nn_images" in path:
            return nn_lists
        if path.endswith("set.pickle"):
            return examples
        raise AssertionError(path)

    ns = {"open": lambda path, mode="r": path, "print": lambda *a, **k: None,
          "pickle": types.SimpleNamespace(load=fake_load), "json": types.SimpleNamespace(load=fake_load)}
    for src in run:
        exec(compile(src, nb_file, "exec"), ns)
    return ns[filtered_name], ns["coco_name_to_num"]


def main():
    os.makedirs(OUT, exist_ok=True)
    for tag, (nb_file, list_name, filtered_name, split) in NOTEBOOKS.items():
        seed, n_img, n_q = (11, 300, 2000) if tag == "val" else (12, 120, 900)
        coco, knn, q_img, qids, pairs, examples, nn_lists = synthetic(seed, n_img, n_q, split)
        kept, name_to_num = run_notebook(nb_file, list_name, filtered_name, examples, pairs, nn_lists)
        assert len(kept) > 50, len(kept)
        np.savez_compressed(
            os.path.join(OUT, f"builder_{tag}.npz"),
            # inputs
            coco=coco, knn=knn, q_img=q_img, qids=qids, pairs=pairs,
            # what the notebook kept, in its order
            kept_qid=np.asarray([ex["question_id"] for ex in kept], dtype=np.int64),
            kept_comp_qid=np.asarray([ex["comp"]["question_id"] for ex in kept], dtype=np.int64),
            kept_knn_index=np.asarray([ex["comp"]["knn_index"] for ex in kept], dtype=np.int64),
            kept_image=np.asarray([name_to_num(ex["image_name"]) for ex in kept], dtype=np.int64),
            kept_knns=np.asarray([[name_to_num(n) for n in ex["knns"]] for ex in kept], dtype=np.int64))
        print(tag, "examples", n_q, "kept", len(kept))


if __name__ == "__main__":
    sys.dont_write_bytecode = True
    main()
