"""TEST INFRASTRUCTURE (oracle): plain-Python restatement of the example builder join of the reference's notebooks
(vqacx_valset_builder.ipynb cells 4, 10-13; vqacx_trainset_builder.ipynb cell 18). Only tests/ may import this.

Pinning: oracle/gen_golden_builder.py EXECUTES the notebooks' own cell code (extracted from the notebook JSON) on seeded
synthetic inputs and writes tests/golden/builder_{val,train}.npz; tests/test_builder.py checks this restatement (and the
CUDA join) against those fixtures.

The notebooks hold the logic inline (no function to import), so this restates it on in-memory structures:
  comp_pairs : list of (q1, q2) question-id pairs          (cell 4:  comp_q[q1] = q2; comp_q[q2] = q1)
  examples   : list of dicts with 'question_id', 'image'    (cell 10: q_id_to_example)
  knns       : {image: [24 images sorted by distance]}      (cell 6)
and keeps an example iff q1 and its complement q2 are both in the intersection of the two question-id sets and
the complement's image is among the 24 neighbours of the example's image (cell 13), recording
`comp.knn_index = knns[i1].index(i2)` and `knns = knns[i1]`."""


def build_examples(examples, comp_pairs, knns):
    comp_q = {}
    for q1, q2 in comp_pairs:                                   # cell 4
        comp_q[q1] = q2
        comp_q[q2] = q1
    qid_to_example = {ex["question_id"]: ex for ex in examples}  # cell 10
    intersection = set(comp_q.keys()) & set(qid_to_example.keys())   # cell 12
    out = []
    for ex in examples:                                         # cell 13
        q1, i1 = ex["question_id"], ex["image"]
        if q1 in intersection:
            q2 = comp_q[q1]
            if q2 in intersection:
                i2 = qid_to_example[q2]["image"]
                if i2 in knns[i1]:
                    out.append({"question_id": q1, "image": i1, "knns": list(knns[i1]),
                                "comp_knn_index": knns[i1].index(i2), "comp_question_id": q2})
    return out


def to_arrays(examples, comp_pairs):
    """The host half of the join (dictionary work on ids): per-question image row and complementary-question index
    (-1 where q1 or its complement falls outside the intersection), in `examples` order."""
    import numpy as np

    pos = {ex["question_id"]: i for i, ex in enumerate(examples)}
    comp_q = {}
    for q1, q2 in comp_pairs:
        comp_q[q1] = q2
        comp_q[q2] = q1
    q_img = np.array([ex["image"] for ex in examples], dtype=np.int32)
    q_comp = np.full(len(examples), -1, dtype=np.int32)
    for i, ex in enumerate(examples):
        q2 = comp_q.get(ex["question_id"])
        if q2 is not None and q2 in pos:
            q_comp[i] = pos[q2]
    return q_img, q_comp
