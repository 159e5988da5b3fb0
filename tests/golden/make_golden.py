"""Regenerates tests/golden/packing_golden.json by importing the REFERENCE's own
utils/data_utils.py (it runs unmodified under Python 3).  Run only in the build
container, where /root/reference exists:

    python tests/golden/make_golden.py

The strings include the ones the reference's dssm_cnn_v2/test.py:8-22 exercises
("statue of liberty", "new york").  Vocabulary ids are assigned in sorted
order so the fixture is reproducible (the reference enumerates a Python set,
dssm_cnn_v2/data_helpers.py:116-118, which is not stable across processes).
The batch section replays the body of DataHelpers.load_data_generator
(dssm_cnn_v2/data_helpers.py:160-197) with the reference's DataUtils calls;
data_helpers.py itself is Python-2 syntax and cannot be imported.
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, "/root/reference")
from utils.data_utils import DataUtils  # noqa: E402  (the reference's module)

du = DataUtils()
HERE = os.path.dirname(os.path.abspath(__file__))

STRINGS = [
    "statue of liberty",
    "new york",
    "Statue  of   Liberty!!",
    "  München: größte Straße – 5€ #1 (top) $9.99 ",
    "",
    "a",
    "tab\tand\nnewline",
    "UPPER lower MiXeD 123_abc",
    "x" * 40,
    "the quick brown fox jumps over the lazy dog " * 3,
]
CASES = [("word", 5, 3), ("word", 20, 3), ("char", 5, 3), ("char", 30, 3),
         ("ngram", 5, 3), ("ngram", 12, 3), ("ngram", 8, 2), ("word", None, 3)]

out = {"splits": [], "batches": []}
for mode, cutoff, n in CASES:
    for s in STRINGS:
        single = du.get_text_feature_splits(s, cutoff=cutoff, mode=mode, ngram=n)
        listed = du.get_text_feature_splits([s, s + " z"], cutoff=cutoff, mode=mode, ngram=n)
        rec = {"mode": mode, "cutoff": cutoff, "ngram": n, "text": s,
               "single": single, "listed": listed, "clean": du.clean_str(s)}
        if cutoff is not None:
            rec["padded"] = du.pad_sentences(single, cutoff)
        out["splits"].append(rec)

LINES = [
    {"q": "statue of liberty", "doc_corr": "Statue of Liberty National Monument, New York",
     "doc_incorr": ["new york pizza", "liberty mutual insurance", "statue park budapest"]},
    {"q": "münchen wetter", "doc_corr": "Wetter München: 16-Tage Trend",
     "doc_incorr": ["berlin wetter", "münchen hbf", "wetter.com"]},
    {"q": "skipped line", "doc_corr": "has too few negatives", "doc_incorr": ["only one"]},
    {"q": "cheap flights", "doc_corr": "Cheap Flights, Airline Tickets & Airfare Deals",
     "doc_incorr": ["cheap hotels", "flight simulator", "tickets for concerts"]},
]
J = 3
for mode, Lq, Ld in [("word", 4, 7), ("char", 10, 25), ("ngram", 6, 15)]:
    vocab_set = {"<PAD/>", "<UNK/>", " "}
    for d in LINES:
        if len(d["doc_incorr"]) == J:
            s_list = [d["q"], d["doc_corr"]] + d["doc_incorr"]
            vocab_set.update(du.build_vocab(du.get_text_feature_splits(s_list, mode=mode)))
    vocab = {w: i for i, w in enumerate(sorted(vocab_set))}
    q = np.empty((0, 0), dtype=np.int32)
    pos = np.empty((0, 0), dtype=np.int32)
    negs = [np.empty((0, 0), dtype=np.int32) for _ in range(J)]
    for d in LINES:
        if len(d["doc_incorr"]) != J:
            continue
        xa = du.build_input_data(du.pad_sentences(
            du.get_text_feature_splits(d["q"], cutoff=Lq, mode=mode), Lq), vocab)
        q = xa if q.shape[0] == 0 else np.vstack((q, xa))
        xa = du.build_input_data(du.pad_sentences(
            du.get_text_feature_splits(d["doc_corr"], cutoff=Ld, mode=mode), Ld), vocab)
        pos = xa if pos.shape[0] == 0 else np.vstack((pos, xa))
        for i in range(J):
            xa = du.build_input_data(du.pad_sentences(
                du.get_text_feature_splits(d["doc_incorr"][i], cutoff=Ld, mode=mode), Ld), vocab)
            negs[i] = xa if negs[i].shape[0] == 0 else np.vstack((negs[i], xa))
    out["batches"].append({
        "mode": mode, "Lq": Lq, "Ld": Ld, "J": J,
        "lines": [json.dumps(d, ensure_ascii=False) for d in LINES],
        "vocab": vocab,
        "q": q.tolist(), "pos": pos.tolist(), "negs": [n.tolist() for n in negs],
        "y": np.ones(len(q)).tolist(),
    })

with open(os.path.join(HERE, "packing_golden.json"), "w", encoding="utf-8") as f:
    json.dump(out, f, ensure_ascii=False, indent=0, sort_keys=True)
print("wrote", len(out["splits"]), "split cases,", len(out["batches"]), "batches")
