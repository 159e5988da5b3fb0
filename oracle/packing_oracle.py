"""CPU oracle for the token-packing half of the CLSM path -- TEST INFRASTRUCTURE ONLY.

Restates what the reference does between a JSON line
{'q', 'doc_corr', 'doc_incorr': [J]} and the padded integer batch it yields:
  clean ......... utils/data_utils.py:13-19   (regex -> space, strip, lower)
  split/cutoff .. utils/data_utils.py:21-55   (word / ngram / char)
  right-pad ..... utils/data_utils.py:57-68   ('<PAD/>')
  id lookup ..... utils/data_utils.py:83-92   (dict.get -> None on OOV)
  batch format .. dssm_cnn_v2/data_helpers.py:128-197

Pinned: tests/golden/packing_golden.json was produced by importing the
reference's own utils/data_utils.py under Python 3 (tests/golden/make_golden.py)
and this restatement is checked against it in tests/test_oracle_packing.py.
"""
from __future__ import annotations

import json
import re

import numpy as np

PAD = "<PAD/>"
_KEEP = re.compile(r"[^\wäöüß€#\n.$]", re.UNICODE)   # data_utils.py:11


def clean(text):
    return _KEEP.sub(" ", text).strip().lower()


def features(text, mode, cutoff=None, ngram=3):
    """One string -> list of features, truncated to cutoff (data_utils.py:21-55)."""
    t = clean(text)
    if mode == "word":
        feats = t.split(" ")
    elif mode == "ngram":
        feats = [t[i:i + ngram] for i in range(len(t) - ngram + 1)]
    elif mode == "char":
        feats = list(t)
    else:
        raise ValueError(mode)
    return feats[:cutoff]


def pack_row(text, vocab, mode, length, ngram=3):
    f = features(text, mode, length, ngram)
    f = f + [PAD] * (length - len(f))
    return [vocab.get(w) for w in f]


def pack_batch(lines, vocab, mode, Lq, Ld, J, ngram=3):
    """JSON lines -> ([q (B,Lq), pos (B,Ld), neg_0..neg_{J-1} (B,Ld)], ones(B)).
    Lines whose doc_incorr length != J are skipped (data_helpers.py:163)."""
    q, pos, negs = [], [], [[] for _ in range(J)]
    for line in lines:
        d = json.loads(line)
        if len(d["doc_incorr"]) != J:
            continue
        q.append(pack_row(d["q"], vocab, mode, Lq, ngram))
        pos.append(pack_row(d["doc_corr"], vocab, mode, Ld, ngram))
        for i in range(J):
            negs[i].append(pack_row(d["doc_incorr"][i], vocab, mode, Ld, ngram))
    xs = [np.array(q), np.array(pos)] + [np.array(n) for n in negs]
    return xs, np.ones(len(q))
