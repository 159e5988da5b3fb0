"""Pre-tokenised corpus: DataHelpers.pretokenize + load_packed_generator must yield exactly what
load_data_generator yields -- itself pinned to the reference's utils/data_utils.py by tests/golden -- batch for
batch, including batches thinned or emptied by records with the wrong number of negatives."""
import json
import os

import numpy as np
import pytest

from dnn_page_vectors_b200 import Configuration, DataHelpers

WORDS = ["alpha", "beta", "gamma", "delta", "epsilon", "zeta", "eta", "theta", "iota", "kappa", "B200", "x-ray", "don't", "(paren)"]


def make_corpus(tmp_path, n, J, level, seed=0, bad_every=7, bad_run=None):
    rng = np.random.RandomState(seed)
    conf = Configuration(experiment_root_directory=str(tmp_path), feature_level=level,
                         num_negative_examples=J, query_length=6, document_length=17)
    os.makedirs(conf.data_dir, exist_ok=True)

    def text(k):
        return " ".join(WORDS[i] for i in rng.randint(0, len(WORDS), size=k))
    lines = []
    for r in range(n):
        nneg = J
        if bad_every and r % bad_every == 3:
            nneg = J - 1 if J > 0 else 1
        if bad_run and bad_run[0] <= r < bad_run[1]:
            nneg = J + 1
        lines.append(json.dumps({"q": text(rng.randint(1, 9)), "doc_corr": text(rng.randint(1, 30)),
                                 "doc_incorr": [text(rng.randint(1, 30)) for _ in range(nneg)]}))
    for path in conf.input_file_list:
        with open(path, "w", encoding="utf-8") as f:
            f.write("\n".join(lines) + "\n")
    return conf


@pytest.mark.parametrize("level", ["word", "char", "ngram"])
@pytest.mark.parametrize("n,bs,J,bad_run", [(50, 8, 3, None), (33, 16, 2, (16, 32)), (5, 128, 4, None), (64, 16, 1, None)])
def test_packed_generator_equals_json_generator(tmp_path, level, n, bs, J, bad_run):
    conf = make_corpus(tmp_path, n, J, level, bad_run=bad_run)
    dh = DataHelpers(conf)
    _, vocab = dh.generate_vocabulary_set()
    packed = dh.pretokenize(vocab, mode="training", chunk_lines=7, workers=3 if level == "word" else 1)
    meta = json.load(open(os.path.join(packed, "meta.json")))
    assert meta["lines"] == n and meta["J"] == J
    ref = list(dh.load_data_generator(vocab, mode="training", batch_size=bs, nb_epochs=1))
    got = list(dh.load_packed_generator(packed, batch_size=bs, nb_epochs=1))
    assert len(ref) == len(got) == 3 * ((n + bs - 1) // bs)          # nb_epochs + 2 passes
    for (xr, yr), (xg, yg) in zip(ref, got):
        assert len(xr) == len(xg) == 2 + J
        for a, b in zip(xr, xg):
            assert b.dtype == np.int32 and a.shape == b.shape
            np.testing.assert_array_equal(a, b)
        np.testing.assert_array_equal(yr, yg)
    if bad_run:          # a whole batch of skipped records comes out empty in both
        assert any(len(y) == 0 for _, y in got)


def test_pretokenize_requires_mode(tmp_path):
    conf = make_corpus(tmp_path, 4, 2, "word")
    dh = DataHelpers(conf)
    _, vocab = dh.generate_vocabulary_set()
    with pytest.raises(Exception):
        dh.pretokenize(vocab)
    # validation file goes to its own directory
    p = dh.pretokenize(vocab, mode="validation")
    assert p.endswith(".packed") and os.path.exists(os.path.join(p, "q.npy"))
