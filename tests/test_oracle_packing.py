"""Packing parity (CPU): golden vectors produced by the REFERENCE's utils/data_utils.py
(tests/golden/make_golden.py) against (a) the oracle restatement and (b) the product's host
code -- DataUtils, DataUtils.pack and the DataHelpers generator."""
import json
import os

import numpy as np
import pytest

from dnn_page_vectors_b200 import Configuration, DataHelpers, DataUtils
from oracle import packing_oracle as PO

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "packing_golden.json"), encoding="utf-8") as f:
    GOLD = json.load(f)


def test_golden_has_the_reference_test_py_strings():
    texts = {r["text"] for r in GOLD["splits"]}
    assert "statue of liberty" in texts and "new york" in texts     # dssm_cnn_v2/test.py:8-22


@pytest.mark.parametrize("rec", GOLD["splits"], ids=lambda r: "%s-%s-%d" % (r["mode"], r["cutoff"], len(r["text"])))
def test_splits_match_reference(rec):
    du = DataUtils()
    assert PO.clean(rec["text"]) == rec["clean"]
    assert du.clean_str(rec["text"]) == rec["clean"]
    assert [PO.features(rec["text"], rec["mode"], rec["cutoff"], rec["ngram"])] == rec["single"]
    got = du.get_text_feature_splits(rec["text"], cutoff=rec["cutoff"], mode=rec["mode"], ngram=rec["ngram"])
    assert got == rec["single"]
    got = du.get_text_feature_splits([rec["text"], rec["text"] + " z"], cutoff=rec["cutoff"], mode=rec["mode"],
                                     ngram=rec["ngram"])
    assert got == rec["listed"]
    if "padded" in rec:
        assert du.pad_sentences(rec["single"], rec["cutoff"]) == rec["padded"]


@pytest.mark.parametrize("gb", GOLD["batches"], ids=lambda g: g["mode"])
def test_batches_match_reference(gb, tmp_path):
    xs, y = PO.pack_batch(gb["lines"], gb["vocab"], gb["mode"], gb["Lq"], gb["Ld"], gb["J"])
    np.testing.assert_array_equal(xs[0], np.array(gb["q"]))
    np.testing.assert_array_equal(xs[1], np.array(gb["pos"]))
    for a, b in zip(xs[2:], gb["negs"]):
        np.testing.assert_array_equal(a, np.array(b))
    np.testing.assert_array_equal(y, np.array(gb["y"]))

    # product generator on the same lines
    conf = Configuration(experiment_root_directory=str(tmp_path), feature_level=gb["mode"],
                         num_negative_examples=gb["J"], query_length=gb["Lq"], document_length=gb["Ld"])
    os.makedirs(conf.data_dir)
    for path in conf.input_file_list:
        with open(path, "w", encoding="utf-8") as f:
            f.write("\n".join(gb["lines"]) + "\n")
    dh = DataHelpers(conf)
    gen = dh.load_data_generator(gb["vocab"], mode="training", batch_size=128, nb_epochs=1)
    x, yy = next(gen)
    assert len(x) == 2 + gb["J"] and x[0].dtype == np.int32
    np.testing.assert_array_equal(x[0], np.array(gb["q"]))
    np.testing.assert_array_equal(x[1], np.array(gb["pos"]))
    for a, b in zip(x[2:], gb["negs"]):
        np.testing.assert_array_equal(a, np.array(b))
    np.testing.assert_array_equal(yy, np.ones(len(gb["q"])))
    # the generator makes nb_epochs + 2 passes (data_helpers.py:148)
    assert sum(1 for _ in gen) == 2


def test_generator_requires_mode():
    with pytest.raises(Exception):
        next(DataHelpers(Configuration(feature_level="word")).load_data_generator({}, mode=None))


def test_vocab_and_embedding_init(tmp_path):
    gb = GOLD["batches"][0]                                   # word level
    conf = Configuration(experiment_root_directory=str(tmp_path), feature_level="word",
                         num_negative_examples=gb["J"], query_length=gb["Lq"], document_length=gb["Ld"])
    os.makedirs(conf.data_dir)
    os.makedirs(conf.vectors_directory)
    for path in conf.input_file_list:
        with open(path, "w", encoding="utf-8") as f:
            f.write("\n".join(gb["lines"]) + "\n")
    D = 4
    with open(conf.word_vectors_file, "w", encoding="utf-8") as f:
        f.write("3 4\n")
        f.write("liberty 0.1 0.2 0.3 0.4\n")
        f.write("wetter 1 2 3 4 5\n")                          # longer line: first D floats
        f.write("notinvocab 9 9 9 9\n")
    dh = DataHelpers(conf)
    np.random.seed(1337)
    W, vocab = dh.get_vocab_index_embedding_weights(D, False, False, False)
    assert vocab == gb["vocab"]                               # sorted ids, PAD/UNK/' ' included
    assert W.shape == (len(vocab) + 1, D) and W.dtype == np.float64
    np.testing.assert_allclose(W[vocab["liberty"]], [0.1, 0.2, 0.3, 0.4])
    np.testing.assert_allclose(W[vocab["wetter"]], [1, 2, 3, 4])
    assert (W[-1] == 0).all()                                 # spare row (data_helpers.py:56)
    others = np.delete(W[:-1], [vocab["liberty"], vocab["wetter"]], axis=0)
    assert (np.abs(others) <= 0.25).all() and (others != 0).any()
    # pickles written and reloadable
    W2, vocab2 = dh.get_vocab_index_embedding_weights(D, False, True, True)
    np.testing.assert_array_equal(W, W2)
    assert vocab2 == vocab


def test_char_level_returns_no_weights(tmp_path):
    gb = GOLD["batches"][1]
    conf = Configuration(experiment_root_directory=str(tmp_path), feature_level="char",
                         num_negative_examples=gb["J"])
    os.makedirs(conf.data_dir)
    for path in conf.input_file_list:
        with open(path, "w", encoding="utf-8") as f:
            f.write("\n".join(gb["lines"]) + "\n")
    W, vocab = DataHelpers(conf).get_vocab_index_embedding_weights(100, False)
    assert W == [] and vocab == gb["vocab"]                   # data_helpers.py:203,218


def test_configuration_literals():
    c = Configuration()
    assert (c.feature_level, c.query_length, c.document_length, c.num_negative_examples) == ("char", 250, 5000, 3)
    w = Configuration(feature_level="word")
    assert (w.query_length, w.document_length) == (20, 975)
    n = Configuration(feature_level="ngram")
    assert (n.query_length, n.document_length, n.default_ngram) == (45, 2000, 3)
    assert c.word_vectors_file.endswith("fast_model_ns.vec")
