"""Vocabulary, word-vector initialisation and batch generator of the CLSM path.

Same public names and return formats as the reference's ``DataHelpers``
(dssm_cnn_v2/data_helpers.py):

  get_vocab_index_embedding_weights(...) -> (ndarray (V+1, D) | [], dict)   :202-222
  generate_vocabulary_set(masking)       -> (set, dict)                      :77-124
  load_word_embeddings_compact(...)      -> ndarray (V+1, D) float64         :17-74
  load_data_generator(...)               -> yields ([q, pos, neg_0..], y)    :128-197

Differences, all deliberate: ids are assigned over the SORTED vocabulary (the
reference enumerates a Python set, :116-118, which is not reproducible across
processes); table rows follow the ids of ``vocab_index_dict`` (the reference
fills rows in ``list(vocab_set)`` order, :19,58, which only coincides with the
ids inside one process); batches are built as one contiguous int32 block per
input instead of row-wise ``np.vstack`` (:171-192).
"""
from __future__ import annotations

import json
import os
from itertools import islice

import joblib
import numpy as np

from .config import Configuration
from .data_utils import PAD_WORD, UNK_WORD, DataUtils


def _mask_tag(masking):
    return "masked" if masking else "non_masked"


class DataHelpers(object):
    def __init__(self, conf=None):
        self.conf = conf if conf is not None else Configuration()
        self.du = DataUtils()

    # ------------------------------------------------------------------ :77-124
    def generate_vocabulary_set(self, masking=False):
        conf = self.conf
        vocab_set = {PAD_WORD, UNK_WORD, " "}
        for path in conf.input_file_list:
            with open(path, encoding="utf-8") as fo:
                for line in fo:
                    data = json.loads(line)
                    if len(data["doc_incorr"]) != conf.num_negative_examples:
                        continue
                    texts = [data["q"], data["doc_corr"]] + list(data["doc_incorr"])
                    for feats in self.du.get_text_feature_splits(
                            texts, mode=conf.feature_level, ngram=conf.default_ngram):
                        vocab_set.update(feats)
        first = 1 if masking else 0
        vocab_index_dict = {w: i for i, w in enumerate(sorted(vocab_set), start=first)}
        if conf.create_data_dump:
            os.makedirs(conf.pickle_files_dir, exist_ok=True)
            tag = _mask_tag(masking)
            joblib.dump(vocab_set, conf.vocab_set_file.format(tag))
            joblib.dump(vocab_index_dict, conf.vocab_index_file.format(tag))
        return vocab_set, vocab_index_dict

    # ------------------------------------------------------------------ :17-74
    def load_word_embeddings_compact(self, embedding_dim, vocab_set, masking=False,
                                     use_pickled=True, vocab_index_dict=None):
        conf = self.conf
        weights_path = conf.embedding_weights_file_tpl.format(_mask_tag(masking))
        if use_pickled:
            return joblib.load(weights_path)

        first = 1 if masking else 0
        if vocab_index_dict is None:
            vocab_index_dict = {w: i for i, w in enumerate(sorted(vocab_set), start=first)}
        wanted = set(vocab_index_dict)
        found = {}
        with open(conf.word_vectors_file, "r", encoding="utf-8", errors="replace") as f:
            for line in f:
                parts = line.strip().split()
                if len(parts) < embedding_dim:      # header / short line (:41)
                    continue
                if parts[0] in wanted:
                    try:
                        found[parts[0]] = np.asarray(
                            [float(x) for x in parts[1:embedding_dim + 1]])
                    except ValueError:
                        pass                         # the reference prints and goes on (:49-50)
        n_rows = len(vocab_index_dict) + 1           # +1: spare zero row (:56)
        weights = np.zeros((n_rows, embedding_dim))
        for word, i in sorted(vocab_index_dict.items(), key=lambda kv: kv[1]):
            vec = found.get(word)
            if vec is not None and vec.shape[0] == embedding_dim:
                weights[i, :] = vec
            else:
                weights[i, :] = np.random.uniform(-0.25, 0.25, embedding_dim)   # :62
        if conf.create_data_dump:
            os.makedirs(conf.pickle_files_dir, exist_ok=True)
            joblib.dump(weights, weights_path)
        return weights

    # ------------------------------------------------------------------ :128-197
    def load_data_generator(self, vocab_index_dict, mode=None, batch_size=128, nb_epochs=1):
        if mode is None:
            raise Exception("Please provide mode as either 'training' or 'validation'")
        conf = self.conf
        if mode == "training":
            path = conf.model_training_data
        elif mode == "validation":
            path = conf.model_validation_data
        else:
            path = ""
        J = conf.num_negative_examples
        level, ng = conf.feature_level, conf.default_ngram
        for _ in range(nb_epochs + 2):               # :148
            with open(path, "r", encoding="utf-8") as fin:
                while True:
                    rows = list(islice(fin, batch_size))
                    if not rows:
                        break
                    recs = [d for d in map(json.loads, rows) if len(d["doc_incorr"]) == J]
                    q = self.du.pack([d["q"] for d in recs], vocab_index_dict,
                                     conf.query_length, level, ng)
                    pos = self.du.pack([d["doc_corr"] for d in recs], vocab_index_dict,
                                       conf.document_length, level, ng)
                    negs = [self.du.pack([d["doc_incorr"][i] for d in recs], vocab_index_dict,
                                         conf.document_length, level, ng) for i in range(J)]
                    yield [q, pos] + negs, np.ones(len(recs))

    # ------------------------------------------------------------------ :202-222
    def get_vocab_index_embedding_weights(self, embedding_dim, embedding_weights_masking,
                                          load_embeddings_pickled=False, load_vocab_pickled=False):
        conf = self.conf
        tag = _mask_tag(embedding_weights_masking)
        if load_vocab_pickled:
            vocab_index_dict = joblib.load(conf.vocab_index_file.format(tag))
            vocab_set = joblib.load(conf.vocab_set_file.format(tag))
        else:
            vocab_set, vocab_index_dict = self.generate_vocabulary_set(
                masking=embedding_weights_masking)
        embedding_weights = []
        if conf.feature_level == "word":             # :218
            embedding_weights = self.load_word_embeddings_compact(
                embedding_dim, vocab_set, masking=embedding_weights_masking,
                use_pickled=load_embeddings_pickled, vocab_index_dict=vocab_index_dict)
        return embedding_weights, vocab_index_dict
