"""Host-side text -> padded token-id packing.

Mirrors the reference's ``DataUtils`` (utils/data_utils.py:8-92) name for name --
``clean_str``, ``get_text_feature_splits``, ``pad_sentences``, ``build_vocab``,
``build_input_data`` -- so callers of the reference's data path can switch
without edits.  ``pack`` is the added fast path the trainer uses: it goes from
raw strings straight to one contiguous int32 (n, length) array ready for a
pinned-memory upload, instead of one (1, L) array per row stacked with
``np.vstack`` (dssm_cnn_v2/data_helpers.py:171-192).
"""
from __future__ import annotations

import re
from collections import Counter

import numpy as np

PAD_WORD = "<PAD/>"
UNK_WORD = "<UNK/>"


class DataUtils(object):
    # utils/data_utils.py:11 -- everything that is not a word character or one
    # of a few kept symbols becomes a single space.
    _drop = re.compile(r"[^\wäöüß€#\n.$]", re.UNICODE)

    def __init__(self):
        self.rgx = self._drop

    # -- utils/data_utils.py:13-19
    def clean_str(self, string):
        return self._drop.sub(" ", string).strip().lower()

    def _split_one(self, text, mode, ngram):
        cleaned = self.clean_str(text)
        if mode == "word":
            return cleaned.split(" ")
        if mode == "ngram":
            return [cleaned[i:i + ngram] for i in range(max(len(cleaned) - ngram + 1, 0))]
        if mode == "char":
            return list(cleaned)
        raise ValueError("mode must be 'word', 'ngram' or 'char', got %r" % (mode,))

    # -- utils/data_utils.py:21-55
    def get_text_feature_splits(self, training_data, cutoff=None, mode="word", ngram=3):
        items = training_data if isinstance(training_data, list) else [training_data]
        return [self._split_one(t, mode, ngram)[:cutoff] for t in items]

    # -- utils/data_utils.py:57-68
    def pad_sentences(self, sentences, sequence_length, padding_word=PAD_WORD):
        return [list(s) + [padding_word] * (sequence_length - len(s)) for s in sentences]

    # -- utils/data_utils.py:70-81 (returns the inverse vocabulary, most common first)
    def build_vocab(self, sentences):
        counts = Counter()
        for s in sentences:
            counts.update(s)
        return [w for w, _ in counts.most_common()]

    # -- utils/data_utils.py:83-92 (OOV -> None, i.e. an object array, as the reference)
    def build_input_data(self, sentences, vocabulary, return_array=True):
        ids = [[vocabulary.get(w) for w in s] for s in sentences]
        return np.array(ids) if return_array else ids

    # -- added: one call from strings to a contiguous int32 block
    def pack(self, texts, vocabulary, length, mode="word", ngram=3, out=None):
        """texts: list[str] -> int32 (len(texts), length), truncated to ``length``
        then right-padded with the id of '<PAD/>' (same result as
        build_input_data(pad_sentences(get_text_feature_splits(t, length, mode), length))).
        Raises KeyError on an out-of-vocabulary feature (the reference would
        produce an object array holding None, utils/data_utils.py:87)."""
        n = len(texts)
        if out is None:
            out = np.empty((n, length), dtype=np.int32)
        out[...] = vocabulary[PAD_WORD]
        for r, t in enumerate(texts):
            feats = self._split_one(t, mode, ngram)[:length]
            if feats:
                out[r, :len(feats)] = [vocabulary[w] for w in feats]
        return out
