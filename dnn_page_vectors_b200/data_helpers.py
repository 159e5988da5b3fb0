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


_PACK = {}


def _pack_init(vocab_index_dict, spec):
    _PACK["vocab"], _PACK["spec"], _PACK["du"] = vocab_index_dict, spec, DataUtils()


def _pack_chunk(rows):
    """rows: JSON lines -> (number of lines, indices of the kept ones, [q, pos, neg_0..] int32 blocks).  Runs in the
    pretokenize() worker processes (and in-process when workers == 1)."""
    J, Lq, Ld, level, ng = _PACK["spec"]
    vocab, du = _PACK["vocab"], _PACK["du"]
    recs, kept = [], []
    for k, row in enumerate(rows):
        d = json.loads(row)
        if len(d["doc_incorr"]) == J:
            recs.append(d)
            kept.append(k)
    arrays = [du.pack([d["q"] for d in recs], vocab, Lq, level, ng),
              du.pack([d["doc_corr"] for d in recs], vocab, Ld, level, ng)]
    for i in range(J):
        arrays.append(du.pack([d["doc_incorr"][i] for d in recs], vocab, Ld, level, ng))
    return len(rows), kept, arrays


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

    # ------------------------------------------------------------------ added
    def pretokenize(self, vocab_index_dict, mode=None, out_dir=None, chunk_lines=4096, workers=1):
        """One-off conversion of the JSON-lines file of `mode` into int32 arrays on disk, so that epochs after the
        first never touch JSON, regexes or dict lookups again (the reference's generator redoes all three for
        every line of every epoch, data_helpers.py:148-196).

        Writes `<out_dir>/{q,pos,neg0..negJ-1}.npy` (one contiguous (n, L) int32 matrix per model input, rows =
        the records load_data_generator would keep), `line.npy` (int64 source line number of each kept record:
        batch boundaries of the original generator are reproduced from it) and `meta.json`.  Returns out_dir.
        workers > 1 tokenises the chunks in that many processes (the result does not depend on it)."""
        if mode not in ("training", "validation"):
            raise Exception("Please provide mode as either 'training' or 'validation'")
        conf = self.conf
        path = conf.model_training_data if mode == "training" else conf.model_validation_data
        out_dir = out_dir or (path + ".packed")
        os.makedirs(out_dir, exist_ok=True)
        J, Lq, Ld = conf.num_negative_examples, conf.query_length, conf.document_length
        level, ng = conf.feature_level, conf.default_ngram
        names = ["q", "pos"] + ["neg%d" % i for i in range(J)]
        parts = {name: [] for name in names}
        lines, n_lines = [], 0

        def chunks():
            with open(path, "r", encoding="utf-8") as fin:
                while True:
                    rows = list(islice(fin, chunk_lines))
                    if not rows:
                        return
                    yield rows

        spec = (J, Lq, Ld, level, ng)
        if workers > 1:
            import multiprocessing as mp
            # fork, not spawn: spawn re-imports the caller's __main__, which a library cannot ask of its callers.  The
            # workers run only json + NumPy code on their chunk, nothing that needs the parent's threads or CUDA state.
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore", DeprecationWarning)
                pool = mp.get_context("fork").Pool(workers, initializer=_pack_init, initargs=(vocab_index_dict, spec))
            results = pool.imap(_pack_chunk, chunks())          # ordered: chunk i's arrays come out i-th
        else:
            pool = None
            _pack_init(vocab_index_dict, spec)
            results = map(_pack_chunk, chunks())
        try:
            for n_rows, kept, arrays in results:
                lines.extend(n_lines + k for k in kept)
                n_lines += n_rows
                for name, arr in zip(names, arrays):
                    parts[name].append(arr)
        finally:
            if pool is not None:
                pool.close()
                pool.join()
        for name, blocks in parts.items():
            width = Lq if name == "q" else Ld
            arr = np.concatenate(blocks, axis=0) if blocks else np.empty((0, width), dtype=np.int32)
            np.save(os.path.join(out_dir, name + ".npy"), np.ascontiguousarray(arr, dtype=np.int32))
        np.save(os.path.join(out_dir, "line.npy"), np.asarray(lines, dtype=np.int64))
        with open(os.path.join(out_dir, "meta.json"), "w") as f:
            json.dump({"source": os.path.abspath(path), "records": len(lines), "lines": n_lines, "J": J, "Lq": Lq, "Ld": Ld,
                       "feature_level": level, "ngram": ng, "vocab_size": len(vocab_index_dict)}, f)
        return out_dir

    def load_packed_generator(self, packed_dir, batch_size=128, nb_epochs=1, pinned=False):
        """Same yields as load_data_generator(mode, batch_size, nb_epochs) -- `([q, pos, neg_0..], ones)`, the
        same batch boundaries (batch b = the kept records among source lines [b*batch_size, (b+1)*batch_size)),
        the same nb_epochs + 2 passes -- read from the arrays pretokenize() wrote (memory-mapped).
        pinned=True yields contiguous int32 torch tensors in pinned host memory instead of NumPy arrays
        (two rotating buffer sets): CLSM.train_on_batch copies those to the device without repacking."""
        with open(os.path.join(packed_dir, "meta.json")) as f:
            meta = json.load(f)
        names = ["q", "pos"] + ["neg%d" % i for i in range(meta["J"])]
        arrs = [np.load(os.path.join(packed_dir, n + ".npy"), mmap_mode="r") for n in names]
        line = np.load(os.path.join(packed_dir, "line.npy"))
        n_batches = (meta["lines"] + batch_size - 1) // batch_size
        cuts = np.searchsorted(line, np.arange(n_batches + 1, dtype=np.int64) * batch_size)
        bufs, turn = None, 0
        if pinned:
            import torch
            bufs = [[torch.empty((batch_size, a.shape[1]), dtype=torch.int32).pin_memory() for a in arrs] for _ in range(2)]
        for _ in range(nb_epochs + 2):               # data_helpers.py:148
            for b in range(n_batches):
                lo, hi = int(cuts[b]), int(cuts[b + 1])
                if pinned:
                    out = []
                    for a, t in zip(arrs, bufs[turn]):
                        t.numpy()[:hi - lo] = a[lo:hi]
                        out.append(t[:hi - lo])
                    turn ^= 1
                    yield out, np.ones(hi - lo)
                else:
                    yield [np.array(a[lo:hi]) for a in arrs], np.ones(hi - lo)

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
