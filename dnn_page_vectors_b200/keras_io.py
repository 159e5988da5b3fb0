"""Checkpoint interchange with the reference's Keras-1 files.

The reference persists its model with Keras 1.x ``ModelCheckpoint('weights.{epoch:02d}.hdf5')``,
``model.save_weights('cnn_dssm_model_weights.h5')`` and ``model.save(...)``
(dssm_cnn_v2/cnn_dssm_tf.py:193,199-206).  Such a file holds, per top-level layer, a group with the layer's
weights; the two towers are the ``Sequential`` containers ``sequential_1`` (query) and ``sequential_2`` (document),
whose weights Keras stores flattened in creation order:

    embedding_N_W (V+1, D) | convolution1d_M_W, convolution1d_M_b per window | dense_N_W (|K|F, H) | dense_N_b (H,)

A Keras-1 ``Convolution1D`` kernel is 4-D: ``(k, 1, D, F)`` with the TensorFlow dimension ordering,
``(F, D, k, 1)`` with the Theano ordering of the first 1.0.x releases; under the Theano BACKEND the kernel is
applied flipped along the window axis.  Very early files name weights ``param_0 ...`` -- then only the
order identifies them.  All of that is handled here, in ONE place, for three containers:

  * ``.npz``  -- the dependency-free interchange format (flat keys ``<group>/<weight>`` + ``__layer_names__``);
  * ``.h5`` / ``.hdf5`` -- through h5py when it is importable (it is not part of this image: the functions then
    raise ImportError naming scripts/keras_h5_to_npz.py, which converts in both directions wherever h5py exists);
  * in-memory dicts ``{group: [(name, array), ...]}`` -- what both of the above read into.

``CLSM.save_weights`` / ``load_weights`` / ``ModelCheckpoint`` pick the container by file extension.
"""
from __future__ import annotations

import re
from collections import OrderedDict

import numpy as np

GROUPS = ("sequential_1", "sequential_2")          # query tower, document tower (cnn_dssm_tf.py:135-137)


# ----------------------------------------------------------------------------- model -> Keras-1 naming
def export_layers(weights, n_windows):
    """weights: CLSM.get_weights() ({'query': {...}, 'doc': {...}} in reference layouts) ->
    OrderedDict {group: [(keras_weight_name, float32 array)]} with Keras-1 names, numbering and 4-D conv kernels
    (k, 1, D, F) as the TensorFlow dimension ordering stores them."""
    out = OrderedDict()
    for t, (group, tag) in enumerate(zip(GROUPS, ("query", "doc"))):
        w = weights[tag]
        items = [("embedding_%d_W" % (t + 1), np.asarray(w["E"], dtype=np.float32))]
        for i in range(n_windows):
            n = t * n_windows + i + 1
            kern = np.asarray(w["convW%d" % i], dtype=np.float32)
            items.append(("convolution1d_%d_W" % n, kern[:, None, :, :]))
            items.append(("convolution1d_%d_b" % n, np.asarray(w["convb%d" % i], dtype=np.float32)))
        items.append(("dense_%d_W" % (t + 1), np.asarray(w["Wd"], dtype=np.float32)))
        items.append(("dense_%d_b" % (t + 1), np.asarray(w["bd"], dtype=np.float32)))
        out[group] = items
    return out


# ----------------------------------------------------------------------------- Keras-1 naming -> model
def _conv_kernel(a, theano_kernels):
    """Any Keras-1 Convolution1D kernel layout -> (k, D, F) in cross-correlation order."""
    a = np.asarray(a, dtype=np.float32)
    if a.ndim == 4 and a.shape[1] == 1:            # tf ordering (k, 1, D, F)
        a = a[:, 0]
    elif a.ndim == 4 and a.shape[3] == 1:          # th ordering (F, D, k, 1)
        a = np.transpose(a[:, :, :, 0], (2, 1, 0))
    elif a.ndim != 3:
        raise ValueError("unrecognised Convolution1D kernel shape %r" % (a.shape,))
    if theano_kernels:                             # Keras-Theano conv2d flips the filter (filter_flip=True)
        a = a[::-1]
    return np.ascontiguousarray(a)


def _tower_from_items(items, theano_kernels):
    """[(name, array)] of one Sequential tower -> {'E', 'convW0', 'convb0', ..., 'Wd', 'bd'}."""
    names = [n for n, _ in items]
    arrays = [np.asarray(a) for _, a in items]
    if all(re.fullmatch(r"param_\d+", n) for n in names):
        order = sorted(range(len(names)), key=lambda i: int(names[i].split("_")[1]))
        arrays = [arrays[i] for i in order]         # positional: E, (W, b) per window, Wd, bd
        roles = ["E"] + ["cw", "cb"] * ((len(arrays) - 3) // 2) + ["Wd", "bd"]
    else:
        def key(n):
            m = re.fullmatch(r"(embedding|convolution1d|dense)_(\d+)_(W|b)", n)
            if m is None:
                raise ValueError("unrecognised Keras-1 weight name %r" % n)
            return ({"embedding": 0, "convolution1d": 1, "dense": 2}[m.group(1)], int(m.group(2)), m.group(3) == "b")
        order = sorted(range(len(names)), key=lambda i: key(names[i]))
        arrays = [arrays[i] for i in order]
        roles = []
        for i in order:
            kind, _, is_b = key(names[i])
            roles.append("E" if kind == 0 else ("cb" if is_b else "cw") if kind == 1 else ("bd" if is_b else "Wd"))
    out, nw = {}, 0
    for role, a in zip(roles, arrays):
        if role == "E":
            out["E"] = np.asarray(a, dtype=np.float32)
        elif role == "cw":
            out["convW%d" % nw] = _conv_kernel(a, theano_kernels)
        elif role == "cb":
            out["convb%d" % nw] = np.asarray(a, dtype=np.float32)
            nw += 1
        else:
            out[role] = np.asarray(a, dtype=np.float32)
    for need in ("E", "Wd", "bd", "convW0", "convb0"):
        if need not in out:
            raise ValueError("tower is missing %s (got %s)" % (need, names))
    return out


def import_layers(layers, theano_kernels=False):
    """{group: [(name, array)]} (any superset: groups without weights are ignored) -> {'query': {...}, 'doc': {...}}
    for CLSM.set_weights.  The first two groups that own weights are the query and the document tower, in that
    order (cnn_dssm_tf.py:135-137 builds them in that order)."""
    towers = [(g, it) for g, it in layers.items() if len(it) > 0]
    if len(towers) < 2:
        raise ValueError("expected two towers with weights, found %d" % len(towers))
    return {"query": _tower_from_items(towers[0][1], theano_kernels),
            "doc": _tower_from_items(towers[1][1], theano_kernels)}


# ----------------------------------------------------------------------------- .npz
def save_npz(path, layers):
    flat = {"__layer_names__": np.array(list(layers), dtype="S")}
    for g, items in layers.items():
        flat["__weight_names__/" + g] = np.array([n for n, _ in items], dtype="S")
        for n, a in items:
            flat[g + "/" + n] = np.asarray(a)
    with open(path, "wb") as f:                    # np.savez would append '.npz' to other extensions
        np.savez(f, **flat)


def load_npz(path):
    with np.load(path, allow_pickle=False) as z:
        out = OrderedDict()
        for g in [x.decode() for x in z["__layer_names__"]]:
            names = [x.decode() for x in z["__weight_names__/" + g]]
            out[g] = [(n, z[g + "/" + n]) for n in names]
    return out


# ----------------------------------------------------------------------------- .h5 (h5py, when present)
_NO_H5PY = ("h5py is not installed here.  Convert the file where it is (any machine with h5py):\n"
            "    python scripts/keras_h5_to_npz.py to-npz weights.hdf5 weights.npz      # then load_weights('weights.npz')\n"
            "    python scripts/keras_h5_to_npz.py to-h5 weights.npz weights.hdf5       # after save_weights('weights.npz')")


def load_h5(path):
    """Keras-1 save_weights / ModelCheckpoint / model.save file -> {group: [(name, array)]}."""
    try:
        import h5py
    except ImportError as e:
        raise ImportError(_NO_H5PY) from e
    with h5py.File(path, "r") as f:
        root = f["model_weights"] if "model_weights" in f else f       # model.save() nests the weights
        out = OrderedDict()
        for g in [n.decode() if isinstance(n, bytes) else n for n in root.attrs["layer_names"]]:
            grp = root[g]
            names = [n.decode() if isinstance(n, bytes) else n for n in grp.attrs["weight_names"]]
            out[g] = [(n, np.asarray(grp[n])) for n in names]
    return out


def save_h5(path, layers):
    try:
        import h5py
    except ImportError as e:
        raise ImportError(_NO_H5PY) from e
    with h5py.File(path, "w") as f:
        f.attrs["layer_names"] = [g.encode() for g in layers]
        for g, items in layers.items():
            grp = f.create_group(g)
            grp.attrs["weight_names"] = [n.encode() for n, _ in items]
            for n, a in items:
                grp.create_dataset(n, data=np.asarray(a))


def container_of(path):
    p = str(path).lower()
    if p.endswith(".npz"):
        return "npz"
    if p.endswith((".h5", ".hdf5", ".hdf")):
        return "h5"
    return "torch"
