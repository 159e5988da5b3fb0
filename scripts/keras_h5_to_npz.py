#!/usr/bin/env python
"""Keras-1 HDF5 weights <-> the .npz interchange of dnn_page_vectors_b200 (run where h5py is installed; needs only
numpy + h5py, not this package).

    python keras_h5_to_npz.py to-npz weights.03.hdf5 weights.03.npz     # reference checkpoint -> CLSM.load_weights
    python keras_h5_to_npz.py to-h5  weights.npz cnn_dssm_model_weights.h5   # CLSM.save_weights -> Keras-1 load_weights

The HDF5 side is what Keras 1.x writes (dssm_cnn_v2/cnn_dssm_tf.py:193,204-206): root attribute layer_names, one group
per layer with attribute weight_names and one dataset per weight; model.save() nests the same under /model_weights."""
import sys
import h5py
import numpy as np

mode, src, dst = sys.argv[1:4]
if mode == "to-npz":
    flat = {}
    with h5py.File(src, "r") as f:
        root = f["model_weights"] if "model_weights" in f else f
        layers = [n.decode() if isinstance(n, bytes) else n for n in root.attrs["layer_names"]]
        flat["__layer_names__"] = np.array(layers, dtype="S")
        for g in layers:
            names = [n.decode() if isinstance(n, bytes) else n for n in root[g].attrs["weight_names"]]
            flat["__weight_names__/" + g] = np.array(names, dtype="S")
            for n in names:
                flat[g + "/" + n] = np.asarray(root[g][n])
    with open(dst, "wb") as out:
        np.savez(out, **flat)
elif mode == "to-h5":
    with np.load(src, allow_pickle=False) as z, h5py.File(dst, "w") as f:
        layers = [x.decode() for x in z["__layer_names__"]]
        f.attrs["layer_names"] = [g.encode() for g in layers]
        for g in layers:
            names = [x.decode() for x in z["__weight_names__/" + g]]
            grp = f.create_group(g)
            grp.attrs["weight_names"] = [n.encode() for n in names]
            for n in names:
                grp.create_dataset(n, data=z[g + "/" + n])
else:
    sys.exit("usage: keras_h5_to_npz.py to-npz|to-h5 SRC DST")
