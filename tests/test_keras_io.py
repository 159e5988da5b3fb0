"""Keras-1 checkpoint interchange, CPU: layer / weight naming as the reference's HDF5 files have it
(dssm_cnn_v2/cnn_dssm_tf.py:193,199-206), both Convolution1D kernel layouts, the Theano flip, the 'param_N' names
of the earliest 1.0.x files, and the .npz container."""
import numpy as np
import pytest

from dnn_page_vectors_b200 import keras_io as K


def tower(rng, V, D, ks, F, H):
    w = {"E": rng.randn(V, D).astype(np.float32), "Wd": rng.randn(len(ks) * F, H).astype(np.float32),
         "bd": rng.randn(H).astype(np.float32)}
    for i, k in enumerate(ks):
        w["convW%d" % i] = rng.randn(k, D, F).astype(np.float32)
        w["convb%d" % i] = rng.randn(F).astype(np.float32)
    return w


def same(a, b):
    assert set(a) == set(b)
    for t in a:
        assert set(a[t]) == set(b[t])
        for k in a[t]:
            np.testing.assert_array_equal(a[t][k], b[t][k])


def test_export_uses_keras1_names_numbering_and_4d_kernels():
    rng = np.random.RandomState(0)
    W = {"query": tower(rng, 11, 6, (3, 4), 5, 7), "doc": tower(rng, 11, 6, (3, 4), 5, 7)}
    L = K.export_layers(W, 2)
    assert list(L) == ["sequential_1", "sequential_2"]
    assert [n for n, _ in L["sequential_1"]] == ["embedding_1_W", "convolution1d_1_W", "convolution1d_1_b",
                                                 "convolution1d_2_W", "convolution1d_2_b", "dense_1_W", "dense_1_b"]
    assert [n for n, _ in L["sequential_2"]][1] == "convolution1d_3_W"          # layers are numbered in creation order
    assert dict(L["sequential_1"])["convolution1d_1_W"].shape == (3, 1, 6, 5)   # (k, 1, D, F): tf dimension ordering
    same(K.import_layers(L), W)


def test_npz_round_trip(tmp_path):
    rng = np.random.RandomState(1)
    W = {"query": tower(rng, 9, 4, (2,), 3, 5), "doc": tower(rng, 9, 4, (2,), 3, 5)}
    p = str(tmp_path / "weights.03.npz")
    K.save_npz(p, K.export_layers(W, 1))
    same(K.import_layers(K.load_npz(p)), W)
    assert K.container_of(p) == "npz" and K.container_of("w.hdf5") == "h5" and K.container_of("w.pt") == "torch"


def test_theano_ordering_flip_param_names_and_foreign_groups():
    """A file as an early Keras-1 / Theano run writes it: kernels (F, D, k, 1) stored FLIPPED along the window axis,
    weights named param_N, weightless layers (inputs, lambdas, merge) listed between the towers."""
    rng = np.random.RandomState(2)
    W = {"query": tower(rng, 9, 4, (3, 4), 3, 5), "doc": tower(rng, 9, 4, (3, 4), 3, 5)}
    L = K.export_layers(W, 2)

    def theano_items(items, rename):
        out = []
        for i, (n, a) in enumerate(items):
            if a.ndim == 4:
                a = np.transpose(a[:, 0][::-1], (2, 1, 0))[:, :, :, None]
            out.append(("param_%d" % i if rename else n, a))
        return out
    layers = {"input_1": [], "sequential_1": theano_items(L["sequential_1"], True), "lambda_1": [],
              "sequential_2": theano_items(L["sequential_2"], False), "merge_1": []}
    same(K.import_layers(layers, theano_kernels=True), W)
    with pytest.raises(ValueError):
        K.import_layers({"sequential_1": L["sequential_1"]})                  # one tower only
    with pytest.raises(ValueError):
        K.import_layers({"a": [("mystery_W", np.zeros(3))], "b": L["sequential_2"]})


def test_hdf5_route_needs_h5py_and_says_how_to_convert(tmp_path):
    try:
        import h5py  # noqa: F401
    except ImportError:
        with pytest.raises(ImportError, match="keras_h5_to_npz"):
            K.load_h5(str(tmp_path / "x.h5"))
        return
    rng = np.random.RandomState(3)
    W = {"query": tower(rng, 9, 4, (3,), 3, 5), "doc": tower(rng, 9, 4, (3,), 3, 5)}
    p = str(tmp_path / "w.h5")
    K.save_h5(p, K.export_layers(W, 1))
    same(K.import_layers(K.load_h5(p)), W)
