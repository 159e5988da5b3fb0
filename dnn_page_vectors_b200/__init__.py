"""dnn_page_vectors_b200 -- the CLSM (convolutional DSSM) path of collawolley/dnn_page_vectors
`dssm_cnn_v2`, rebuilt as hand-written sm_100a CUDA kernels behind a C ABI.

Host-only pieces (Configuration, DataUtils, DataHelpers) import without a GPU or the shared
library; everything that computes (CLSM, Tower, RTF, DataParallel) needs libclsm_b200.so and a
CUDA device and fails loudly without them.
"""
from .config import Configuration, ModelParams          # noqa: F401
from .data_utils import DataUtils                        # noqa: F401
from .data_helpers import DataHelpers                    # noqa: F401

__all__ = ["Configuration", "ModelParams", "DataUtils", "DataHelpers", "CLSM", "Tower", "RTF",
           "ModelCheckpoint", "DataParallel"]


def __getattr__(name):
    if name in ("CLSM", "Tower", "RTF", "ModelCheckpoint", "History"):
        from . import model
        return getattr(model, name)
    if name in ("DataParallel", "shard_range"):
        from . import parallel
        return getattr(parallel, name)
    raise AttributeError(name)
