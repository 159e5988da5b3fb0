"""Configuration of the CLSM path.

``Configuration`` keeps the attribute names of the reference's class
(dssm_cnn_v2/config.py:13-91) -- paths, ``feature_level``,
``num_negative_examples``, ``query_length`` / ``document_length``, vector-file
choice -- so code written against it keeps working.  What the reference
hard-codes (root directory, timestamp, feature level) is a constructor argument
here with the reference's literal as default.

``ModelParams`` carries the module-level constants the reference scripts
duplicate (dssm_cnn_v2/cnn_dssm_tf.py:28-44).
"""
from __future__ import annotations

import json
import os
import time
from dataclasses import asdict, dataclass, field

_LENGTHS = {           # dssm_cnn_v2/config.py:80-91  (query, document)
    "word": (20, 975),
    "char": (250, 5000),
    "ngram": (45, 2000),
}


class Configuration(object):
    def __init__(self, experiment_root_directory="/ebs/project_data",
                 feature_level="char", word_vectors="fast",
                 num_negative_examples=3, reuse_experiment_timestamp=True,
                 timestamp="2016-09-16T23-04-38", query_length=None,
                 document_length=None):
        if feature_level not in _LENGTHS:
            raise ValueError("feature_level must be one of %s" % sorted(_LENGTHS))
        self.experiment_root_directory = experiment_root_directory
        self.reuse_experiment_timestamp = reuse_experiment_timestamp
        self.experiment_name = "dssm_cnn_v2"

        if not reuse_experiment_timestamp:           # config.py:26-36
            stamp_file = os.path.join(experiment_root_directory, "_TIMESTAMP")
            if os.path.exists(stamp_file):
                with open(stamp_file) as f:
                    timestamp = json.load(f)["project_timestamp"]
            else:
                timestamp = time.strftime("%Y-%m-%dT%H-%M-%S")
                os.makedirs(experiment_root_directory, exist_ok=True)
                with open(stamp_file, "w") as f:
                    json.dump({"project_timestamp": timestamp}, f)

        self.feature_level = feature_level           # config.py:38
        self.default_ngram = 3
        self.data_path = os.path.join(experiment_root_directory, self.experiment_name,
                                      timestamp, feature_level)
        self.data_dir = os.path.join(self.data_path, "data")
        self.trained_model_dir = os.path.join(self.data_path, "model")
        self.pickle_files_dir = os.path.join(self.data_path, "pickled_files")
        self.vectors_directory = os.path.join(experiment_root_directory, "vectors")

        self.word_vectors = word_vectors             # config.py:51-56
        if word_vectors == "word2vec":
            self.word_vectors_file = os.path.join(self.vectors_directory, "vectors_wholecorpus100.txt")
        elif word_vectors == "fast":
            self.word_vectors_file = os.path.join(self.vectors_directory, "fast_model_ns.vec")
        else:
            raise ValueError("word_vectors must be 'word2vec' or 'fast'")

        self.input_dataset = os.path.join(self.data_dir, "input_dataset_new.txt")
        self.model_training_data = os.path.join(self.data_dir, "model_training_data_new.txt")
        self.model_validation_data = os.path.join(self.data_dir, "model_validation_data_new.txt")
        self.input_file_list = [self.model_training_data, self.model_validation_data]

        self.create_data_dump = True
        self.vocab_set_file = os.path.join(self.pickle_files_dir, "vocab_set_{}.pkl")
        self.vocab_index_file = os.path.join(self.pickle_files_dir, "vocab_index_dict_{}.pkl")
        self.embedding_weights_file_tpl = os.path.join(
            self.pickle_files_dir, "we_embedding_weights_compact_{}.pkl")
        self.num_negative_examples = num_negative_examples   # config.py:76
        self.train_validation_split = 0.2

        lq, ld = _LENGTHS[feature_level]
        self.query_length = lq if query_length is None else query_length
        self.document_length = ld if document_length is None else document_length


@dataclass
class ModelParams:
    """dssm_cnn_v2/cnn_dssm_tf.py:28-44 literals as defaults."""
    embedding_dim: int = 100
    batch_size: int = 128
    nb_epoch: int = 15
    filter_sizes: tuple = (3, 4)
    num_filters: int = 150
    dropout_prob: tuple = (0.25, 0.5)
    hidden_dims: int = 150
    J: int = 3
    GAMMA: float = 10.0
    num_train_samples: int = 16000
    num_validation_samples: int = 4000
    activation: str = "relu"         # cnn_dssm_tf.py:88,126; 'tanh' = CLSM paper variant
    precision: str = "bf16"          # 'bf16' (tcgen05 operands, fp32 accumulate) or 'fp32'
    lr: float = 1e-3                 # Keras-1 Adam defaults (cnn_dssm_tf.py:182)
    beta_1: float = 0.9
    beta_2: float = 0.999
    epsilon: float = 1e-8
    seed: int = 1337                 # cnn_dssm_tf.py:20

    def to_dict(self):
        d = asdict(self)
        d["filter_sizes"] = list(self.filter_sizes)
        d["dropout_prob"] = list(self.dropout_prob)
        return d
