"""Host feed rate: pairs/s the JSON generator (the reference's scheme: parse, clean, split, pad, map every
line of every epoch) and the pre-tokenised generator deliver at the cfg2 shape (J=4, Lq=20, Ld=100).  CPU only.
    python scripts/feed_bench.py [n_records]"""
import json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dnn_page_vectors_b200 import Configuration, DataHelpers

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
rng = np.random.RandomState(0)
words = ["w%d" % i for i in range(20000)]
with tempfile.TemporaryDirectory() as root:
    conf = Configuration(experiment_root_directory=root, feature_level="word", num_negative_examples=4, query_length=20, document_length=100)
    os.makedirs(conf.data_dir)
    sent = lambda k: " ".join(words[i] for i in rng.zipf(1.3, size=k) % len(words))
    lines = [json.dumps({"q": sent(12), "doc_corr": sent(80), "doc_incorr": [sent(80) for _ in range(4)]}) for _ in range(n)]
    for path in conf.input_file_list:
        open(path, "w").write("\n".join(lines) + "\n")
    dh = DataHelpers(conf)
    _, vocab = dh.generate_vocabulary_set()
    t = time.perf_counter(); cnt = sum(len(y) for _, y in dh.load_data_generator(vocab, "training", 1024, nb_epochs=0)); dt_json = time.perf_counter() - t
    t = time.perf_counter(); packed = dh.pretokenize(vocab, "training"); dt_pack = time.perf_counter() - t
    nw = os.cpu_count() or 1
    t = time.perf_counter(); dh.pretokenize(vocab, "training", out_dir=os.path.join(root, "packed_mp"), chunk_lines=256, workers=nw); dt_mp = time.perf_counter() - t
    t = time.perf_counter(); cnt2 = sum(len(y) for _, y in dh.load_packed_generator(packed, 1024, nb_epochs=8)); dt_mm = time.perf_counter() - t
    print(json.dumps({"records": n, "json_generator_pairs_per_s": cnt / dt_json, "pretokenize_pairs_per_s": n / dt_pack,
                      "pretokenize_%d_workers_pairs_per_s" % nw: n / dt_mp,
                      "packed_generator_pairs_per_s": cnt2 / dt_mm}))
