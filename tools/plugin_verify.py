"""the 7B Q4_K_M verify step (65-token tree batch, n_kv 4096) through the reference's llama_decode with every layer on libggml-b200.so —
for launch lists (ncu) and A/B timing of the plug-in.  usage: python tools/plugin_verify.py [n_decodes=5]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench
import fixtures as fx, fullsize_run as fr, gguf_fixture as gf, refround as rr
import tempfile
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5
variant = "native" if rr.native_ok() else "ref"
rr.lib(variant=variant)
tgt = fr.random_model(bench.HP, "q4_k_m", 7)
tmp = tempfile.mkdtemp(prefix="eb_plug_")
path = gf.write_gguf(tgt, os.path.join(tmp, "m.gguf"), n_ctx_train=8192)
plug = os.path.join(ROOT, "eagle-in-llama.cpp_b200", "ggml_backend", "libggml-b200.so")
assert rr.lib().rr_load_backend(plug.encode()) >= 2
mp = rr.Model(path, n_gpu_layers=99, variant=variant)
os.remove(path); os.rmdir(tmp)
rng = np.random.default_rng(11)
par = fx.driver_like_tree(65, bench.TREE["n_branch"], rng); depth = fx.tree_depths(par); seqs, n_seq = fx.tree_seq_ids(par)
cp = rr.Context(mp, 4096 + 256, n_batch=512, n_seq_max=max(8, n_seq), n_threads=len(os.sched_getaffinity(0)))
for l in range(bench.HP["L"]):
    for which in (0, 1):
        cp.kv_write(l, which, 0, rng.standard_normal((bench.N_PAST0, bench.HP["E"]), np.float32).astype(np.float16))
tokens = rng.integers(3, bench.HP["V"], 65).astype(np.int32)
ts = []
for _ in range(n):
    cp.kv_set_prefix(bench.N_PAST0, n_seq)
    ts.append(cp.time_decode(tokens, bench.N_PAST0 + depth, seqs))
print("plugin verify step us:", ts, "median", float(np.median(ts[1:] if n > 1 else ts)))
