"""N3: time eb_model_load_gguf on a Vicuna-7B-shaped Q4_K_M GGUF (4.0 GB, written with the reference's gguf-py) against the tensor-by-tensor
upload from host memory (eb_model_set_tensor) of the same bytes.  usage: python tools/gguf_load_bench.py"""
import os, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import bench
from __graft_entry__ import load_package
import fixtures as fx, fullsize_run as fr, gguf_fixture as gf
eb = load_package(); eb.init(0)
tgt = fr.random_model(bench.HP, "q4_k_m", 7)
tmp = tempfile.mkdtemp(prefix="eb_n3_")
path = gf.write_gguf(tgt, os.path.join(tmp, "m.gguf"), n_ctx_train=8192)
print("file", os.path.getsize(path) / 1e9, "GB", eb.gguf_probe(path))
for rep in range(3):
    m = eb.Model.from_gguf(path)
    print(f"eb_model_load_gguf: {m.load_bytes / 1e9:.2f} GB in {m.load_seconds:.3f} s = {m.load_bytes / 1e9 / m.load_seconds:.2f} GB/s (page cache warm: {rep > 0})")
    if rep < 2:
        m.free()
t0 = time.perf_counter(); r = fx.to_gpu_model(eb, tgt); t1 = time.perf_counter() - t0
print(f"eb_model_set_tensor per tensor from pageable host memory: {t1:.3f} s = {m.load_bytes / 1e9 / t1:.2f} GB/s")
same = all(np.array_equal(m.get_tensor(l, w, raw.nbytes), r.get_tensor(l, w, raw.nbytes)) for l, d in enumerate(tgt["layers"]) for w, (t, raw) in d.items() if l in (0, 31))
print("layers 0 and 31 identical in both models:", same)
# the reference's own loader (llama_model_load_from_file: mmap + one ggml_backend_tensor_set per tensor) uploading the same file through the plug-in
try:
    import refround as rr
    plug = os.path.join(ROOT, "eagle-in-llama.cpp_b200", "ggml_backend", "libggml-b200.so")
    if rr.have() and rr.lib().rr_load_backend(plug.encode()) >= 2:
        for rep in range(2):
            t0 = time.perf_counter(); mp = rr.Model(path, n_gpu_layers=99); t2 = time.perf_counter() - t0
            print(f"reference loader -> libggml-b200.so (set_tensor per tensor): {t2:.3f} s = {m.load_bytes / 1e9 / t2:.2f} GB/s")
            mp.free()
except Exception as e:
    print("reference loader not timed:", e)
os.remove(path); os.rmdir(tmp)
