"""Is eb_session_prefill reproducible?  Prefills the same 4031-token prompt several times (7B Q4_K_M synthetic target, the full-size test's
shape) and prints a hash of the last token's logits and of the K rows of three layers per repetition; every repetition must print the same
line.  EB_PREFILL_CHUNK picks the rows per pass (read once per process).  usage: EB_PREFILL_CHUNK=80 python tools/prefill_determinism.py [reps=4]"""
import hashlib, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import bench
from __graft_entry__ import load_package
eb = load_package(); eb.init(0)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
tgt, head = bench.build_gpu_models(eb)
s = eb.Session(tgt, None, n_ctx=4096 + 256, use_graph=False, **bench.TREE)
toks = np.random.default_rng(1).integers(3, 32000, 4031).astype(np.int32)
lines = []
for rep in range(reps):
    s.set_n_past(0)
    lg, ft = s.prefill(toks)
    h = [hashlib.sha1(lg.tobytes()).hexdigest()[:10]]
    for layer in (0, 15, 31):
        h.append(hashlib.sha1(s.kv_read(0, layer, 0, 0, 4031).tobytes()).hexdigest()[:10])
    lines.append(" ".join(h)); print(f"rep {rep} chunk {os.environ.get('EB_PREFILL_CHUNK', '128')}: logits / K layer 0 / 15 / 31: {lines[-1]}", flush=True)
print("REPRODUCIBLE" if len(set(lines)) == 1 else "DIFFERS BETWEEN REPETITIONS")
