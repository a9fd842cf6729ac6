import sys, os, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests"); sys.path.insert(0, "oracle")
from __graft_entry__ import load_package
import fixtures as fx, pyoracle as po
from oracle_round import CpuStream
eb = load_package(); eb.init(0)
hp = dict(E=256, H=2, F=512, V=1024, L=2)
tgt = fx.make_model(**hp, ftype="q4_k_m", seed=31)
head = fx.make_model(**{**hp, "L": 1}, ftype=po.Q4_0, seed=32, is_eagle=True)
cfg = dict(n_branch=4, n_draft=12, max_depth=5, top_k=10, p_split=0.1, max_nodes=16)
po.set_fattn_acc_f32(1)
mt = fx.to_gpu_model(eb, tgt); mh = fx.to_gpu_model(eb, head, target=mt)
cpu = CpuStream(po.oracle(), tgt, head, 128)
s = eb.Session(mt, mh, n_ctx=128, use_graph=False, **cfg)
prompt = np.random.default_rng(1).integers(3, hp["V"], 12).astype(np.int32)
lg_c, _ = cpu.prefill(prompt)
lg, ft = s.forward(0, prompt, np.arange(len(prompt)))
hid = np.zeros((len(prompt), hp["E"]), np.float32); hid[1:] = ft[:-1]
s.forward(1, prompt, np.arange(len(prompt)), hidden=hid)
s.commit(0, len(prompt)); s.set_root(cpu.root_token, ft[-1])
want, got = cpu.round(**cfg), s.round()
n = got.n_nodes
print("want tokens", want["tokens"].tolist()); print("got  tokens", list(got.tokens[:n]))
print("want parent", want["parent"].tolist()); print("got  parent", list(got.parent[:n]))
for ids, v in want["cand"]:
    print("cand", list(ids[:4]), np.round(v[:4], 4), "min gap", float(np.min(-np.diff(v))))
