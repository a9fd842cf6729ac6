"""Generates tests/golden/*.npz by running the REFERENCE's own ggml-cpu code (oracle/_ref,
compiled from /root/reference by oracle/Makefile) on seeded synthetic inputs.  Run in the build
container (needs /root/reference):   python tests/golden/make_golden.py
The vectors pin oracle/eagle_oracle.c (CPU tests) and the CUDA path (GPU tests)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
sys.path.insert(0, os.path.join(HERE, ".."))
import fixtures as fx  # noqa: E402
import pyoracle as po  # noqa: E402


def main():
    po.build(ref=True)
    r = po.ref()
    rng = np.random.default_rng(20261104)
    out = {}
    # ---- formats + mul_mat
    N, K, T = 32, 512, 3
    W = rng.normal(0, 0.02, (N, K)).astype(np.float32)
    X = rng.normal(0, 1, (T, K)).astype(np.float32)
    out["mm_x"] = X
    for t in fx.QTYPES:
        n = po.TYPE_NAMES[t]
        raw = fx.quantize(t, W)
        out[f"w_{n}"] = raw
        out[f"deq_{n}"] = r.dequantize(t, raw, N * K)
        out[f"mm_{n}"] = r.mul_mat(t, raw, N, K, X)
    out["act_q8_0"] = r.quantize_act(po.Q8_0, X[0])
    out["act_q8_K"] = r.quantize_act("q8_K", X[0])
    # ---- glue ops
    x = rng.normal(0, 1, (5, 512)).astype(np.float32)
    out["rms_x"] = x; out["rms_y"] = r.rms_norm(x, 1e-5)
    q = rng.normal(0, 1, (6, 2, 128)).astype(np.float32)
    pos = np.array([0, 1, 7, 100, 1000, 4000], np.int32)
    out["rope_x"] = q; out["rope_pos"] = pos; out["rope_y"] = r.rope_norm(q, pos, 10000.0)
    # ---- tree-masked flash attention
    n_past, Tq, H, D = 70, 9, 2, 128
    par = fx.tree_parents(Tq, rng)
    n_kv = 128
    k = rng.normal(0, 1, (n_kv, H, D)).astype(np.float16); v = rng.normal(0, 1, (n_kv, H, D)).astype(np.float16)
    qa = rng.normal(0, 1, (Tq, H, D)).astype(np.float32)
    vis = fx.vis_matrix(par, n_past, n_kv)
    out["fa_q"] = qa; out["fa_k"] = k.view(np.uint16); out["fa_v"] = v.view(np.uint16); out["fa_parent"] = par
    out["fa_n_past"] = np.int32(n_past); out["fa_out"] = r.flash_attn(qa, k.view(np.uint16), v.view(np.uint16), vis, 1 / np.sqrt(D))
    np.savez_compressed(os.path.join(HERE, "ops.npz"), **out)

    # ---- tiny target + head: chain prefill then a tree batch
    hp = dict(E=256, H=2, F=512, V=512, L=2)
    tgt = fx.make_model(**hp, ftype="q4_k_m", seed=11)
    head = fx.make_model(**{**hp, "L": 1}, ftype=po.Q4_0, seed=12, is_eagle=True)
    n_ctx = 64
    mt = fx.to_cpu_model(r, tgt); mh = fx.to_cpu_model(r, head, target=mt)
    kvt, kvh = r._g("kv_new")(hp["L"], hp["E"], n_ctx), r._g("kv_new")(1, hp["E"], n_ctx)
    prompt = rng.integers(3, hp["V"], 10).astype(np.int32)
    P = len(prompt)
    causal = np.tril(np.ones((P, P), np.uint8)); vis = np.zeros((P, n_ctx), np.uint8); vis[:, :P] = causal
    lg_p, ft_p = r.forward(mt, kvt, prompt, np.arange(P, dtype=np.int32), None, 0, n_ctx, vis, hp["E"], hp["V"])
    hid = np.zeros((P, hp["E"]), np.float32); hid[1:] = ft_p[:-1]          # head cell p: (token_p, feature_{p-1})
    lgh_p, fth_p = r.forward(mh, kvh, prompt, np.arange(P, dtype=np.int32), hid, 0, n_ctx, vis, hp["E"], hp["V"])
    Tt = 7
    par = fx.tree_parents(Tt, rng, 3)
    toks = rng.integers(3, hp["V"], Tt).astype(np.int32)
    pos = (P + fx.tree_depths(par)).astype(np.int32)
    vis_t = fx.vis_matrix(par, P, n_ctx)
    lg_t, ft_t = r.forward(mt, kvt, toks, pos, None, P, n_ctx, vis_t, hp["E"], hp["V"])
    hid_t = rng.normal(0, 1, (Tt, hp["E"])).astype(np.float32)
    lgh_t, fth_t = r.forward(mh, kvh, toks, pos, hid_t, P, n_ctx, vis_t, hp["E"], hp["V"])
    m = {"prompt": prompt, "tree_tokens": toks, "tree_parent": par, "tree_pos": pos, "tree_hidden": hid_t,
         "tgt_prefill_logits": lg_p, "tgt_prefill_feat": ft_p, "head_prefill_logits": lgh_p, "head_prefill_feat": fth_p,
         "tgt_tree_logits": lg_t, "tgt_tree_feat": ft_t, "head_tree_logits": lgh_t, "head_tree_feat": fth_t}
    for name, f in (("tgt", tgt), ("head", head)):
        for which, (t, raw) in f["globals"].items():
            m[f"{name}.g{which}.t{t}"] = raw
        for l, d in enumerate(f["layers"]):
            for which, (t, raw) in d.items():
                m[f"{name}.l{l}.{which}.t{t}"] = raw
    np.savez_compressed(os.path.join(HERE, "tiny_model.npz"), **m)
    print("golden written:", {k: os.path.getsize(os.path.join(HERE, k)) for k in ("ops.npz", "tiny_model.npz")})


if __name__ == "__main__":
    main()
