"""One full-size tree-verify forward of a BASELINE config: engine (GPU) vs the reference's llama_decode on ggml-cpu, in the ggml
variant named on the command line ("ref" = unmodified; "f32acc" = the attribution build of oracle/Makefile whose ONLY change is
the f32 flash-attention V accumulator).  Run as a script by tests/test_fullsize.py (one ggml build per process); prints one JSON
line.  Test infrastructure.

Both sides get the same committed prefix: the prompt is prefilled by the ENGINE (a 4K-token CPU prefill of a 7B model takes
minutes) and its K/V cells are copied bit for bit into the reference's cache tensors; everything the verify step itself does is
the reference's own code on one side and the engine on the other."""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE); sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))

import fixtures as fx          # noqa: E402
import gguf_fixture as gf      # noqa: E402
import parity as pa            # noqa: E402
import pyoracle as po          # noqa: E402
import refround as rr          # noqa: E402
from conftest import load_package  # noqa: E402
from golden_io import nmse     # noqa: E402

HP7B = dict(E=4096, H=32, F=11008, V=32000, L=32)
CONFIGS = {
    "config1": dict(name="config 1: Vicuna-7B shape Q4_0, 26-node tree, 512-token prompt", ftype=po.Q4_0, T=27, n_prefix=512, n_ctx=1024, seed=101),
    "config2": dict(name="config 2: Vicuna-7B shape Q4_K_M, 64-node tree, 4K KV", ftype="q4_k_m", T=65, n_prefix=4031, n_ctx=4352, seed=102),
    # diagnostics (not part of the test suite): one factor of config 2 at a time
    "diagA": dict(name="diag A: Q4_K_M, T 65, 512-token prompt", ftype="q4_k_m", T=65, n_prefix=512, n_ctx=1024, seed=102),
    "diagB": dict(name="diag B: Q4_0, T 65, 4K KV", ftype=po.Q4_0, T=65, n_prefix=4031, n_ctx=4352, seed=102),
    "diagC": dict(name="diag C: Q4_K_M, T 27, 4K KV", ftype="q4_k_m", T=27, n_prefix=4031, n_ctx=4352, seed=102),
}


def random_model(hp, ftype, seed, branch_gain=None):
    """well-formed random blocks of the config's type mixture (quantising 6.7e9 Gaussian weights with the reference quantiser would
    take minutes; any block bytes are a valid tensor for both sides).  Conditioning: a deep stack of i.i.d. random layers is
    chaotic — with std-0.02 embeddings and std-0.02 output projections every layer rewrites the residual stream, and ANY 1e-7
    difference (a different f32 summation order, one activation rounding flip) grows ~1.3x per layer: the reference's own two
    accumulator variants then disagree by 1e-1 on the logits of a 32-layer model (tests/test_fullsize.py reports that number).  So
    the fixture is initialised the way deep transformers are: unit-scale embeddings and residual-branch output projections
    (attn_output, ffn_down) scaled by branch_gain = 1/sqrt(2 L) (GPT-2 / Megatron init), which keeps the per-layer update a
    fraction of the stream — as it is in a trained checkpoint."""
    rng = np.random.default_rng(seed)
    E, F, V, L = hp["E"], hp["F"], hp["V"], hp["L"]
    if branch_gain is None:
        branch_gain = 1.0 / np.sqrt(2.0 * L)

    def mat(rows, cols, t, scale=0.02):
        return (t, fx.random_blocks(t, rows * cols, rng, scale))

    def vec(n):
        return (po.F32, (1 + rng.normal(0, 0.02, n)).astype(np.float32).view(np.uint8).reshape(-1).copy())

    def lt(which, l):
        return fx.q4_k_m_type(which, l, L) if ftype == "q4_k_m" else ftype

    base = po.Q4_K if ftype == "q4_k_m" else ftype
    out_t = po.Q6_K if ftype in ("q4_k_m", po.Q4_0) else ftype
    g = {po.T_TOK_EMBD: mat(V, E, base, 1.0), po.T_OUTPUT_NORM: vec(E), po.T_OUTPUT: mat(V, E, out_t)}
    layers = []
    for l in range(L):
        layers.append({po.L_ATTN_NORM: vec(E), po.L_WQ: mat(E, E, lt(po.L_WQ, l)), po.L_WK: mat(E, E, lt(po.L_WK, l)), po.L_WV: mat(E, E, lt(po.L_WV, l)),
                       po.L_WO: mat(E, E, lt(po.L_WO, l), 0.02 * branch_gain), po.L_FFN_NORM: vec(E), po.L_GATE: mat(F, E, lt(po.L_GATE, l)),
                       po.L_UP: mat(F, E, lt(po.L_UP, l)), po.L_DOWN: mat(E, F, lt(po.L_DOWN, l), 0.02 * branch_gain)})
    return {"hparams": dict(**hp, eps=1e-5, freq_base=10000.0, is_eagle=False), "globals": g, "layers": layers}


def run(cfg_id, variant, tmp_dir):
    cfg = CONFIGS[cfg_id]
    eb = load_package(); eb.init(0)
    t0 = time.time()
    hp = HP7B
    tgt = random_model(hp, cfg["ftype"], cfg["seed"])
    path = gf.write_gguf(tgt, os.path.join(tmp_dir, f"tgt7b_{cfg_id}.gguf"), n_ctx_train=8192)
    rng = np.random.default_rng(cfg["seed"] + 1)
    T, n_prefix, n_ctx = cfg["T"], cfg["n_prefix"], cfg["n_ctx"]
    par = fx.driver_like_tree(T, 8, rng); depth = fx.tree_depths(par)
    seqs, n_seq = fx.tree_seq_ids(par)
    prompt = rng.integers(3, hp["V"], n_prefix).astype(np.int32)
    tokens = rng.integers(3, hp["V"], T).astype(np.int32)
    pos = (n_prefix + depth).astype(np.int32)
    mg = fx.to_gpu_model(eb, tgt)
    s = eb.Session(mg, None, n_ctx=n_ctx, max_nodes=max(T, 16), use_graph=False)
    s.prefill(prompt)                                                        # engine prefill in 128-row chunks (eb_session_prefill)
    m = rr.Model(path, variant=variant)
    c = rr.Context(m, n_ctx, n_batch=512, n_seq_max=max(8, n_seq), n_threads=os.cpu_count() or 8)
    for l in range(hp["L"]):
        for which in (0, 1):
            c.kv_write(l, which, 0, s.kv_read(0, l, which, 0, n_prefix))
    c.kv_set_prefix(n_prefix, n_seq)
    os.remove(path)
    t1 = time.time()
    lo, taps = c.decode_tapped(tokens, pos, seqs)
    t2 = time.time()
    lg, fg, pl_g = s.forward_tapped(0, tokens, pos, parent=par)
    pl_o = pa.planes_from_ref_taps(taps, m.tensor_type)
    dirty, primary, per_site = pa.classify_rows(pl_g, pl_o, po.tree_vis(par).astype(bool))
    rel = pa.row_rel(lg, lo)
    clean = ~dirty
    first = next((i for i, n in enumerate(per_site) if n), -1)
    dev = float(rel.max())
    bad, undecided, _ = pa.greedy_report(lg, lo, max(2 * dev, pa.TIE_REL))
    k_g = s.kv_read(0, 0, 0, n_prefix, T); v_g = s.kv_read(0, 0, 1, n_prefix, T)
    k_o = np.stack([c.kv_row(0, 0, n_prefix + i, hp["E"]) for i in range(T)]); v_o = np.stack([c.kv_row(0, 1, n_prefix + i, hp["E"]) for i in range(T)])
    out = dict(config=cfg["name"], variant=variant, T=T, n_kv=n_prefix + T, rel_max=dev, rel_median=float(np.median(rel)), nmse=nmse(lg, lo),
               rows_differing_planes=int(dirty.sum()), rows_primary_flip=int(primary.sum()), first_differing_site=first, n_sites=len(per_site),
               clean_rel_max=float(rel[clean].max()) if clean.any() else None, differing_rel_max=float(rel[dirty].max()) if dirty.any() else None,
               greedy_identical=int(T - undecided - bad), greedy_undecided=int(undecided), greedy_mismatch=int(bad),
               kv_layer0_bit_identical=[float((k_g == k_o).mean()), float((v_g == v_o).mean())],
               setup_s=round(t1 - t0, 1), reference_decode_s=round(t2 - t1, 2), host_threads=os.cpu_count())
    if os.environ.get("EB_FULLSIZE_DEBUG"):
        frac = [float((a != b).mean()) for a, b in zip(pl_g, pl_o)]
        big = [float((np.abs(a.astype(np.int32) - b) > 1).mean()) for a, b in zip(pl_g, pl_o)]
        print("element fraction differing per site:", np.round(frac, 4).tolist(), file=sys.stderr)
        print("element fraction differing by more than 1:", np.round(big, 4).tolist(), file=sys.stderr)
        d0 = (pl_g[0] != pl_o[0])
        print("site0 differing rows", np.nonzero(d0.any(1))[0].tolist(), "elements", int(d0.sum()), file=sys.stderr)
        r0 = int(np.nonzero(d0.any(1))[0][0]) if d0.any() else 0
        cols = np.nonzero(d0[r0])[0][:8]
        print("row", r0, "cols", cols.tolist(), "gpu", pl_g[0][r0, cols].tolist(), "ref", pl_o[0][r0, cols].tolist(), "ref x", taps[0][1][r0, cols].tolist(), file=sys.stderr)
        t, raw = tgt["globals"][po.T_TOK_EMBD]
        emb = po.ref().get_rows(t, raw, hp["E"], hp["V"], tokens)
        x = po.ref().rms_norm(emb, 1e-5) * tgt["layers"][0][po.L_ATTN_NORM][1].view(np.float32)[None, :]
        print("reference tap site0 == recomputed norm(embedding):", np.array_equal(x.astype(np.float32), taps[0][1]), float(np.abs(x - taps[0][1]).max()), file=sys.stderr)
        print("tap names[:8]", [n for n, _ in taps[:8]], file=sys.stderr)
    np.savez(os.path.join(tmp_dir, f"logits_{cfg_id}_{variant}.npz"), engine=lg, reference=lo)
    s.free(); c.free(); m.free()
    return out


if __name__ == "__main__":
    res = run(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "/tmp")
    print("FULLSIZE_JSON " + json.dumps(res))
