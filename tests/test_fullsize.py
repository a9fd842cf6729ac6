"""GPU: one full-size tree-verify forward per BASELINE config against the UNMODIFIED reference (its own libllama / llama_decode
on ggml-cpu, oracle/_ref) — 32 layers, E 4096, F 11008, 32 heads, V 32000, the config's quantisation mix, tree shape and KV
length.  The committed prefix is synthetic on both sides (identical f16 K/V bytes written into the reference's cache tensors and
into the engine's cache: a 512..4K-token CPU prefill of a 7B model is minutes of host time, and the prefix is an INPUT of the
verify step, not part of it); everything the step itself does is the reference's code on one side and the engine on the other.

Reported per config (pytest -s / the driver's log): logits max-rel and NMSE over the 27 / 65 rows, rows whose int8 activation
planes differ anywhere in the 32 layers (engine tap vs the reference's cb_eval tap), greedy-token agreement.  The reference keeps
its flash-attention V accumulator in f16 (R/ggml/src/ggml-cpu/ggml-cpu.c:11154-11168; the engine accumulates in f32 — the one
documented deviation), so against it every row inherits an O(1e-3) perturbation at the first wo input: the bar here is the
north star's 1e-2 on the logits of every row, the reference's whole-graph NMSE bar 2e-3, and greedy ids identical wherever the
reference's own top-2 margin exceeds twice the measured deviation.
"""
import os
import time

import numpy as np
import pytest

import fixtures as fx
import gguf_fixture as gf
import parity as pa
import pyoracle as po
import refround as rr
from golden_io import nmse

pytestmark = pytest.mark.gpu
HP7B = dict(E=4096, H=32, F=11008, V=32000, L=32)


def random_model(hp, ftype, seed):
    """well-formed random blocks of the config's type mixture (quantising 6.7e9 Gaussian weights with the reference quantiser would
    take minutes; any block bytes are a valid tensor for both sides)"""
    rng = np.random.default_rng(seed)
    E, F, V, L = hp["E"], hp["F"], hp["V"], hp["L"]

    def mat(rows, cols, t):
        return (t, fx.random_blocks(t, rows * cols, rng))

    def vec(n):
        return (po.F32, (1 + rng.normal(0, 0.02, n)).astype(np.float32).view(np.uint8).reshape(-1).copy())

    def lt(which, l):
        return fx.q4_k_m_type(which, l, L) if ftype == "q4_k_m" else ftype

    base = po.Q4_K if ftype == "q4_k_m" else ftype
    out_t = po.Q6_K if ftype in ("q4_k_m", po.Q4_0) else ftype
    g = {po.T_TOK_EMBD: mat(V, E, base), po.T_OUTPUT_NORM: vec(E), po.T_OUTPUT: mat(V, E, out_t)}
    layers = []
    for l in range(L):
        layers.append({po.L_ATTN_NORM: vec(E), po.L_WQ: mat(E, E, lt(po.L_WQ, l)), po.L_WK: mat(E, E, lt(po.L_WK, l)), po.L_WV: mat(E, E, lt(po.L_WV, l)),
                       po.L_WO: mat(E, E, lt(po.L_WO, l)), po.L_FFN_NORM: vec(E), po.L_GATE: mat(F, E, lt(po.L_GATE, l)),
                       po.L_UP: mat(F, E, lt(po.L_UP, l)), po.L_DOWN: mat(E, F, lt(po.L_DOWN, l))})
    return {"hparams": dict(**hp, eps=1e-5, freq_base=10000.0, is_eagle=False), "globals": g, "layers": layers}


@pytest.mark.parametrize("cfg", [dict(name="config 1: 7B Q4_0, 26-node tree, 512-token prefix", ftype=po.Q4_0, T=27, n_prefix=512, n_ctx=1024, seed=101),
                                 dict(name="config 2: 7B Q4_K_M, 64-node tree, 4K KV", ftype="q4_k_m", T=65, n_prefix=4031, n_ctx=4352, seed=102)],
                         ids=["config1", "config2"])
def test_fullsize_verify_forward_vs_reference(eb, tmp_path, cfg):
    if not rr.have():
        pytest.fail("oracle/_ref (the reference build) is missing: run __graft_entry__.build() where /root/reference is present")
    t0 = time.time()
    hp = HP7B
    tgt = random_model(hp, cfg["ftype"], cfg["seed"])
    path = gf.write_gguf(tgt, str(tmp_path / "tgt7b.gguf"), n_ctx_train=8192)
    rng = np.random.default_rng(cfg["seed"] + 1)
    T, n_prefix, n_ctx = cfg["T"], cfg["n_prefix"], cfg["n_ctx"]
    par = fx.driver_like_tree(T, 8, rng); depth = fx.tree_depths(par)
    seqs, n_seq = fx.tree_seq_ids(par)
    tokens = rng.integers(3, hp["V"], T).astype(np.int32)
    pos = (n_prefix + depth).astype(np.int32)
    # ---- engine
    mg = fx.to_gpu_model(eb, tgt)
    s = eb.Session(mg, None, n_ctx=n_ctx, max_nodes=max(T, 16), use_graph=False)
    # ---- reference
    m = rr.Model(path)
    c = rr.Context(m, n_ctx, n_batch=512, n_seq_max=max(8, n_seq), n_threads=os.cpu_count() or 8)
    for l in range(hp["L"]):
        for which in (0, 1):
            rows = rng.standard_normal((n_prefix, hp["E"]), np.float32).astype(np.float16)
            s.kv_write(0, l, which, 0, rows); c.kv_write(l, which, 0, rows)
    s.set_n_past(n_prefix); c.kv_set_prefix(n_prefix, n_seq)
    t1 = time.time()
    lo, taps = c.decode_tapped(tokens, pos, seqs)
    t2 = time.time()
    lg, fg, pl_g = s.forward_tapped(0, tokens, pos, parent=par)
    pl_o = pa.planes_from_ref_taps(taps, m.tensor_type)
    dirty, per_site = pa.flipped_rows(pl_g, pl_o)
    rel = pa.row_rel(lg, lo)
    first = next((i for i, n in enumerate(per_site) if n), -1)
    bad, undecided, _ = pa.greedy_report(lg, lo, 2 * float(rel.max()))
    print(f"[fullsize] {cfg['name']}: T {T}, n_kv {n_prefix + T}; logits vs the unmodified reference: max rel {rel.max():.2e}, median {np.median(rel):.2e}, "
          f"NMSE {nmse(lg, lo):.2e}; rows with differing int8 planes {int(dirty.sum())} of {T} (first differing site {first} of {len(per_site)}); greedy: "
          f"{T - undecided - bad} of {T} rows identical, {undecided} undecided at twice the measured deviation, {bad} mismatches; setup {t1 - t0:.0f} s, reference decode {t2 - t1:.1f} s")
    assert rel.max() <= pa.NORTH_STAR_REL, float(rel.max())
    assert nmse(lg, lo) <= 2e-3
    assert bad == 0
    # the K / V rows the forward stored for the tree cells (RoPE'd K, V) against the reference's cache tensors: f16, to the last bit
    # wherever the row's layer-0 planes agree (layer 0 only: deeper layers inherit the attention deviation)
    k_g = s.kv_read(0, 0, 0, n_prefix, T); v_g = s.kv_read(0, 0, 1, n_prefix, T)
    k_o = np.stack([c.kv_row(0, 0, n_prefix + i, hp["E"]) for i in range(T)]); v_o = np.stack([c.kv_row(0, 1, n_prefix + i, hp["E"]) for i in range(T)])
    same_k = float((k_g == k_o).mean()); same_v = float((v_g == v_o).mean())
    print(f"[fullsize] layer-0 K/V cells of the tree: {100 * same_k:.3f} % / {100 * same_v:.3f} % of f16 values bit-identical to the reference's cache")
    assert same_k > 0.99 and same_v > 0.99      # an f32 difference of 1e-6 relative moves ~0.2 % of the f16 roundings by one ulp
    s.free(); c.free(); m.free()
