"""CPU: the speculative-step bookkeeping of the oracle (tree mask rule, accept walk, compaction,
top-k order) against the known answers the reference's tests hold and against properties."""
import numpy as np

import fixtures as fx
import pyoracle as po


def test_top_k_known_answer():
    # R/tests/test-sampling.cpp:317-320: top_k({0.1,0.2,0.3,0.4}, k=1) keeps 0.4 ; k=3 keeps 0.4,0.3,0.2
    probs = np.array([0.1, 0.2, 0.3, 0.4], np.float32)
    ids, vals, p = po.top_k(np.log(probs), 1)
    assert ids.tolist() == [3] and abs(p[0] - 1.0) < 1e-6
    ids, vals, p = po.top_k(np.log(probs), 3)
    assert ids.tolist() == [3, 2, 1]
    assert np.allclose(p, np.array([0.4, 0.3, 0.2]) / 0.9, atol=1e-6)


def test_top_k_ties_and_order():
    x = np.zeros(1000, np.float32); x[[7, 500, 900]] = 5.0
    ids, vals, _ = po.top_k(x, 4)
    assert ids.tolist() == [7, 500, 900, 0]
    rng = np.random.default_rng(0)
    x = rng.normal(0, 1, 32000).astype(np.float32)
    ids, vals, p = po.top_k(x, 10)
    assert ids.tolist() == np.argsort(-x, kind="stable")[:10].tolist()
    assert abs(p.sum() - 1) < 1e-5 and (np.diff(vals) <= 0).all()


def test_tree_mask_equals_reference_seq_id_rule():
    """compact form (prefix + ancestor bits) == the KV-cell rule of llama_set_inputs
    (R/src/llama-context.cpp:161-193) for a batch built the way the tree driver builds it."""
    rng = np.random.default_rng(3)
    for trial in range(20):
        T, n_past = int(rng.integers(2, 40)), int(rng.integers(0, 50))
        # grow branches like speculative-eagle.cpp: each node is created by one branch (its first
        # seq id); when a branch splits, every node on its path gains the new seq id
        par = np.full(T, -1, np.int32); first_seq = np.zeros(T, np.int32); seqs = [set([0])] + [set() for _ in range(T - 1)]
        tips = {0: 0}; n_seq = 1
        for i in range(1, T):
            s = int(rng.choice(list(tips)))
            if rng.random() < 0.4 and n_seq < 60:            # split s -> new branch continuing from the same tip
                new = n_seq; n_seq += 1
                a = tips[s]
                while a >= 0:
                    seqs[a].add(new); a = par[a]
                tips[new] = tips[s]; s = new
            par[i] = tips[s]; first_seq[i] = s; seqs[i] = {s}; tips[s] = i
        depth = fx.tree_depths(par)
        n_kv = n_past + T
        cell_pos = np.concatenate([np.arange(n_past), n_past + depth]).astype(np.int32)
        cell_seq = np.zeros(n_kv, np.uint64)
        cell_seq[:n_past] = np.uint64(2 ** 64 - 1)           # committed prefix carries every seq id (seq_cp 0 -> s, :648-652)
        for i in range(T):
            cell_seq[n_past + i] = np.uint64(sum(1 << s for s in seqs[i]))
        vis_ref = po.build_kq_mask(cell_pos, cell_seq, n_past + depth, first_seq)
        assert np.array_equal(vis_ref, fx.vis_matrix(par, n_past, n_kv))
        # the product's host-side entry for a tree-shaped llama_batch (token i carries [first_seq[i]] + the branches that
        # split off below it) gives the same sets as the oracle's rule and as the parent form
        from conftest import load_package
        eb = load_package()
        seq_ids = [[int(first_seq[i])] + sorted(x for x in seqs[i] if x != first_seq[i]) for i in range(T)]
        bits = eb.tree_bits_from_batch(n_past + depth, seq_ids)
        assert np.array_equal(bits, eb.tree_bits_from_parent(par))
        got = np.array([[(bits[j, i >> 5] >> np.uint32(i & 31)) & 1 for i in range(T)] for j in range(T)], np.uint8)
        assert np.array_equal(got, vis_ref[:, n_past:])


def test_accept_walk():
    #        0
    #      / | \
    #     1  2  3        tokens 10 11 12
    #    /|     |
    #   4 5     6        tokens 20 21 22
    par = np.array([-1, 0, 0, 0, 1, 1, 3], np.int32)
    tok = np.array([5, 10, 11, 12, 20, 21, 22], np.int32)
    best = np.array([12, 0, 0, 22, 0, 0, 99], np.int32)
    path, bonus = po.accept_walk(tok, par, best)
    assert path.tolist() == [0, 3, 6] and bonus == 99
    best = np.array([10, 21, 0, 0, 0, 77, 0], np.int32)
    path, bonus = po.accept_walk(tok, par, best)
    assert path.tolist() == [0, 1, 5] and bonus == 77
    best = np.array([13, 0, 0, 0, 0, 0, 0], np.int32)       # nothing matches: only the root
    path, bonus = po.accept_walk(tok, par, best)
    assert path.tolist() == [0] and bonus == 13
    tok2 = tok.copy(); tok2[2] = 10                          # duplicate sibling tokens: first child wins
    path, _ = po.accept_walk(tok2, par, np.array([10, 0, 0, 0, 0, 0, 0], np.int32))
    assert path.tolist() == [0, 1]


def test_kv_compaction_moves_accepted_rows():
    o = po.oracle()
    L, E, n_ctx, n_past = 2, 256, 32, 5
    kv = o._g("kv_new")(L, E, n_ctx)
    ks, vs = o.kv_arrays(kv, L, E, n_ctx)
    rng = np.random.default_rng(0)
    for a in ks + vs:
        a[:] = rng.integers(0, 65535, a.shape, dtype=np.uint16)
    before = [a.copy() for a in ks + vs]
    path = np.array([0, 3, 6], np.int32)
    o._g("kv_compact")(kv, n_past, po._ptr(path), 3)
    for a, b in zip(ks + vs, before):
        assert np.array_equal(a[:n_past], b[:n_past])
        for i, p in enumerate(path):
            assert np.array_equal(a[n_past + i], b[n_past + p])
    o._g("kv_compact")(kv, n_past, po._ptr(np.array([0, 1, 2], np.int32)), 3)   # idempotent once compacted
    for a, b in zip(ks + vs, before):
        for i, p in enumerate(path):
            assert np.array_equal(a[n_past + i], b[n_past + p])


import pytest


@pytest.mark.skipif(not po.have_ref(), reason="oracle/_ref not built")
def test_round_composition_oracle_vs_reference_ops():
    """the same round rules driven by (a) the C restatement and (b) the reference's own ggml-cpu
    ops give the same draft trees, accepted paths and bonus tokens."""
    from oracle_round import CpuStream
    hp = dict(E=256, H=2, F=512, V=1024, L=2)
    tgt = fx.make_model(**hp, ftype="q4_k_m", seed=31)
    head = fx.make_model(**{**hp, "L": 1}, ftype=po.Q4_0, seed=32, is_eagle=True)
    cfg = dict(n_branch=4, n_draft=12, max_depth=5, top_k=10, p_split=0.1, max_nodes=16)
    a, b = CpuStream(po.oracle(), tgt, head, 128), CpuStream(po.ref(), tgt, head, 128)
    prompt = np.random.default_rng(1).integers(3, hp["V"], 12).astype(np.int32)
    a.prefill(prompt); b.prefill(prompt)
    assert a.root_token == b.root_token
    for rnd in range(4):
        ra, rb = a.round(**cfg), b.round(**cfg)
        assert ra["tokens"].tolist() == rb["tokens"].tolist() and ra["parent"].tolist() == rb["parent"].tolist()
        assert ra["path"].tolist() == rb["path"].tolist() and ra["bonus"] == rb["bonus"]
        assert len(ra["tokens"]) >= 2 and a.n_past == b.n_past
