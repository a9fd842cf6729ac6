"""CPU: the round-level pieces of the oracle pinned to the REFERENCE's own C++ (libllama + common compiled by oracle/Makefile
into oracle/_ref/, driven through oracle/ref_round.cpp): the loader on our GGUF fixtures, llama-quantize's type mixture,
llama_decode on a tree-shaped llama_batch, the KQ mask of llama_set_inputs, the sampler chains (top-k order incl. ties, softmax
over the kept candidates), the chain accept walk common_sampler_sample_and_accept_n and the KV sequence operations + defrag."""
import os
import subprocess

import numpy as np
import pytest

import fixtures as fx
import gguf_fixture as gf
import pyoracle as po
import refround as rr

pytestmark = pytest.mark.skipif(not rr.have(), reason="oracle/_ref/libref_round.so not built (needs /root/reference at build time)")

HP = dict(E=256, H=2, F=512, V=1024, L=2)


@pytest.fixture(scope="module")
def tiny(tmp_path_factory):
    d = tmp_path_factory.mktemp("gguf")
    tgt = fx.make_model(**HP, ftype="q4_k_m", seed=31)
    path = gf.write_gguf(tgt, str(d / "tgt_q4_k_m.gguf"))
    m = rr.Model(path)
    yield tgt, m, d
    m.free()


def test_gguf_fixture_loads_in_reference_loader(tiny, tmp_path):
    tgt, m, _ = tiny
    assert m.V == HP["V"]
    assert m.tensor_type("token_embd.weight") == po.Q4_K and m.tensor_type("output.weight") == po.Q6_K
    for ftype in (po.Q4_0, po.Q8_0):
        fxm = fx.make_model(**HP, ftype=ftype, seed=5)
        mm = rr.Model(gf.write_gguf(fxm, str(tmp_path / f"t{ftype}.gguf")))
        assert mm.tensor_type("blk.1.ffn_down.weight") == ftype
        mm.free()
    head = fx.make_model(**{**HP, "L": 1}, ftype=po.Q4_0, seed=32, is_eagle=True)
    mh = rr.Model(gf.write_gguf(head, str(tmp_path / "head.gguf")))     # arch `eagle`: fc.weight / fc.bias, no attn_norm / output_norm
    assert mh.tensor_type("fc.weight") == po.Q4_0 and mh.tensor_type("fc.bias") == po.F32 and mh.tensor_type("blk.0.attn_norm.weight") == -1
    mh.free()


def test_llama_quantize_q4_k_m_mixture(tmp_path):
    """the reference's own llama-quantize on an f16 fixture: Q6_K for output and for attn_v / ffn_down in the use_more_bits layers
    (R/src/llama-quant.cpp:129-338) — the mixture fixtures.q4_k_m_type and the bench's synthetic 7B assume."""
    L = 8
    src = fx.make_model(**{**HP, "L": L}, ftype=po.F16, seed=9)
    f16 = gf.write_gguf(src, str(tmp_path / "f16.gguf"))
    out = str(tmp_path / "q4km.gguf")
    exe = os.path.join(rr.REF, "llama-quantize")
    r = subprocess.run([exe, f16, out, "Q4_K_M", "2"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    m = rr.Model(out)
    assert m.tensor_type("output.weight") == po.Q6_K and m.tensor_type("token_embd.weight") == po.Q4_K
    for l in range(L):
        for which, name in ((po.L_WV, "attn_v"), (po.L_DOWN, "ffn_down"), (po.L_WQ, "attn_q"), (po.L_GATE, "ffn_gate")):
            assert m.tensor_type(f"blk.{l}.{name}.weight") == fx.q4_k_m_type(which, l, L), (l, name)
    m.free()


def _tree_case(rng, T, n_prompt):
    par = fx.tree_parents(T, rng); depth = fx.tree_depths(par)
    seqs, n_seq = fx.tree_seq_ids(par)
    prompt = rng.integers(3, HP["V"], n_prompt).astype(np.int32)
    tokens = rng.integers(3, HP["V"], T).astype(np.int32)
    return par, depth, seqs, n_seq, prompt, tokens


def test_llama_decode_tree_batch_equals_graph_harness_and_oracle(tiny):
    """llama_decode (R/src/llama.cpp:11916) on a tree-shaped llama_batch == the same forward through oracle/ref_harness.cpp (the
    ggml graphs the golden vectors were generated with): bit for bit; the C restatement follows within the op tolerance."""
    tgt, m, _ = tiny
    rng = np.random.default_rng(4)
    T, n_prompt, n_ctx = 19, 9, 256
    par, depth, seqs, n_seq, prompt, tokens = _tree_case(rng, T, n_prompt)
    c = rr.Context(m, n_ctx, n_seq_max=16)
    c.decode(prompt, np.arange(n_prompt))
    for s in range(1, n_seq):
        c.seq_cp(0, s)                                   # the prefix belongs to every branch (speculative-eagle.cpp:648-652)
    lg = c.decode(tokens, n_prompt + depth, seqs)
    for path, tol in ((po.ref(), 0.0), (po.oracle(), 3e-2)):
        mm = fx.to_cpu_model(path, tgt)
        kv = path._g("kv_new")(HP["L"], HP["E"], n_ctx)
        path.forward(mm, kv, prompt, np.arange(n_prompt), None, 0, n_ctx, fx.vis_matrix(np.arange(n_prompt) - 1, 0, n_ctx), HP["E"], HP["V"])
        l2, _ = path.forward(mm, kv, tokens, n_prompt + depth, None, n_prompt, n_ctx, fx.vis_matrix(par, n_prompt, n_ctx), HP["E"], HP["V"])
        err = float(np.abs(l2 - lg).max() / np.abs(lg).max())
        print(f"{path.kind}: logits vs llama_decode rel {err:.3g}")
        assert err <= tol
    c.free()


def test_kq_mask_of_llama_set_inputs(tiny):
    """the mask libllama builds for the tree batch (R/src/llama-context.cpp:136-201) == the oracle's rule == the product's compact
    128-bit row sets (eb_tree_bits_from_batch) == prefix + ancestors-or-self"""
    from conftest import load_package
    tgt, m, _ = tiny
    rng = np.random.default_rng(11)
    for T, n_prompt in ((7, 5), (27, 12), (65, 30)):
        par, depth, seqs, n_seq, prompt, tokens = _tree_case(rng, T, n_prompt)
        c = rr.Context(m, 256, n_seq_max=max(8, n_seq))
        c.decode(prompt, np.arange(n_prompt))
        for s in range(1, n_seq):
            c.seq_cp(0, s)
        c.decode(tokens, n_prompt + depth, seqs, want_logits=False)
        mask = c.kq_mask()                               # [pad(T, 64)][n_kv], 0 = visible, -inf = masked
        n_kv = mask.shape[1]
        assert mask.shape[0] % 64 == 0 and mask.shape[0] >= T and n_kv % 256 == 0
        vis = (mask[:T] == 0).astype(np.uint8)
        assert np.isinf(mask[T:]).all() and ((mask == 0) | np.isinf(mask)).all()
        want = fx.vis_matrix(par, n_prompt, n_kv)
        assert np.array_equal(vis, want)
        cell_pos, cell_seq = c.kv_cells()
        q_seq0 = np.array([s[0] for s in seqs], np.int32)
        assert np.array_equal(po.build_kq_mask(cell_pos[:n_kv], cell_seq[:n_kv], n_prompt + depth, q_seq0), want)
        bits = load_package().tree_bits_from_batch(n_prompt + depth, seqs)
        tv = np.array([[(bits[j, i >> 5] >> (i & 31)) & 1 for i in range(T)] for j in range(T)], np.uint8)
        assert np.array_equal(tv, want[:, n_prompt:n_prompt + T])
        c.free()


def _logit_cases(rng, V, rows):
    x = rng.normal(0, 2, (rows, V)).astype(np.float32)
    x[1, rng.integers(0, V, 40)] = 3.25                  # heavy ties at the top
    x[2] = np.round(x[2] * 2) / 2                        # coarse grid: ties everywhere
    x[3, :] = 0.0; x[3, [V - 1, 5, 77]] = 1.0           # three-way tie, the rest equal
    x[4] = np.float32(-1.5)                              # all equal
    return x


def test_top_k_order_and_probs_vs_reference_sampler(tiny):
    """eo_top_k == the draft sampler chain of common_speculative_init run by the reference (top-k 10 via llama_sampler_top_k_impl =
    std::partial_sort on logit >, then softmax over the kept candidates), including rows full of ties"""
    tgt, m, _ = tiny
    rng = np.random.default_rng(21)
    rows = 8
    c = rr.Context(m, 64)
    c.decode(rng.integers(3, HP["V"], rows).astype(np.int32), np.arange(rows))
    x = _logit_cases(rng, HP["V"], rows)
    c.set_logits(x)
    n_tied = 0
    for r in range(rows):
        for k in (1, 10, 16):
            ids, lg, pr = c.draft_candidates(r, k)
            oi, ov, op_ = po.top_k(x[r], k)
            assert np.array_equal(lg, ov), (r, k)                        # the kept VALUES always agree
            assert np.allclose(pr, op_, rtol=1e-6, atol=1e-7)
            # ids: std::partial_sort is unstable, so inside a run of equal logits the reference's order is libstdc++'s heap order; the
            # oracle (and the GPU kernel) take the lowest ids.  Wherever the kept logits are distinct the ids must be identical.
            distinct = np.concatenate([[True], np.diff(lg) != 0]) & np.concatenate([np.diff(lg) != 0, [x[r][x[r] <= lg[-1]].size == 0 or np.sort(x[r])[::-1][min(k, HP["V"] - 1)] != lg[-1]]])
            assert np.array_equal(ids[distinct], oi[distinct]), (r, k)
            n_tied += int((~distinct).sum())
            assert np.array_equal(x[r][ids], lg)                          # the reference's ids do carry those logits
    print(f"top-k: {n_tied} tied slots compared by value only")
    c.free()


def test_accept_walk_chain_vs_reference(tiny):
    """eo_accept_walk on chain-shaped trees == common_sampler_sample_and_accept_n (R/common/sampling.cpp:423-458) with the greedy
    recipe, over logits whose argmax agrees with the draft for a prefix of every length"""
    tgt, m, _ = tiny
    rng = np.random.default_rng(33)
    n_draft = 7
    c = rr.Context(m, 64)
    c.decode(rng.integers(3, HP["V"], n_draft + 1).astype(np.int32), np.arange(n_draft + 1))
    for agree in range(n_draft + 1):
        x = rng.normal(0, 1, (n_draft + 1, HP["V"])).astype(np.float32)
        draft = rng.integers(3, HP["V"], n_draft).astype(np.int32)
        for i in range(agree):
            x[i, draft[i]] = 9.0                         # the target agrees with draft[i] after row i
        best = x.argmax(1).astype(np.int32)
        if agree < n_draft:
            assert best[agree] != draft[agree]
        c.set_logits(x)
        ref_ids = c.accept_n(draft)
        tokens = np.concatenate([[1], draft]).astype(np.int32)           # chain tree: node 0 = id_last, node i = draft[i-1]
        path, bonus = po.accept_walk(tokens, np.arange(n_draft + 1) - 1, best)
        ours = [int(tokens[p]) for p in path[1:]] + [bonus]              # accepted draft tokens + the target's own next token
        assert ref_ids.tolist() == ours and len(ref_ids) == agree + 1
    c.free()


def test_kv_compaction_vs_reference_seq_ops_and_defrag(tiny):
    """eo_kv_compact (accepted cells moved to the committed prefix) == what the reference ends up with after the tree driver's
    sequence operations (R/examples/speculative/speculative-eagle.cpp:459-472 -> R/src/llama-kv-cache.cpp:368,431,478) followed by a
    forced defrag (R/src/llama.cpp:10941): the same position -> K/V row map, nothing else left in the cache"""
    tgt, m, _ = tiny
    rng = np.random.default_rng(8)
    T, n_prompt, n_ctx, E, L = 23, 10, 256, HP["E"], HP["L"]
    o = po.oracle()
    for trial in range(4):
        par, depth, seqs, n_seq, prompt, tokens = _tree_case(rng, T, n_prompt)
        c = rr.Context(m, n_ctx, n_seq_max=max(8, n_seq))
        c.decode(prompt, np.arange(n_prompt))
        for s in range(1, n_seq):
            c.seq_cp(0, s)
        c.decode(tokens, n_prompt + depth, seqs, want_logits=False)
        rows_before = [[[c.kv_row(l, w, cell, E) for cell in range(n_prompt + T)] for w in (0, 1)] for l in range(L)]
        # accepted path: walk down from the root along random children
        path = [0]
        while True:
            ch = [j for j in range(T) if par[j] == path[-1]]
            if not ch or rng.random() < 0.25:
                break
            path.append(int(rng.choice(ch)))
        n_acc = len(path)
        s_keep = seqs[path[-1]][0]                       # the branch of the last accepted node
        n_past_new = n_prompt + n_acc
        c.seq_rm(s_keep, n_past_new, -1); c.seq_keep(s_keep); c.seq_cp(s_keep, 0); c.seq_keep(0)
        c.defrag()
        pos, seq = c.kv_cells()
        live = np.nonzero(pos >= 0)[0]
        assert sorted(pos[live].tolist()) == list(range(n_past_new)) and (seq[live] == 1).all()
        # oracle: the same rows in an eo_kv, compacted
        kv = o._g("kv_new")(L, E, n_ctx)
        ks, vs = o.kv_arrays(kv, L, E, n_ctx)
        for l in range(L):
            for cell in range(n_prompt + T):
                ks[l][cell] = rows_before[l][0][cell]; vs[l][cell] = rows_before[l][1][cell]
        pa = np.array(path, np.int32)
        o._g("kv_compact")(kv, n_prompt, po._ptr(pa), n_acc)
        for cell in live:
            p = int(pos[cell])
            for l in range(L):
                assert np.array_equal(c.kv_row(l, 0, int(cell), E), ks[l][p]) and np.array_equal(c.kv_row(l, 1, int(cell), E), vs[l][p])
        o._g("kv_free")(kv)
        c.free()
