"""Synthetic model fixtures shared by the CPU and GPU tests (random-init weights, fixed seeds).

Weights are N(0, 0.02^2) (norm weights 1 + N(0, 0.02^2)) quantised with the reference's own
quantiser when oracle/_ref is present (ggml_quantize_chunk, what llama-quantize calls); without
it, well-formed random blocks are generated instead — any block bytes are a valid tensor for
both the oracle and the GPU path.
"""
import numpy as np

import pyoracle as po

QTYPES = (po.Q4_0, po.Q8_0, po.Q4_K, po.Q6_K)


def random_blocks(t, n_elems, rng, scale=0.02):
    be, bb = po.BLOCK[t]
    nb = n_elems // be
    raw = rng.integers(0, 256, (nb, bb), dtype=np.uint8)
    d = (scale * (2.2 + 0.6 * rng.random(nb))).astype(np.float32)
    if t == po.Q4_K:
        # centred sub-blocks, as a quantiser produces them for zero-mean weights: value = d*sc*q - dmin*m with dmin = 8 d and
        # m = round(7.5 sc / 8), i.e. ~ d*sc*(q - 7.5).  (Independent random sc and m give every 32-weight sub-block its own large
        # offset; a 32-layer model of such matrices amplifies an f32 rounding difference ~4x per layer — chaotic, useless for parity.)
        sc = rng.integers(24, 64, (nb, 8), dtype=np.uint8)
        m = np.rint(sc * (7.5 / 8)).astype(np.uint8)
        dd = (d / (100 * 4.6)).astype(np.float16)          # dequantised std ~ 0.022, like the other formats
        raw[:, 0:2] = dd.view(np.uint8).reshape(nb, 2)
        raw[:, 2:4] = (dd.astype(np.float32) * 8).astype(np.float16).view(np.uint8).reshape(nb, 2)
        # 6-bit packing of get_scale_min_k4 (R/ggml/src/ggml-quants.c:631): j < 4: sc = q[j] & 63, m = q[j+4] & 63;
        # j >= 4: sc = (q[j+4] & 15) | (q[j-4] >> 6) << 4, m = (q[j+4] >> 4) | (q[j] >> 6) << 4
        q = np.zeros((nb, 12), np.uint8)
        q[:, 0:4] = (sc[:, 0:4] & 63) | ((sc[:, 4:8] >> 4) << 6)
        q[:, 4:8] = (m[:, 0:4] & 63) | ((m[:, 4:8] >> 4) << 6)
        q[:, 8:12] = (sc[:, 4:8] & 15) | ((m[:, 4:8] & 15) << 4)
        raw[:, 4:16] = q
    elif t == po.Q6_K:
        raw[:, 208:210] = (d / (32 * 96)).astype(np.float16).view(np.uint8).reshape(nb, 2)
    elif t == po.Q4_0:
        raw[:, 0:2] = (d / 8).astype(np.float16).view(np.uint8).reshape(nb, 2)
    elif t == po.Q8_0:
        raw[:, 0:2] = (d / 127).astype(np.float16).view(np.uint8).reshape(nb, 2)
    return raw.reshape(-1)


def quantize(t, w, rng=None):
    """f32 [rows][cols] -> GGUF block bytes"""
    w = np.ascontiguousarray(w, np.float32)
    if t == po.F32:
        return w.view(np.uint8).reshape(-1).copy()
    if t == po.F16:
        return w.astype(np.float16).view(np.uint8).reshape(-1).copy()
    rows, cols = w.shape
    if po.have_ref():
        raw = np.empty(rows * po.row_size(t, cols), np.uint8)
        po.ref()._g("quantize")(t, po._ptr(w), po._ptr(raw), rows, cols)
        return raw
    return random_blocks(t, rows * cols, rng or np.random.default_rng(0))


def q4_k_m_type(which, layer, n_layer):
    """llama-quantize's Q4_K_M mixture (R/src/llama-quant.cpp:129-338): Q6_K for `output` and for
    attn_v / ffn_down in the use_more_bits layers, Q4_K elsewhere."""
    more = layer < n_layer // 8 or layer >= 7 * n_layer // 8 or (layer - n_layer // 8) % 3 == 2
    if which in (po.L_WV, po.L_DOWN) and more:
        return po.Q6_K
    return po.Q4_K


def make_model(E, H, F, V, L, ftype, seed, is_eagle=False, std=0.02):
    """-> dict: {'hparams':..., 'globals': {which: (type, raw)}, 'layers': [{which: (type, raw)}]}
    ftype: one of Q4_0 / Q8_0 / 'q4_k_m'"""
    rng = np.random.default_rng(seed)

    def mat(rows, cols, t):
        return (t, quantize(t, rng.normal(0, std, (rows, cols)).astype(np.float32), rng))

    def vec(n):
        return (po.F32, (1 + rng.normal(0, std, n)).astype(np.float32).view(np.uint8).reshape(-1).copy())

    def lt(which, l):
        if ftype == "q4_k_m":
            return q4_k_m_type(which, l, L)
        return ftype

    base = po.Q4_K if ftype == "q4_k_m" else ftype
    out_t = po.Q6_K if ftype in ("q4_k_m", po.Q4_0) else ftype      # `output` is Q6_K for Q4_0 and Q4_K_M files
    g = {po.T_TOK_EMBD: mat(V, E, base)}
    if is_eagle:
        g[po.T_FC] = mat(E, 2 * E, base)
        g[po.T_FC_BIAS] = (po.F32, rng.normal(0, std, E).astype(np.float32).view(np.uint8).reshape(-1).copy())
    else:
        g[po.T_OUTPUT_NORM] = vec(E)
        g[po.T_OUTPUT] = mat(V, E, out_t)
    layers = []
    for l in range(L):
        d = {}
        if not is_eagle:
            d[po.L_ATTN_NORM] = vec(E)
        d[po.L_WQ] = mat(E, E, lt(po.L_WQ, l)); d[po.L_WK] = mat(E, E, lt(po.L_WK, l)); d[po.L_WV] = mat(E, E, lt(po.L_WV, l))
        d[po.L_WO] = mat(E, E, lt(po.L_WO, l)); d[po.L_FFN_NORM] = vec(E)
        d[po.L_GATE] = mat(F, E, lt(po.L_GATE, l)); d[po.L_UP] = mat(F, E, lt(po.L_UP, l)); d[po.L_DOWN] = mat(E, F, lt(po.L_DOWN, l))
        layers.append(d)
    return {"hparams": dict(E=E, H=H, F=F, V=V, L=L, eps=1e-5, freq_base=10000.0, is_eagle=is_eagle), "globals": g, "layers": layers}


def to_cpu_model(path, fx, target=None):
    """instantiate the fixture in a CPU path (oracle or ref); keeps the raw buffers alive"""
    hp = fx["hparams"]
    m = path.model_new(hp["E"], hp["H"], hp["F"], hp["V"], hp["L"], hp["eps"], hp["freq_base"], hp["is_eagle"])
    for which, (t, raw) in fx["globals"].items():
        path._g("model_set_global")(m, which, t, po._ptr(raw))
    for l, d in enumerate(fx["layers"]):
        for which, (t, raw) in d.items():
            path._g("model_set_layer")(m, l, which, t, po._ptr(raw))
    if target is not None:
        path._g("model_set_lm_head_from")(m, target)
    return m


def to_gpu_model(eb, fx, target=None):
    hp = fx["hparams"]
    m = eb.Model(hp["E"], hp["H"], hp["F"], hp["V"], hp["L"], hp["eps"], hp["freq_base"], hp["is_eagle"])
    for which, (t, raw) in fx["globals"].items():
        m.set_tensor(-1, which, t, raw)
    for l, d in enumerate(fx["layers"]):
        for which, (t, raw) in d.items():
            m.set_tensor(l, which, t, raw)
    if target is not None:
        m.share_lm_head(target)
    return m


def tree_parents(T, rng, n_branch=4):
    """random tree in creation order: parent[i] < i"""
    par = np.full(T, -1, np.int32)
    for i in range(1, T):
        par[i] = rng.integers(max(0, i - n_branch), i)
    return par


def driver_like_tree(T, n_branch, rng):
    """a verify tree shaped like the reference tree driver's (-np n_branch): the root, branches that split off within the first
    levels (p_split), then every branch extends as a chain, one node per branch per depth level, until T nodes exist"""
    par = [-1]; tips = [0]
    while len(par) < T:
        nxt = []
        n_now = len(tips)                                  # branches alive at this depth (+ the ones opened below)
        for tip in tips:
            if len(par) >= T:
                break
            par.append(tip); nxt.append(len(par) - 1)
            if n_now < n_branch and len(par) < T and rng.random() < 0.5:
                par.append(tip); nxt.append(len(par) - 1); n_now += 1
        tips = nxt
    return np.array(par[:T], np.int32)


def tree_seq_ids(par):
    """seq-id sets as the reference's tree driver assigns them (R/examples/speculative/speculative-eagle.cpp:556-625): one branch
    per root-to-leaf path; a node's FIRST id is the branch that created it, followed by the branches that split off below it."""
    T = len(par)
    children = [[j for j in range(T) if par[j] == i] for i in range(T)]
    seqs = [[] for _ in range(T)]; nxt = [1]

    def walk(i, sid):
        if not seqs[i]:
            seqs[i].insert(0, sid)
        for n, ch in enumerate(children[i]):
            s = sid if n == 0 else nxt[0]
            if n:
                nxt[0] += 1
            walk(ch, s)
            seqs[i].extend(x for x in seqs[ch] if x not in seqs[i])
    walk(0, 0)
    return seqs, nxt[0]


def tree_depths(par):
    d = np.zeros(len(par), np.int32)
    for i in range(1, len(par)):
        d[i] = d[par[i]] + 1
    return d


def vis_matrix(par, n_past, n_kv):
    """[T][n_kv] visibility: committed prefix + ancestors-or-self (cells n_past + node)"""
    tv = po.tree_vis(par)
    T = len(par)
    vis = np.zeros((T, n_kv), np.uint8)
    vis[:, :n_past] = 1
    vis[:, n_past:n_past + T] = tv
    return vis
