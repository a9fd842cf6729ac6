"""A speculative round composed from oracle (or reference-harness) calls: the rule set of
DESIGN.md §round, mirroring R/examples/speculative/speculative-eagle.cpp:395-656 with the EAGLE
head as the draft model.  TEST INFRASTRUCTURE (checker for eb_session_round)."""
import numpy as np

import fixtures as fx
import pyoracle as po


class CpuStream:
    def __init__(self, path, tgt_fx, head_fx, n_ctx):
        self.p = path
        self.hp = tgt_fx["hparams"]
        self.mt = fx.to_cpu_model(path, tgt_fx)
        self.mh = fx.to_cpu_model(path, head_fx, target=self.mt)
        self._keep = (tgt_fx, head_fx)
        self.n_ctx = n_ctx
        self.kvt = path._g("kv_new")(self.hp["L"], self.hp["E"], n_ctx)
        self.kvh = path._g("kv_new")(head_fx["hparams"]["L"], self.hp["E"], n_ctx)
        self.Lh = head_fx["hparams"]["L"]
        self.n_past = 0
        self.root_token = 0
        self.root_feat = np.zeros(self.hp["E"], np.float32)

    def fwd(self, head, tokens, pos, par_abs_bits, hidden, cell0):
        """par_abs_bits: visibility rows over tree cells [T][n_tree]; cells start at n_past"""
        T = len(tokens)
        vis = np.zeros((T, self.n_ctx), np.uint8)
        vis[:, :self.n_past] = 1
        vis[:, self.n_past:self.n_past + par_abs_bits.shape[1]] = par_abs_bits
        m, kv = (self.mh, self.kvh) if head else (self.mt, self.kvt)
        return self.p.forward(m, kv, tokens, pos, hidden, cell0, self.n_ctx, vis, self.hp["E"], self.hp["V"])

    def prefill(self, prompt):
        P = len(prompt)
        vis = np.tril(np.ones((P, P), np.uint8))
        lg, ft = self.fwd(False, prompt, np.arange(P), vis, None, 0)
        hid = np.zeros((P, self.hp["E"]), np.float32); hid[1:] = ft[:-1]
        self.fwd(True, prompt, np.arange(P), vis, hid, 0)
        self.n_past = P
        self.root_token = int(po.oracle()._g("argmax")(po._ptr(np.ascontiguousarray(lg[-1])), self.hp["V"]))
        self.root_feat = ft[-1].copy()
        return lg, ft

    def round(self, n_branch, n_draft, max_depth, top_k, p_split, max_nodes, fixed_accept=-1, p_min=0.0):
        E = self.hp["E"]
        tokens, parent, depth = [self.root_token], [-1], [0]
        drafting = [False] * n_branch; last_node = [0] * n_branch; row = [0] * n_branch
        drafting[0] = True; n_seq = 1
        node_feat = {}
        batch = [0]                                           # node ids of the next head forward
        done = False
        cand_log = []
        for d in range(max_depth + 1):
            if not batch:
                break
            tv = po.tree_vis(np.array(parent, np.int32))
            toks = np.array([tokens[n] for n in batch], np.int32)
            pos = np.array([self.n_past + depth[n] for n in batch], np.int32)
            hid = np.stack([self.root_feat if parent[n] < 0 else node_feat[parent[n]] for n in batch])
            lg, ft = self.fwd(True, toks, pos, tv[batch], hid, self.n_past + batch[0])
            for r, n in enumerate(batch):
                node_feat[n] = ft[r]
            if d == max_depth or done:
                break
            new, n_seq = expand_level(drafting, last_node, row, n_seq, len(tokens), lambda b: po.top_k(lg[row[b]], top_k), n_branch, n_draft, top_k,
                                      p_split, max_nodes, p_min, cand_log)
            first = len(tokens)
            for i, (tok, par, br) in enumerate(new):
                tokens.append(tok); parent.append(par); depth.append(depth[par] + 1)
                last_node[br] = first + i; row[br] = i
            batch = list(range(first, first + len(new)))
            if not new or not any(drafting) or len(tokens) > n_draft:
                done = True
        T = len(tokens)
        par = np.array(parent, np.int32)
        lg, ft = self.fwd(False, np.array(tokens, np.int32), self.n_past + np.array(depth, np.int32), po.tree_vis(par), None, self.n_past)
        best = lg.argmax(1).astype(np.int32)
        if fixed_accept >= 0:
            path, cur = [0], 0
            while len(path) - 1 < fixed_accept:
                ch = [c for c in range(cur + 1, T) if parent[c] == cur]
                if not ch:
                    break
                cur = ch[0]; path.append(cur)
            path = np.array(path, np.int32); bonus = int(best[cur])
        else:
            path, bonus = po.accept_walk(np.array(tokens, np.int32), par, best)
        for kv in (self.kvt, self.kvh):
            po.oracle()._g("kv_compact")(kv, self.n_past, po._ptr(np.ascontiguousarray(path)), len(path)) if self.p.p == "eo_" else _compact_ref(self, kv, path)
        self.root_feat = ft[path[-1]].copy()
        self.n_past += len(path); self.root_token = bonus
        return dict(tokens=np.array(tokens, np.int32), parent=par, path=path, bonus=bonus, logits=lg, cand=cand_log)


def expand_level(drafting, last_node, row, n_seq, n_tokens, cands_of, n_branch, n_draft, top_k, p_split, max_nodes, p_min=0.0, cand_log=None):
    """One level of the tree expansion (R/examples/speculative/speculative-eagle.cpp:527-600): every drafting branch looks at its candidate
    list (ids, values, probabilities, best first: cands_of(b)), splits off a new branch for each further candidate above p_split while
    branches are left, and every branch of the split takes one candidate as its next token.  Mutates drafting / last_node / row of the new
    branches; -> ([(token, parent node, branch)] in batch order, n_seq).  max_nodes / p_min are this repo's additions (caps that do not bind
    in the reference's configuration; R/common/speculative.cpp:283).  Pinned against the reference driver's own log in
    tests/test_drivers.py::test_tree_expansion_replays_the_reference_drivers_log."""
    new = []
    skip = [False] * n_branch
    for b in range(n_branch):
        if not drafting[b] or skip[b]:
            continue
        ids, vals, pr = cands_of(b)
        sa = [b]; checked = []
        for f in range(1, min(8, top_k)):
            if n_seq < n_branch:
                checked.append(f)                 # a p_split comparison that was actually evaluated
            if n_seq < n_branch and pr[f] > p_split:
                nb = n_seq; n_seq += 1
                drafting[nb] = True; skip[nb] = True; last_node[nb] = last_node[b]; row[nb] = row[b]
                sa.append(nb)
            else:
                break
        if cand_log is not None:
            cand_log.append((np.array(ids).copy(), np.array(vals).copy(), len(sa), checked))
        for i, br in enumerate(sa):
            if n_tokens + len(new) >= max_nodes:
                drafting[br] = False
                continue
            new.append((int(ids[i]), last_node[br], br))
            if n_tokens + len(new) > n_draft:
                drafting[br] = False
        if p_min > 0 and pr[0] < p_min:               # R/common/speculative.cpp:283
            drafting[b] = False
    return new, n_seq


def _compact_ref(stream, kv, path):
    L = stream.hp["L"] if kv == stream.kvt else stream.Lh
    ks, vs = stream.p.kv_arrays(kv, L, stream.hp["E"], stream.n_ctx)
    for a in ks + vs:
        for i, p in enumerate(path):
            a[stream.n_past + i] = a[stream.n_past + p]
