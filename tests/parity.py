"""Shared machinery of the model-level parity tests (test infrastructure).

Bars (BASELINE.json north_star): logits within 1e-2 relative; greedy ids identical, ties within tolerance flagged.

Why a flip detector.  Every GEMM input is quantised to the CPU backend's int8 grid (Q8_K / Q8_0), on the GPU exactly as in
the reference, so all integer dot products are equal as long as the int8 planes are equal.  The f32 glue around them (norms,
residual adds, softmax) is summed in a different ORDER on the GPU; a value that lands within an ulp of a .5 boundary of the grid
can round the other way — a *rounding flip* — after which that row carries a perturbation of one quantisation step (~1e-2 of the
logit range on the small fixtures).  Instead of hiding this under a loose tolerance, the tests DETECT it: the engine's activation
tap (eb_session_set_tap) and the oracle's tap (eo_set_tap / the reference's cb_eval through oracle/ref_round.cpp) deliver the int8
plane of every quantisation site; a row is *clean* iff all of its planes are identical on both sides.  Clean rows must meet the
north-star 1e-2 (they sit near 1e-6); flipped rows are counted, bounded and held to FLIP_REL.
"""
import numpy as np

import pyoracle as po

NORTH_STAR_REL = 1e-2      # BASELINE.json: "logits within 1e-2 relative"
FLIP_REL = 3e-2            # a row that suffered a detected rounding flip (one int8 step somewhere in its history)
TIE_REL = 2e-3             # = the per-op tolerance of tests/test_gpu_ops.py: |top1 - top2| <= TIE_REL * |top1| is a flagged tie
MAX_TIE_FRAC = 0.10        # at most this share of rows may be excused as ties (Gaussian logits: ~2-3 % expected)


def row_rel(a, b):
    """per-row max |a - b| / max |b_row| (b = the reference side)"""
    return np.abs(a - b).max(1) / np.maximum(np.abs(b).max(1), 1e-30)


def site_widths(hp, logits=True):
    E, F, L = hp["E"], hp["F"], hp["L"]
    return ([2 * E] if hp["is_eagle"] else []) + [E, E, E, F] * L + ([E] if logits else [])


def flipped_rows(planes_a, planes_b):
    # (simple form, used where both sides are whole-model reference runs)
    """-> bool [T]: row t has at least one differing int8 value at any site; and the per-site count of differing rows"""
    T = planes_a[0].shape[0]
    bad = np.zeros(T, bool); per_site = []
    for a, b in zip(planes_a, planes_b):
        assert a.shape == b.shape, (a.shape, b.shape)
        d = (a != b).any(1)
        per_site.append(int(d.sum())); bad |= d
    return bad, per_site


def quantize_planes(x, kq):
    """f32 [T][K] -> int8 [T][K] with the REFERENCE's quantiser when oracle/_ref is present (quantize_row_q8_K / q8_0 as the CPU
    mul_mat's from_float calls them), else the C restatement (pinned bit-exact to it in tests/test_oracle.py)"""
    path = po.ref() if po.have_ref() else po.oracle()
    T, K = x.shape
    out = np.empty((T, K), np.int8)
    for t in range(T):
        raw = path.quantize_act("q8_K" if kq else po.Q8_0, x[t])
        if kq:
            out[t] = raw.reshape(-1, po.Q8_K_BYTES)[:, 4:260].reshape(-1).view(np.int8)
        else:
            out[t] = raw.reshape(-1, 34)[:, 2:].reshape(-1).view(np.int8)
    return out


def _is_site(name):
    if name in ("fc.weight", "output.weight"):
        return True
    return name.startswith("blk.") and name.split(".")[2] in ("attn_q", "attn_output", "ffn_up", "ffn_down")


def planes_from_ref_taps(taps, type_of):
    """reference-side planes in the engine's site order from oracle/ref_round.cpp's MUL_MAT taps [(weight name, x [T][K])];
    type_of(name) -> ggml type of that weight"""
    planes = []
    for name, x in taps:
        if not _is_site(name):
            continue
        t = type_of(name)
        planes.append(quantize_planes(x, t in (po.Q4_K, po.Q6_K)))
    return planes


def greedy_report(a, b, tie=TIE_REL):
    """a vs b (b = reference side) -> (mismatches on decided rows, flagged tie rows, rows)"""
    bad = ties = 0
    for ra, rb in zip(a, b):
        top = np.sort(rb)[-2:]
        if top[1] - top[0] <= tie * abs(top[1]):
            ties += 1
        elif int(ra.argmax()) != int(rb.argmax()):
            bad += 1
    return bad, ties, len(a)


def classify_rows(planes_g, planes_o, vis=None, is_eagle=False):
    """-> (dirty [T], primary [T], per-site differing-row counts).  dirty: the row has a differing int8 value at some site.
    primary: the row's FIRST difference is a rounding flip of its own — not explained by attention over an already-dirty row
    (attention mixes rows: once row r differs, layer l+1's K/V of r differ and every row that sees r inherits a perturbation at
    its wo input; that is propagation of r's flip, not a new one).  vis [T][T]: row t sees row r (None: rows are independent)."""
    T = planes_g[0].shape[0]
    dirty = np.zeros(T, bool); primary = np.zeros(T, bool); per_site = []
    off = 1 if is_eagle else 0
    for si, (a, b) in enumerate(zip(planes_g, planes_o)):
        assert a.shape == b.shape, (a.shape, b.shape)
        d = (a != b).any(1)
        per_site.append(int(d.sum()))
        new = d & ~dirty
        attn_site = si >= off and (si - off) % 4 == 1 and si < len(planes_g) - 1
        if attn_site and vis is not None:
            others = np.array(vis, bool).copy(); np.fill_diagonal(others, False)
            contaminated = (others & dirty[None, :]).any(1)
            primary |= new & ~contaminated
        else:
            primary |= new
        dirty |= d
    return dirty, primary, per_site


def check_forward(tag, lg, lo, planes_g=None, planes_o=None, max_flip_frac=0.2, clean_rel=NORTH_STAR_REL, flip_rel=FLIP_REL, vis=None, is_eagle=False):
    """the model-level bar.  Rows whose int8 activation planes are identical on both sides at every site must agree to clean_rel;
    rows with a difference (a detected rounding flip of their own, or one inherited through attention) to flip_rel; the share of
    rows with a PRIMARY flip is bounded by max_flip_frac.  Returns the numbers (also printed: `pytest -s` shows them)."""
    rel = row_rel(lg, lo)
    T = len(rel)
    if planes_g is not None:
        dirty, primary, per_site = classify_rows(planes_g, planes_o, vis, is_eagle)
    else:
        dirty, primary, per_site = np.zeros(T, bool), np.zeros(T, bool), []
    clean = ~dirty
    bad, ties, _ = greedy_report(lg, lo)
    rep = dict(tag=tag, rows=T, dirty=int(dirty.sum()), flipped=int(primary.sum()), clean_max_rel=float(rel[clean].max()) if clean.any() else 0.0,
               flipped_max_rel=float(rel[dirty].max()) if dirty.any() else 0.0, greedy_bad=bad, ties=ties)
    print(f"[parity] {tag}: rows {T}, primary flips {rep['flipped']} ({100.0 * rep['flipped'] / T:.1f} %), rows with differing planes {rep['dirty']}, "
          f"clean max rel {rep['clean_max_rel']:.2e}, differing-row max rel {rep['flipped_max_rel']:.2e}, greedy mismatches {bad}, undecided at {TIE_REL:.0e}: {ties}"
          + (f", differing rows per site {per_site}" if per_site else ""))
    assert rep["clean_max_rel"] <= clean_rel, rep
    assert rep["flipped_max_rel"] <= flip_rel, rep
    assert rep["flipped"] <= max(2, max_flip_frac * T), rep
    # greedy ids: identical on every decided row whose planes are identical; a differing row may legitimately move a near-decision
    bad_clean, _, _ = greedy_report(lg[clean], lo[clean]) if clean.any() else (0, 0, 0)
    assert bad_clean == 0, rep
    return rep


# ---- speculative rounds: every decision of a round has a margin in the ORACLE's own numbers; a GPU / oracle divergence is only
# excusable when the first differing decision was closer than FLIP_REL (what one detected rounding flip can move a logit by)
def round_margin(want, p_split):
    """smallest relative margin (in units of the row's top logit) of any decision the oracle took in this round that shapes the
    result: the order of the draft candidates that were USED (+ the first unused one), the p_split comparisons that were evaluated
    (R/examples/speculative/speculative-eagle.cpp:556-597), the target's argmax at the nodes of the accepted path"""
    m = np.inf
    for ids, vals, n_used, checked in want["cand"]:
        v = np.asarray(vals, np.float64)
        scale = max(abs(v[0]), 1e-30)
        hi = min(n_used + 1, len(v))
        if hi > 1:
            m = min(m, float(np.min(-np.diff(v[:hi])) / scale))
        p = np.exp(v - v[0]); p /= p.sum()
        for f in checked:                                   # ln p moves by at most ~2 * (logit shift): same units as the other margins
            m = min(m, float(abs(np.log(p[f] / p_split)) / (2.0 * scale)))
    rows = want["logits"][np.asarray(want["path"], np.int64)]
    top2 = np.sort(rows, axis=1)[:, -2:]
    m = min(m, float(np.min((top2[:, 1] - top2[:, 0]) / np.maximum(np.abs(top2[:, 1]), 1e-30))))
    return m


def same_round(got, want):
    n = got.n_nodes
    return (n == len(want["tokens"]) and list(got.tokens[:n]) == want["tokens"].tolist() and list(got.parent[:n]) == want["parent"].tolist()
            and list(got.path[:got.n_accept]) == want["path"].tolist() and got.bonus == want["bonus"])
