"""SURVEY 8f N3 — GGUF file -> resident model (csrc/gguf_load.cu).  The files are written by tests/gguf_fixture.py with the reference's
vendored gguf-py and are the same ones the reference's own loader accepts (tests/test_ref_round.py, tests/test_dropin.py).
CPU: the header parser (eb_gguf_probe needs no device) on llama / eagle files with and without a vocabulary section, and its error paths.
GPU: eb_model_load_gguf streams the data section into HBM; every tensor read back is bit-identical to the fixture's GGUF bytes (plane-layout
types included) and a forward through the loaded models equals the forward through models filled tensor by tensor."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402

import fixtures as fx  # noqa: E402
import gguf_fixture as gf  # noqa: E402
import pyoracle as po  # noqa: E402

HP = dict(E=256, H=2, F=512, V=1024, L=2)


def _files(tmp_path, vocab):
    tgt = fx.make_model(**HP, ftype="q4_k_m", seed=81)
    head = fx.make_model(**{**HP, "L": 1}, ftype=po.Q4_0, seed=82, is_eagle=True)
    return tgt, head, gf.write_gguf(tgt, str(tmp_path / "t.gguf"), vocab=vocab), gf.write_gguf(head, str(tmp_path / "h.gguf"), vocab=vocab)


@pytest.mark.parametrize("vocab", [False, True])
def test_gguf_probe_reads_the_reference_format(tmp_path, vocab):
    eb = load_package()
    tgt, head, pt, ph = _files(tmp_path, vocab)
    a, b = eb.gguf_probe(pt), eb.gguf_probe(ph)
    assert (a["E"], a["H"], a["F"], a["V"], a["L"], a["is_eagle"]) == (256, 2, 512, 1024, 2, False)
    assert (b["E"], b["H"], b["F"], b["V"], b["L"], b["is_eagle"]) == (256, 2, 512, 1024, 1, True)
    assert abs(a["eps"] - 1e-5) < 1e-12 and a["freq_base"] == 10000.0
    n_t = len(tgt["globals"]) + sum(len(d) for d in tgt["layers"]); n_h = len(head["globals"]) + sum(len(d) for d in head["layers"])
    assert a["n_tensors"] == n_t and b["n_tensors"] == n_h
    assert a["bytes"] == sum(r.nbytes for _, r in tgt["globals"].values()) + sum(r.nbytes for d in tgt["layers"] for _, r in d.values())


def test_gguf_probe_rejects_what_it_cannot_load(tmp_path):
    eb = load_package()
    _, _, pt, _ = _files(tmp_path, False)
    raw = open(pt, "rb").read()
    bad = tmp_path / "bad.gguf"
    bad.write_bytes(b"GGML" + raw[4:])
    with pytest.raises(eb.EagleB200Error, match="magic"):
        eb.gguf_probe(str(bad))
    bad.write_bytes(raw[:len(raw) // 2])                      # truncated data section
    with pytest.raises(eb.EagleB200Error, match="past the end"):
        eb.gguf_probe(str(bad))
    bad.write_bytes(raw[:64])                                 # truncated header
    with pytest.raises(eb.EagleB200Error):
        eb.gguf_probe(str(bad))
    with pytest.raises(eb.EagleB200Error, match="cannot open"):
        eb.gguf_probe(str(tmp_path / "missing.gguf"))


def test_gguf_probe_survives_truncation_and_corruption(tmp_path):
    """every prefix of a valid file and a few hundred single-byte corruptions of its header must end in an error code or a clean answer,
    never in a crash (the parser bounds every read: lengths, counts and offsets come from the file)"""
    eb = load_package()
    _, _, pt, _ = _files(tmp_path, True)
    raw = open(pt, "rb").read()
    hdr = raw.index(b"token_embd.weight") + 4096 if b"token_embd.weight" in raw else 8192
    hdr = min(hdr, len(raw))
    f = tmp_path / "cut.gguf"
    outcomes = {"error": 0, "ok": 0}
    for cut in list(range(0, 200, 7)) + list(range(200, hdr, 97)) + [len(raw) - 1]:
        f.write_bytes(raw[:cut])
        try:
            eb.gguf_probe(str(f)); outcomes["ok"] += 1
        except eb.EagleB200Error:
            outcomes["error"] += 1
    assert outcomes["error"] > 0 and outcomes["ok"] == 0                  # no prefix of the file is a whole file
    rng = np.random.default_rng(0)
    for _ in range(300):
        b = bytearray(raw[:hdr + 64])
        i = int(rng.integers(4, hdr)); b[i] = int(rng.integers(0, 256))
        f.write_bytes(bytes(b) + raw[hdr + 64:])
        try:
            eb.gguf_probe(str(f)); outcomes["ok"] += 1
        except eb.EagleB200Error:
            outcomes["error"] += 1
    assert outcomes["ok"] + outcomes["error"] >= 300


@pytest.mark.gpu
def test_load_gguf_bit_exact_and_same_forward(eb, tmp_path):
    tgt, head, pt, ph = _files(tmp_path, True)
    mt, mh = eb.Model.from_gguf(pt), eb.Model.from_gguf(ph)
    assert (mt.E, mt.H, mt.F, mt.V, mt.L, mt.is_eagle) == (256, 2, 512, 1024, 2, False) and mh.is_eagle and mh.L == 1
    assert mt.load_bytes == eb.gguf_probe(pt)["bytes"] and mt.load_seconds > 0
    for m, f in ((mt, tgt), (mh, head)):
        for which, (t, raw) in f["globals"].items():
            np.testing.assert_array_equal(m.get_tensor(-1, which, raw.nbytes), raw.view(np.uint8).reshape(-1))
        for l, d in enumerate(f["layers"]):
            for which, (t, raw) in d.items():
                np.testing.assert_array_equal(m.get_tensor(l, which, raw.nbytes), raw.view(np.uint8).reshape(-1))
    mh.share_lm_head(mt)
    rt = fx.to_gpu_model(eb, tgt); rh = fx.to_gpu_model(eb, head, target=rt)
    rng = np.random.default_rng(3)
    prompt = rng.integers(3, HP["V"], 12).astype(np.int32)
    outs = []
    for a, b in ((mt, mh), (rt, rh)):
        s = eb.Session(a, b, n_ctx=128, max_nodes=17, use_graph=False)
        lg, ft = s.forward(0, prompt, np.arange(12))
        hid = np.zeros((12, HP["E"]), np.float32); hid[1:] = ft[:-1]
        lh, _ = s.forward(1, prompt, np.arange(12), hidden=hid)
        outs.append((lg, lh)); s.free()
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    for m in (mt, mh, rt, rh):
        m.free()


def _llama_quantize(tmp_path, L=4, seed=83):
    """an f16 fixture through the reference's OWN llama-quantize (C++ gguf writer, Q4_K_M mixture): the file a user of the reference has"""
    import subprocess
    import refround as rr
    exe = os.path.join(rr.REF, "llama-quantize")
    if not os.path.exists(exe):
        pytest.fail("oracle/_ref/llama-quantize is missing: run __graft_entry__.build() where /root/reference is present")
    src = fx.make_model(**{**HP, "L": L}, ftype=po.F16, seed=seed)
    f16 = gf.write_gguf(src, str(tmp_path / "f16.gguf"))
    out = str(tmp_path / "q4km.gguf")
    r = subprocess.run([exe, f16, out, "Q4_K_M", "2"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return src, out


def test_gguf_probe_reads_llama_quantize_output(tmp_path):
    """the header parser on a file written by the reference's C++ gguf writer (R/ggml/src/gguf.cpp via llama-quantize), not by gguf-py"""
    import refround as rr
    eb = load_package()
    L = 4
    src, out = _llama_quantize(tmp_path, L)
    p = eb.gguf_probe(out)
    assert (p["E"], p["H"], p["F"], p["V"], p["L"], p["is_eagle"]) == (256, 2, 512, 1024, L, False)
    assert p["n_tensors"] == len(src["globals"]) + sum(len(d) for d in src["layers"])
    m = rr.Model(out)
    want = 0
    for which, name in gf.GLOBAL_NAMES.items():
        if which in src["globals"]:
            shape = gf._shape(src["hparams"], -1, which)
            want += int(np.prod(shape[:-1], dtype=np.int64)) * po.row_size(m.tensor_type(name), shape[-1])
    for l in range(L):
        for which, name in gf.LAYER_NAMES.items():
            shape = gf._shape(src["hparams"], l, which)
            want += int(np.prod(shape[:-1], dtype=np.int64)) * po.row_size(m.tensor_type(f"blk.{l}.{name}.weight"), shape[-1])
    m.free()
    assert p["bytes"] == want, (p["bytes"], want)


@pytest.mark.gpu
def test_llama_quantize_output_through_own_loader_matches_the_reference_decode(eb, tmp_path):
    """file on disk written by the reference's own tool -> eb_model_load_gguf -> engine forward, against the reference's loader ->
    llama_decode on ggml-cpu of the same file: nothing of this repo's fixture code between the file and either side"""
    import parity as pa
    import refround as rr
    from golden_io import nmse
    _, out = _llama_quantize(tmp_path, 4)
    rng = np.random.default_rng(11)
    prompt = rng.integers(3, HP["V"], 24).astype(np.int32)
    mt = eb.Model.from_gguf(out)
    s = eb.Session(mt, n_ctx=128, max_nodes=17, use_graph=False)
    lg, _ = s.forward(0, prompt, np.arange(24))
    s.free(); mt.free()
    m = rr.Model(out)
    c = rr.Context(m, 128)
    lr = c.decode(prompt, np.arange(24))
    c.free(); m.free()
    rel = pa.row_rel(lg, lr)
    agree = int((lg.argmax(1) == lr.argmax(1)).sum())
    print(f"[gguf] llama-quantize file, own loader + engine vs reference loader + llama_decode: max rel {rel.max():.2e}, NMSE {nmse(lg, lr):.2e}, greedy ids {agree}/24")
    assert nmse(lg, lr) <= 2e-3 and rel.max() <= pa.FLIP_REL
