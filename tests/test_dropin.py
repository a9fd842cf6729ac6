"""GPU: the drop-in boundary met by the REFERENCE's own binaries and library (built by oracle/Makefile into oracle/_ref/):

* `test-backend-ops` (R/tests/test-backend-ops.cpp) — the reference's op-parity harness — run unmodified against
  libggml-b200.so loaded through GGML_BACKEND_PATH (R/ggml/src/ggml-backend-reg.cpp:577-581): every op case the plug-in claims in
  supports_op is compared with ggml-cpu at the harness's own NMSE bars (MUL_MAT / FLASH_ATTN_EXT 5e-4, others 1e-7);
* `llama_decode` (R/src/llama.cpp:11916 -> llama_decode_impl :9486) of the reference's libllama on a tree-shaped llama_batch, once on
  ggml-cpu and once with every layer offloaded to the plug-in (weights uploaded by the reference's loader through
  buffer_i::set_tensor, graph split by the reference's ggml_backend_sched), on GGUF fixtures written by tests/gguf_fixture.py;
* a prompt-sized batch (T = 200 > one verify tile) through the same path (§8f N1: prefill through the plug-in).
"""
import os
import re
import subprocess

import numpy as np
import pytest

import fixtures as fx
import gguf_fixture as gf
import parity as pa
import pyoracle as po
import refround as rr
from golden_io import nmse

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "eagle-in-llama.cpp_b200", "ggml_backend", "libggml-b200.so")
OPS = ["MUL_MAT", "FLASH_ATTN_EXT", "RMS_NORM", "ROPE", "GET_ROWS", "CONCAT", "CPY", "CONT", "DUP", "SOFT_MAX", "ARGMAX", "ARGSORT", "ADD", "MUL", "SCALE", "SILU", "RELU"]


def _need_ref():
    if not rr.have() or not os.path.exists(os.path.join(rr.REF, "test-backend-ops")):
        pytest.fail("oracle/_ref (the reference build) is missing: run __graft_entry__.build() where /root/reference is present")
    assert os.path.exists(PLUGIN), PLUGIN


@pytest.mark.parametrize("op", OPS)
def test_backend_ops_reference_harness(eb, op):
    """test-backend-ops -b B2000 -o <op>: must report OK, and must actually have RUN cases on the plug-in for the hot-path ops"""
    _need_ref()
    env = dict(os.environ, GGML_BACKEND_PATH=PLUGIN)
    r = subprocess.run([os.path.join(rr.REF, "test-backend-ops"), "test", "-b", "B2000", "-o", op], env=env, capture_output=True, text=True, timeout=900)
    out = r.stdout + r.stderr
    clean = re.sub(r"\x1b\[[0-9;]*m", "", out)
    m = re.search(r"(\d+)/(\d+) tests passed", clean)
    ran = len(re.findall(r": OK", clean)); unsupported = clean.count("not supported")
    fails = [l for l in clean.splitlines() if "FAIL" in l or "ERR =" in l or "NMSE =" in l]
    print(f"[dropin] test-backend-ops {op}: {m.group(0) if m else 'no summary'}; ran on plug-in {ran}, not supported {unsupported}")
    assert r.returncode == 0 and m and m.group(1) == m.group(2), "\n".join(fails[:20]) + "\n" + clean[-1500:]
    assert "Backend B200: OK" in clean
    if op in ("MUL_MAT", "FLASH_ATTN_EXT", "RMS_NORM", "ROPE", "GET_ROWS", "CPY", "SOFT_MAX", "ADD", "MUL", "ARGSORT"):
        assert ran > 0, f"no {op} case ran on the plug-in:\n{clean[-1500:]}"


@pytest.fixture(scope="module")
def plugin_loaded(eb):
    _need_ref()
    n = rr.lib().rr_load_backend(PLUGIN.encode())
    assert n >= 2, "ggml_backend_load(libggml-b200.so) failed"
    names = [rr.lib().rr_dev_name(i).decode() for i in range(n)]
    assert "B2000" in names, names
    return names


HP = dict(E=256, H=2, F=512, V=1024, L=2)


@pytest.mark.parametrize("ftype", ["q4_k_m", po.Q4_0, po.Q8_0])
def test_llama_decode_tree_batch_on_plugin_vs_cpu(plugin_loaded, tmp_path, ftype):
    """the reference's llama_decode on a tree-shaped batch: all layers on the plug-in vs ggml-cpu, same GGUF, same calls"""
    tgt = fx.make_model(**HP, ftype=ftype, seed=61)
    path = gf.write_gguf(tgt, str(tmp_path / "tgt.gguf"))
    rng = np.random.default_rng(5)
    T, n_prompt, n_ctx = 27, 12, 256
    par = fx.tree_parents(T, rng); depth = fx.tree_depths(par)
    seqs, n_seq = fx.tree_seq_ids(par)
    prompt = rng.integers(3, HP["V"], n_prompt).astype(np.int32)
    tokens = rng.integers(3, HP["V"], T).astype(np.int32)
    out = {}
    for name, ngl in (("cpu", 0), ("plugin", 99)):
        m = rr.Model(path, n_gpu_layers=ngl)
        c = rr.Context(m, n_ctx, n_seq_max=max(8, n_seq))
        lp, taps_p = c.decode_tapped(prompt, np.arange(n_prompt))
        for s in range(1, n_seq):
            c.seq_cp(0, s)
        lt, taps_t = c.decode_tapped(tokens, n_prompt + depth, seqs)
        out[name] = (lp, lt, pa.planes_from_ref_taps(taps_p, m.tensor_type), pa.planes_from_ref_taps(taps_t, m.tensor_type))
        c.free(); m.free()
    for i, tag in ((0, "prompt"), (1, "tree")):
        a, b = out["plugin"][i], out["cpu"][i]
        rel = pa.row_rel(a, b)
        flipped, per_site = pa.flipped_rows(out["plugin"][2 + i], out["cpu"][2 + i])
        print(f"[dropin] llama_decode {tag} ({ftype}): plugin vs ggml-cpu rows {len(rel)}, max rel {rel.max():.2e}, NMSE {nmse(a, b):.2e}, "
              f"rows with differing int8 planes {int(flipped.sum())} {per_site}")
        # the reference's own whole-graph bar (test_llama, R/tests/test-backend-ops.cpp:3482) and the north star on the unflipped rows;
        # vs the UNMODIFIED reference (f16 V accumulator) every row may differ at the wo input, hence FLIP_REL as the outer bound
        assert nmse(a, b) <= 2e-3 and rel.max() <= pa.FLIP_REL
        if (~flipped).any():
            assert rel[~flipped].max() <= pa.NORTH_STAR_REL
        bad, ties, n = pa.greedy_report(a[~flipped], b[~flipped]) if (~flipped).any() else (0, 0, 0)
        assert bad == 0


def test_llama_decode_prompt_batch_on_plugin(plugin_loaded, tmp_path):
    """T = 200 in one llama_decode (n_ubatch 512): MUL_MAT and FLASH_ATTN_EXT beyond one 128-row verify tile through the plug-in"""
    tgt = fx.make_model(**HP, ftype="q4_k_m", seed=62)
    path = gf.write_gguf(tgt, str(tmp_path / "tgt.gguf"))
    rng = np.random.default_rng(6)
    T = 200
    prompt = rng.integers(3, HP["V"], T).astype(np.int32)
    res = {}
    for name, ngl in (("cpu", 0), ("plugin", 99)):
        m = rr.Model(path, n_gpu_layers=ngl)
        c = rr.Context(m, 512)
        res[name] = c.decode(prompt, np.arange(T))
        c.free(); m.free()
    rel = pa.row_rel(res["plugin"], res["cpu"])
    print(f"[dropin] llama_decode T={T}: plugin vs ggml-cpu max rel {rel.max():.2e}, median {np.median(rel):.2e}, NMSE {nmse(res['plugin'], res['cpu']):.2e}")
    assert nmse(res["plugin"], res["cpu"]) <= 2e-3 and rel.max() <= pa.FLIP_REL
