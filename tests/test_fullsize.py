"""GPU: one full-size tree-verify forward per BASELINE config (32 layers, E 4096, F 11008, 32 heads, V 32000, the config's
quantisation mix, tree shape and KV length) against the reference's own llama_decode on ggml-cpu (oracle/_ref), see
tests/fullsize_run.py.  Two reference builds, one process each:

* "ref"    — the UNMODIFIED reference.  Its CPU flash attention rounds the V accumulator to f16 after every KV cell
             (R/ggml/src/ggml-cpu/ggml-cpu.c:11154-11168); over 4096 cells that accumulator alone costs the reference ~1e-2..1e-1
             of its own attention output (increments below half an f16 ulp are lost), which no tensor-core kernel can reproduce
             bit for bit.  Numbers are reported and bounded per config.
* "f32acc" — the same 16 source files with ONE line changed (the accumulator takes the kernel's own f32 branch, :11171-11187;
             oracle/Makefile).  Against it the engine must agree like an implementation of the same arithmetic: rows whose int8
             activation planes are identical through all 129 sites to 1e-4, every row to the north star's 1e-2, greedy ids identical.
The difference between the two reports is therefore exactly the reference's accumulator rounding, not the engine.
"""
import json
import os
import subprocess
import sys

import pytest

import parity as pa
import refround as rr

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
# unmodified-reference bars per config: NMSE (the reference's whole-graph metric, R/tests/test-backend-ops.cpp:3482) and max row rel
REF_BARS = {"config1": (2e-3, 1e-2), "config2": (5e-2, 3.5e-1)}


def _run(cfg_id, variant, tmp_path):
    if not rr.have() or not os.path.exists(os.path.join(rr.REF, rr.VARIANTS[variant])):
        pytest.fail("oracle/_ref (the reference build) is missing: run __graft_entry__.build() where /root/reference is present")
    r = subprocess.run([sys.executable, os.path.join(HERE, "fullsize_run.py"), cfg_id, variant, str(tmp_path)], capture_output=True, text=True, timeout=1500)
    line = [l for l in r.stdout.splitlines() if l.startswith("FULLSIZE_JSON ")]
    assert r.returncode == 0 and line, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads(line[-1][len("FULLSIZE_JSON "):])
    print(f"[fullsize] {res['config']} vs reference build '{variant}': logits max rel {res['rel_max']:.2e} (median {res['rel_median']:.2e}), NMSE {res['nmse']:.2e}; "
          f"rows with differing int8 planes {res['rows_differing_planes']} of {res['T']} ({res['rows_primary_flip']} primary flips, first differing site "
          f"{res['first_differing_site']} of {res['n_sites']}); clean-row max rel {res['clean_rel_max']}, differing-row max rel {res['differing_rel_max']}; greedy "
          f"{res['greedy_identical']} identical / {res['greedy_undecided']} undecided / {res['greedy_mismatch']} mismatches; layer-0 tree K/V bit-identical "
          f"{100 * res['kv_layer0_bit_identical'][0]:.2f} % / {100 * res['kv_layer0_bit_identical'][1]:.2f} %; reference decode {res['reference_decode_s']} s on {res['host_threads']} threads")
    return res


@pytest.mark.parametrize("cfg_id", ["config1", "config2"])
def test_fullsize_verify_vs_reference_f32_accumulator(eb, tmp_path, cfg_id):
    res = _run(cfg_id, "f32acc", tmp_path)
    assert res["rel_max"] <= pa.NORTH_STAR_REL, res
    assert res["clean_rel_max"] is None or res["clean_rel_max"] <= 1e-4, res
    assert res["nmse"] <= 1e-5 and res["greedy_mismatch"] == 0, res
    assert min(res["kv_layer0_bit_identical"]) > 0.99, res      # an f32 difference of 1e-6 relative moves ~0.2 % of the f16 roundings by an ulp


@pytest.mark.parametrize("cfg_id", ["config1", "config2"])
def test_fullsize_verify_vs_unmodified_reference(eb, tmp_path, cfg_id):
    res = _run(cfg_id, "ref", tmp_path)
    bar_nmse, bar_rel = REF_BARS[cfg_id]
    assert res["nmse"] <= bar_nmse and res["rel_max"] <= bar_rel and res["greedy_mismatch"] == 0, res
    assert min(res["kv_layer0_bit_identical"]) > 0.99, res
