"""Random-init GGUF fixtures for the reference's own loader (test infrastructure).

Writes a `fixtures.make_model` dict as a GGUF file that the reference's libllama loads unchanged:
arch `llama` for the target (keys R/src/llama-model.cpp:410-487,506-508, tensor names
R/src/llama-arch.cpp) and arch `eagle` for the draft head (tensor list R/src/llama-arch.cpp:1301-1326,
shapes R/src/llama-model.cpp:1589-1640); `tokenizer.ggml.model = "no_vocab"` + `<arch>.vocab_size`
so no tokenizer data is needed (R/src/llama-vocab.cpp:1350-1364).  The tensors carry the fixture's
own block bytes (already quantised), or f16 when `as_f16` is set — the input llama-quantize wants.

The vendored gguf-py of the reference is used when /root/reference is present, else the installed
`gguf` package (same writer; the fork's additions are enum entries this file does not need).
"""
import os
import sys

import numpy as np

_VENDORED = "/root/reference/llama.cpp/gguf-py"
if os.path.isdir(_VENDORED) and _VENDORED not in sys.path:
    sys.path.insert(0, _VENDORED)
import gguf  # noqa: E402

import pyoracle as po  # noqa: E402

QT = {po.F32: gguf.GGMLQuantizationType.F32, po.F16: gguf.GGMLQuantizationType.F16, po.Q4_0: gguf.GGMLQuantizationType.Q4_0,
      po.Q8_0: gguf.GGMLQuantizationType.Q8_0, po.Q4_K: gguf.GGMLQuantizationType.Q4_K, po.Q6_K: gguf.GGMLQuantizationType.Q6_K}
GLOBAL_NAMES = {po.T_TOK_EMBD: "token_embd.weight", po.T_OUTPUT_NORM: "output_norm.weight", po.T_OUTPUT: "output.weight",
                po.T_FC: "fc.weight", po.T_FC_BIAS: "fc.bias"}
LAYER_NAMES = {po.L_ATTN_NORM: "attn_norm", po.L_WQ: "attn_q", po.L_WK: "attn_k", po.L_WV: "attn_v", po.L_WO: "attn_output",
               po.L_FFN_NORM: "ffn_norm", po.L_GATE: "ffn_gate", po.L_UP: "ffn_up", po.L_DOWN: "ffn_down"}


def _shape(hp, layer, which):
    E, F, V = hp["E"], hp["F"], hp["V"]
    if layer < 0:
        return {po.T_TOK_EMBD: (V, E), po.T_OUTPUT_NORM: (E,), po.T_OUTPUT: (V, E), po.T_FC: (E, 2 * E), po.T_FC_BIAS: (E,)}[which]
    if which in (po.L_ATTN_NORM, po.L_FFN_NORM):
        return (E,)
    if which in (po.L_GATE, po.L_UP):
        return (F, E)
    if which == po.L_DOWN:
        return (E, F)
    return (E, E)


def _add(w, name, t, raw, shape, dequant=None):
    if t in (po.F32, po.F16):
        arr = raw.view(np.float32 if t == po.F32 else np.float16).reshape(shape)
        w.add_tensor(name, arr)
        return
    if dequant is not None:                 # f16 file for llama-quantize: the dequantised values of the fixture's blocks
        vals = np.empty(int(np.prod(shape)), np.float32)
        dequant(t, raw, vals)
        w.add_tensor(name, vals.astype(np.float16).reshape(shape))
        return
    rows = shape[0]
    w.add_tensor(name, raw.reshape(rows, -1), raw_dtype=QT[t])


def _add_spm_vocab(w, V):
    """a minimal SentencePiece-style vocabulary with byte fallback (R/src/llama-vocab.cpp, tokenizer model "llama"): <unk>, <s>, </s>, the
    256 byte tokens, then unique filler pieces — enough for the reference's drivers to tokenise any prompt and detokenise any id"""
    assert V >= 259 + 16
    toks, scores, types = ["<unk>", "<s>", "</s>"], [0.0, 0.0, 0.0], [2, 3, 3]          # UNKNOWN, CONTROL, CONTROL
    for b in range(256):
        toks.append("<0x%02X>" % b); scores.append(0.0); types.append(6)                 # BYTE
    i = 0
    while len(toks) < V:
        toks.append("\u2581t%d" % i); scores.append(-1000.0 - i); types.append(1); i += 1   # NORMAL, never produced by the tokenizer
    w.add_tokenizer_model("llama")
    w.add_token_list(toks); w.add_token_scores(scores); w.add_token_types(types)
    w.add_bos_token_id(1); w.add_eos_token_id(2); w.add_unk_token_id(0)
    w.add_add_bos_token(True); w.add_add_eos_token(False)


def write_gguf(fx, path, n_ctx_train=4096, as_f16=False, vocab=False):
    """fx: fixtures.make_model dict -> GGUF at `path` (arch llama or eagle by fx['hparams']['is_eagle']); vocab=True adds a byte-fallback
    SPM vocabulary so that the reference's command-line drivers can run on the file"""
    hp = fx["hparams"]
    arch = "eagle" if hp["is_eagle"] else "llama"
    w = gguf.GGUFWriter(path, arch)
    w.add_name("eagle-b200 %s fixture" % arch)
    w.add_uint32(f"{arch}.context_length", n_ctx_train)
    w.add_uint32(f"{arch}.embedding_length", hp["E"])
    w.add_uint32(f"{arch}.block_count", hp["L"])
    w.add_uint32(f"{arch}.feed_forward_length", hp["F"])
    w.add_uint32(f"{arch}.attention.head_count", hp["H"])
    w.add_uint32(f"{arch}.attention.head_count_kv", hp["H"])
    w.add_float32(f"{arch}.attention.layer_norm_rms_epsilon", hp["eps"])
    w.add_uint32(f"{arch}.rope.dimension_count", hp["E"] // hp["H"])
    w.add_float32(f"{arch}.rope.freq_base", hp["freq_base"])
    w.add_uint32(f"{arch}.vocab_size", hp["V"])
    if vocab:
        _add_spm_vocab(w, hp["V"])
    else:
        w.add_string("tokenizer.ggml.model", "no_vocab")
    dq = None
    if as_f16:
        def dq(t, raw, out):
            po.oracle().dequantize_row(t, raw, out) if hasattr(po.oracle(), "dequantize_row") else po.oracle()._g("dequantize_row")(t, po._ptr(raw), po._ptr(out), out.size)
    for which, (t, raw) in fx["globals"].items():
        _add(w, GLOBAL_NAMES[which], t, raw, _shape(hp, -1, which), dq)
    for l, d in enumerate(fx["layers"]):
        for which, (t, raw) in d.items():
            _add(w, f"blk.{l}.{LAYER_NAMES[which]}.weight", t, raw, _shape(hp, l, which), dq)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    return path
