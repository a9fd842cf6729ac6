import os

import numpy as np

import pyoracle as po

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_ops():
    return np.load(os.path.join(GOLD, "ops.npz"))


def load_tiny():
    """-> (golden arrays, target fixture, head fixture)"""
    z = np.load(os.path.join(GOLD, "tiny_model.npz"))
    hp = dict(E=256, H=2, F=512, V=512, L=2, eps=1e-5, freq_base=10000.0)
    fxs = {}
    for name, L, eagle in (("tgt", 2, False), ("head", 1, True)):
        f = {"hparams": {**hp, "L": L, "is_eagle": eagle}, "globals": {}, "layers": [dict() for _ in range(L)]}
        for key in z.files:
            if not key.startswith(name + "."):
                continue
            parts = key.split(".")
            t = int(parts[-1][1:])
            if parts[1].startswith("g"):
                f["globals"][int(parts[1][1:])] = (t, np.ascontiguousarray(z[key]))
            else:
                f["layers"][int(parts[1][1:])][int(parts[2])] = (t, np.ascontiguousarray(z[key]))
        fxs[name] = f
    return z, fxs["tgt"], fxs["head"]


def nmse(a, b):
    """test-backend-ops' metric (R/tests/test-backend-ops.cpp:319, 524-545)"""
    a = np.asarray(a, np.float64).ravel(); b = np.asarray(b, np.float64).ravel()
    return float(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-30))


def rel_max(a, b):
    return float(np.abs(np.asarray(a, np.float64) - b).max() / max(np.abs(b).max(), 1e-30))
