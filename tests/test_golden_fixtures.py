"""Committed known-answer vectors (tests/golden/, made by tests/golden/make_goldens.py).

CPU (-m "not gpu"): the oracle reproduces both files (reference_goldens.json = the reference's own
src/lib.rs vectors at its eps 1e-10; oracle_vectors.npz bit for bit). GPU: the CUDA path through
the C ABI meets the north-star bars against the same files, with no access to /root/reference."""
import json
import os

import numpy as np
import pytest

import jbonsai_b200 as jb
from conftest import NITECH, TOHOKU_HAPPY, TOHOKU_NEUTRAL
from jbonsai_b200 import workload as wl
from labels import SAMPLE_SENTENCE_1, SAMPLE_SENTENCE_2
from oracle_lib import Oracle

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REF = json.load(open(os.path.join(HERE, "reference_goldens.json")))
VEC = np.load(os.path.join(HERE, "oracle_vectors.npz"))
VOICE = {"nitech": NITECH, "tohoku-f01-neutral": TOHOKU_NEUTRAL, "tohoku-f01-happy": TOHOKU_HAPPY}
LABELS = {"SAMPLE_SENTENCE_1": SAMPLE_SENTENCE_1, "SAMPLE_SENTENCE_2": SAMPLE_SENTENCE_2, "[]": []}
CASES = {"nitech_s1": (NITECH, SAMPLE_SENTENCE_1), "nitech_s2": (NITECH, SAMPLE_SENTENCE_2),
         "tohoku_syn0": (TOHOKU_NEUTRAL, wl.synthetic_batch(1, wl.SEED_TOHOKU, 0, 6, 6)[0])}
REF_CASES = [k for k in REF if k != "eps"]


def _configure(obj, case, is_oracle):
    if "speed" in case:
        obj.set("speed", case["speed"])
    for key, w in case.get("weights", {}).items():
        kind, stream = ("duration", 0) if key == "duration" else ("parameter", int(key[-1]))
        (obj.set_weights if is_oracle else obj.set_interpolation_weight)(kind, stream, w)


@pytest.mark.parametrize("name", REF_CASES)
def test_oracle_meets_reference_goldens(name):
    case = REF[name]
    o = Oracle([VOICE[v] for v in case["voices"]])
    _configure(o, case, True)
    for fma in (False, True):
        s = o.synthesize(LABELS[case["labels"]], fma)
        assert len(s) == case["len"]
        for i, v in case["samples"].items():
            assert abs(s[int(i)] - v) < REF["eps"]


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_reproduces_committed_vectors(name):
    voice, labels = CASES[name]
    o = Oracle([voice])
    mcp, lf0, lpf = o.parameters(labels)
    tap = o.synthesize_tap(labels, fma=True)
    fr, si = VEC[name + "/frame_idx"], VEC[name + "/sample_idx"]
    assert mcp.shape[0] == int(VEC[name + "/frames"])
    assert np.array_equal(mcp[fr], VEC[name + "/mcp"]) and np.array_equal(lpf[fr], VEC[name + "/lpf"])
    assert np.array_equal(lf0[fr], VEC[name + "/lf0"])
    assert np.array_equal(np.flatnonzero(tap["pulse"]), VEC[name + "/pulse_pos"])
    assert np.array_equal(tap["wave"][si], VEC[name + "/wave"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", REF_CASES)
def test_gpu_meets_reference_goldens(name):
    case = REF[name]
    e = jb.Engine.load([VOICE[v] for v in case["voices"]])
    _configure(e, case, False)
    s = e.synthesize(LABELS[case["labels"]])
    assert len(s) == case["len"]
    for i, v in case["samples"].items():
        assert abs(s[int(i)] - v) < 1e-6  # north-star waveform bar


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_meets_committed_vectors(name):
    voice, labels = CASES[name]
    e = jb.Engine.load([voice])
    dbg = e.debug_batch([labels])
    fr, si = VEC[name + "/frame_idx"], VEC[name + "/sample_idx"]
    rel = lambda a, b: float((np.abs(a - b) / np.maximum(np.abs(b), 1e-300)).max())
    assert dbg["mcp"].shape[0] == int(VEC[name + "/frames"])             # frame indexing: exact
    assert rel(dbg["mcp"][fr], VEC[name + "/mcp"]) <= 1e-9               # parameters: 1e-9 relative
    assert rel(dbg["lpf"][fr], VEC[name + "/lpf"]) <= 1e-9
    assert rel(dbg["lf0"].reshape(-1)[fr], VEC[name + "/lf0"]) <= 1e-9
    mask = np.unpackbits(VEC[name + "/voiced_mask"])[: dbg["mcp"].shape[0]].astype(bool)
    assert np.array_equal(dbg["lf0"].reshape(-1) != -1e10, mask)         # voiced mask: exact
    assert np.array_equal(np.flatnonzero(dbg["pulse"]), VEC[name + "/pulse_pos"])  # pulse positions: exact
    w = e.synthesize(labels)
    assert np.abs(w[si] - VEC[name + "/wave"]).max() <= 1e-6             # waveform: 1e-6 absolute
    assert abs(w.sum() - float(VEC[name + "/wave_sum"])) <= 1e-6 * len(w)
