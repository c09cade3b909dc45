"""Generates tests/golden/*.json|npz (committed): small known-answer vectors for the hot path.

  reference_goldens.json  the reference's own pinned end-to-end vectors, transcribed from
                          jbonsai src/lib.rs:38-160 (lengths, 14 samples, eps 1e-10). The
                          reference is Rust and cannot run here, so these are the only vectors
                          that come from the reference itself.
  oracle_vectors.npz      vectors from oracle/ (the CPU restatement pinned to the file above):
                          MLPG parameter rows, pulse positions, excitation and waveform samples
                          for three utterances, so that the GPU box can check the CUDA path without
                          /root/reference and a later change of the oracle shows up as a diff.

Run from the repo root after `make -C oracle`:  python tests/golden/make_goldens.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import NITECH, TOHOKU_NEUTRAL  # noqa: E402
from labels import SAMPLE_SENTENCE_1, SAMPLE_SENTENCE_2  # noqa: E402
from oracle_lib import Oracle  # noqa: E402
from jbonsai_b200 import workload as wl  # noqa: E402

REFERENCE = {  # src/lib.rs of jbonsai v0.4.2
    "bonsai": {"lib_rs": "39-47", "voices": ["nitech"], "labels": "SAMPLE_SENTENCE_1", "len": 66480,
               "samples": {"2000": 19.35141137623778, "30000": -980.6757547598129}},
    "bonsai_multi": {"lib_rs": "78-91", "voices": ["tohoku-f01-neutral", "tohoku-f01-happy"],
                     "weights": {"duration": [0.7, 0.3], "parameter0": [0.7, 0.3], "parameter1": [0.7, 0.3],
                                 "parameter2": [1.0, 0.0]},
                     "labels": "SAMPLE_SENTENCE_1", "len": 74880,
                     "samples": {"2000": 2.3158134981607754e-5, "30000": 6459.375032316974}},
    "is_this_bonsai": {"lib_rs": "123-133", "voices": ["nitech"], "labels": "SAMPLE_SENTENCE_2", "len": 100800,
                       "samples": {"2000": 17.15977345625943, "30000": 2566.2058730889985,
                                   "70000": -1898.2890228814217, "100799": -13.514971382534956}},
    "is_this_bonsai_fast": {"lib_rs": "136-147", "voices": ["nitech"], "speed": 1.4,
                            "labels": "SAMPLE_SENTENCE_2", "len": 72000,
                            "samples": {"2000": 15.0481014871396, "30000": -56.77163803227678,
                                        "70000": -9.15409432584658, "71199": 7.840225089163972}},
    "empty": {"lib_rs": "150-160", "voices": ["nitech"], "labels": "[]", "len": 0, "samples": {}},
    "eps": 1e-10,
}

CASES = {  # name -> (voice path, labels)
    "nitech_s1": (NITECH, SAMPLE_SENTENCE_1),
    "nitech_s2": (NITECH, SAMPLE_SENTENCE_2),
    "tohoku_syn0": (TOHOKU_NEUTRAL, wl.synthetic_batch(1, wl.SEED_TOHOKU, 0, 6, 6)[0]),
}


def main():
    with open(os.path.join(HERE, "reference_goldens.json"), "w") as f:
        json.dump(REFERENCE, f, indent=1, sort_keys=True)
    out = {}
    for name, (voice, labels) in CASES.items():
        o = Oracle([voice])
        mcp, lf0, lpf = o.parameters(labels)
        tap = o.synthesize_tap(labels, fma=True)
        T, N = mcp.shape[0], len(tap["wave"])
        fr = np.unique(np.linspace(0, T - 1, 24).astype(np.int64))
        si = np.unique(np.linspace(0, N - 1, 512).astype(np.int64))
        out[name + "/frames"] = np.int64(T)
        out[name + "/frame_idx"] = fr
        out[name + "/mcp"] = mcp[fr]
        out[name + "/lf0"] = lf0[fr]
        out[name + "/lpf"] = lpf[fr]
        out[name + "/voiced_mask"] = np.packbits(lf0 != -1e10)
        out[name + "/pulse_pos"] = np.flatnonzero(tap["pulse"]).astype(np.int64)
        out[name + "/sample_idx"] = si
        out[name + "/gained"] = tap["gained"][si]
        out[name + "/wave"] = tap["wave"][si]
        out[name + "/wave_sum"] = np.float64(tap["wave"].sum())
    np.savez_compressed(os.path.join(HERE, "oracle_vectors.npz"), **out)
    print("wrote", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
