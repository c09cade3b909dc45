"""How many frames of warm-up does the MLSA filter need before its state is independent of where
it started? (sizing of jb_engine_set_mlsa_segmenting / mlsa_seg.cu; TEST TOOLING: uses oracle/.)

For synthetic utterances it filters frames [F0 - W, F0 + 4) from a zero state and compares the last
four frames with the whole-utterance oracle waveform. Results (max over cut positions every 37
frames): nitech, 30 utterances: W=24 -> 4.4e-7, W=32 -> 2.7e-10, W=40 -> 5.0e-11, W=48 -> 3.3e-11;
tohoku-f01-neutral, 24 utterances: W>=24 -> ~1e-10 (rounding-noise floor of two fp64 histories).

  python tools/mlsa_memory_probe.py nitech 6 0
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_lib import Oracle  # noqa: E402
from jbonsai_b200 import workload as wl  # noqa: E402

so = "/tmp/libmlsa_memory_probe.so"
subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-mfma", "-shared", "-o", so,
                       os.path.join(ROOT, "tools", "mlsa_memory_probe.cc"), "-L" + os.path.join(ROOT, "oracle"),
                       "-loracle", "-Wl,-rpath," + os.path.join(ROOT, "oracle")])
L = C.CDLL(so)
D = C.POINTER(C.c_double)
voice = sys.argv[1] if len(sys.argv) > 1 else "nitech"
n_utt = int(sys.argv[2]) if len(sys.argv) > 2 else 6
first = int(sys.argv[3]) if len(sys.argv) > 3 else 0
o = Oracle([wl.VOICES[voice]])
worst = {}
for ui, u in enumerate(wl.synthetic_batch(n_utt, wl.SEED_NITECH, first, 30, 50)):
    tap = o.synthesize_tap(u, fma=True)
    mcp, _, _ = o.parameters(u)
    T, fp = mcp.shape[0], 240
    x, ref, mcpc = np.ascontiguousarray(tap["gained"]), tap["wave"], np.ascontiguousarray(mcp)

    def run(f0, f1):
        out = np.zeros((f1 - f0) * fp)
        L.mlsa_from(f0, f1, fp, C.c_double(0.55), mcpc.ctypes.data_as(D), x.ctypes.data_as(D), out.ctypes.data_as(D))
        return out

    assert np.array_equal(run(0, T), ref)
    for W in (16, 24, 32, 40, 48):
        errs = [np.abs(run(F0 - W, min(F0 + 4, T))[W * fp:] - ref[F0 * fp:min(F0 + 4, T) * fp]).max()
                for F0 in range(W + 5, T - 2, 37)]
        worst.setdefault(W, []).append(max(errs))
    print(ui, "frames", T, {W: f"{v[-1]:.2e}" for W, v in worst.items()}, flush=True)
print("worst:", {W: f"{max(v):.2e}" for W, v in worst.items()})
