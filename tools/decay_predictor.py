"""Which cuts of the time-segmented MLSA filter need a long warm-up? (TEST TOOLING: uses oracle/.)
For cuts every `step` frames of synthetic utterances: output error after W warm-up frames from a zero state, for
several W, next to candidate predictors computed from the coefficient rows of the warm-up region.
  python tools/decay_predictor.py nitech 24"""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_lib import Oracle  # noqa: E402
from jbonsai_b200 import workload as wl  # noqa: E402
so = "/tmp/libmlsa_memory_probe.so"
subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-mfma", "-shared", "-o", so,
                       os.path.join(ROOT, "tools", "mlsa_memory_probe.cc"), "-L" + os.path.join(ROOT, "oracle"),
                       "-loracle", "-Wl,-rpath," + os.path.join(ROOT, "oracle")])
L = C.CDLL(so)
D = C.POINTER(C.c_double)
voice = sys.argv[1] if len(sys.argv) > 1 else "nitech"
n_utt = int(sys.argv[2]) if len(sys.argv) > 2 else 12
alpha = 0.55
o = Oracle([wl.VOICES[voice]])
WS = (16, 20, 24, 28, 32, 40)
rows = []
for ui, u in enumerate(wl.synthetic_batch(n_utt, wl.SEED_NITECH, 0, 30, 50)):
    tap = o.synthesize_tap(u, fma=True)
    mcp, _, _ = o.parameters(u)
    T, fp = mcp.shape[0], 240
    x, ref, mcpc = np.ascontiguousarray(tap["gained"]), tap["wave"], np.ascontiguousarray(mcp)
    b = mcp.copy()
    for i in range(mcp.shape[1] - 2, -1, -1):
        b[:, i] = mcp[:, i] - alpha * b[:, i + 1]
    s1 = np.abs(b[:, 1:]).sum(axis=1)
    s2 = np.abs(b[:, 2:]).sum(axis=1)

    def run(f0, f1):
        out = np.zeros((f1 - f0) * fp)
        L.mlsa_from(f0, f1, fp, C.c_double(alpha), mcpc.ctypes.data_as(D), x.ctypes.data_as(D), out.ctypes.data_as(D))
        return out
    for F0 in range(45 + (ui * 7) % 23, T - 6, 23):
        errs = [np.abs(run(F0 - W, F0 + 4)[W * fp:] - ref[F0 * fp:(F0 + 4) * fp]).max() for W in WS]
        sig = np.abs(ref[(F0 - 40) * fp:F0 * fp]).max()
        rows.append(errs + [s1[F0 - 24:F0].max(), s1[F0 - 24:F0].mean(), s2[F0 - 24:F0].max(), b[F0 - 24:F0, 0].max(), b[F0 - 24:F0, 1].max(), sig]
                    + list(b[F0 - 44:F0, 0]) + list(s1[F0 - 44:F0]))
    print(ui, T, len(rows), flush=True)
R = np.array(rows)
np.save(f"/tmp/decay_rows_{voice}.npy", R)
tol = 5e-8
for wi, W in enumerate(WS):
    bad = R[:, wi] > tol
    print(f"W={W}: cuts over {tol:g}: {bad.sum()} of {len(R)}  (max {R[:, wi].max():.2e})")
names = ["s1max", "s1mean", "s2max", "b0max", "b1max", "sig"]
for wi in (1, 2):
    bad = R[:, wi] > tol
    if not bad.any():
        continue
    for fi, nm in enumerate(names):
        f = R[:, len(WS) + fi]
        thr = f[bad].min()  # flag everything at least as extreme as the mildest bad cut
        print(f"  W={WS[wi]} predictor {nm}: bad range [{f[bad].min():.3g}, {f[bad].max():.3g}], all p50 {np.median(f):.3g} p90 {np.percentile(f, 90):.3g};"
              f" flagging f >= {thr:.3g} catches all bad and flags {np.mean(f >= thr) * 100:.1f} % of cuts")
