"""The MLPG seam of the C ABI (jb_engine_mlpg_create = MlpgAdjust::new + create, src/mlpg_adjust/mod.rs:31-95)
against the oracle's restatement of mask.rs / mlpg.rs, on synthetic window shapes the shipped voices do not
have (band widths 1, 3 and 5), with and without MSD masking and GV."""
import numpy as np
import pytest

import jbonsai_b200 as jb
from conftest import NITECH
import oracle_lib as O

pytestmark = pytest.mark.gpu

PARAM_RTOL = 1e-9
F64_MAX = np.finfo(np.float64).max

WINDOWS = {
    "static": [[1.0]],
    "1-3": [[1.0], [-0.5, 0.0, 0.5]],
    "1-3-3": [[1.0], [-0.5, 0.0, 0.5], [1.0, -2.0, 1.0]],
    "1-5": [[1.0], [-0.2, -0.1, 0.0, 0.1, 0.2]],
    "1-3-5": [[1.0], [-0.5, 0.0, 0.5], [0.25, 0.0, -0.5, 0.0, 0.25]],
}


def rel_err(a, b):
    d = np.abs(a - b)
    return float((d / np.maximum(np.abs(b), 1e-300)).max()) if d.size else 0.0


def stream(seed, S, dim, nwin, msd):
    rng = np.random.default_rng(seed)
    dur = rng.integers(1, 9, S).astype(np.int32)
    mean = rng.standard_normal((S, nwin * dim))
    mean[:, dim:] *= 0.1  # dynamic features are small
    vari = rng.uniform(0.05, 2.0, (S, nwin * dim))
    if msd:
        m = (np.sin(np.arange(S) * 0.9 + seed) > -0.3).astype(np.float64)  # voiced / unvoiced runs
        m[S // 2] = 0.0
    else:
        m = np.full(S, F64_MAX)  # model/mod.rs:113: non-MSD streams carry f64::MAX
    return dur, mean, vari, m


@pytest.fixture(scope="module")
def eng():
    return jb.Engine.load([NITECH])


@pytest.mark.parametrize("shape", list(WINDOWS))
@pytest.mark.parametrize("msd", [False, True])
def test_mlpg_seam_window_shapes(eng, shape, msd):
    win = WINDOWS[shape]
    dim = 3
    dur, mean, vari, m = stream(len(shape) + 7 * msd, 40, dim, len(win), msd)
    want = O.mlpg(dur, dim, win, mean, vari, m, 0.5)
    got = eng.mlpg_create(dur, dim, win, mean, vari, msd=m if msd else None)
    assert got.shape == want.shape == (int(dur.sum()), dim)
    assert np.array_equal(got == -1e10, want == -1e10)  # Mask::fill: same frames masked
    v = want != -1e10
    assert rel_err(got[v], want[v]) <= PARAM_RTOL


@pytest.mark.parametrize("shape", ["static", "1-3-3", "1-3-5"])
@pytest.mark.parametrize("msd", [False, True])
def test_mlpg_seam_with_gv(eng, shape, msd):
    win = WINDOWS[shape]
    dim = 2
    S = 60
    dur, mean, vari, m = stream(11 + len(shape), S, dim, len(win), msd)
    rng = np.random.default_rng(5)
    gv = (rng.uniform(0.5, 1.5, dim), rng.uniform(0.01, 0.1, dim), (np.arange(S) % 7 != 0).astype(np.uint8))
    want = O.mlpg(dur, dim, win, mean, vari, m, 0.5, gv=gv, gv_weight=0.9)
    got = eng.mlpg_create(dur, dim, win, mean, vari, msd=m if msd else None, gv=gv, gv_weight=0.9)
    assert np.array_equal(got == -1e10, want == -1e10)
    v = want != -1e10
    assert rel_err(got[v], want[v]) <= PARAM_RTOL


def test_mlpg_seam_batch_ragged_and_errors(eng):
    win = WINDOWS["1-3-3"]
    dim = 4
    streams = [stream(s, n, dim, 3, True) for s, n in ((1, 25), (2, 1), (3, 60))]
    got = eng.mlpg_create([s[0] for s in streams], dim, win, [s[1] for s in streams], [s[2] for s in streams],
                          msd=[s[3] for s in streams])
    assert len(got) == 3
    for g, (dur, mean, vari, m) in zip(got, streams):
        want = O.mlpg(dur, dim, win, mean, vari, m, 0.5)
        assert g.shape == want.shape and np.array_equal(g == -1e10, want == -1e10)
        v = want != -1e10
        assert rel_err(g[v], want[v]) <= PARAM_RTOL
    with pytest.raises(jb.JbError):  # even window width: not a window (window.rs)
        eng.mlpg_create(streams[0][0], dim, [[1.0], [0.5, 0.5]], streams[0][1][:, : 2 * dim], streams[0][2][:, : 2 * dim])
