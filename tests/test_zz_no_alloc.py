"""jb_mlpg_solve_noalloc (no counterpart in the reference): the MLPG
solve with its work planes in a fixed `static __device__` array instead of cudaMalloc'ed memory, called through the
C ABI and compared with the existing entry point jb_engine_mlpg_create on the same input - twice in a row with
different shapes (the scratch keeps the last call's contents), then the first call again.

Collected last on purpose. Device buffers of the TEST come from the CUDA runtime
(cuda.bindings); the entry point itself allocates nothing, which the test checks with cudaMemGetInfo."""
import numpy as np
import pytest

import jbonsai_b200 as jb
from conftest import NITECH

pytestmark = pytest.mark.gpu

PARAM_RTOL = 1e-9  # the bar of tests/test_gpu_mlpg_seam.py (north star: parameters <= 1e-9 relative)

WIN_133 = [[1.0], [-0.5, 0.0, 0.5], [1.0, -2.0, 1.0]]
WIN_15 = [[1.0], [-0.2, -0.1, 0.0, 0.1, 0.2]]
WIN_STATIC = [[1.0]]


def rel_err(a, b):
    d = np.abs(a - b)
    return float((d / np.maximum(np.abs(b), 1e-300)).max()) if d.size else 0.0


def stream(seed, S, dim, nwin):
    rng = np.random.default_rng(seed)
    dur = rng.integers(1, 9, S).astype(np.int32)
    mean = rng.standard_normal((S, nwin * dim))
    mean[:, dim:] *= 0.1
    vari = rng.uniform(0.05, 2.0, (S, nwin * dim))
    vari[S // 3, 0] = 1e-30  # MeanVari::with_ivar: |v| < 1e-19 -> 1e38
    if nwin > 1:
        vari[S // 2, dim] = 1e25  # |v| > 1e19 -> 0
    return dur, mean, vari


class DeviceBuffers:
    """Plain cudaMalloc'ed buffers of the test (allocated before the calls under test)."""

    def __init__(self, nbytes_each, n):
        from cuda.bindings import runtime as cudart
        self.rt = cudart
        self.ptrs = []
        for _ in range(n):
            err, p = cudart.cudaMalloc(nbytes_each)
            assert int(err) == 0
            self.ptrs.append(int(p))

    def put(self, i, arr):
        a = np.ascontiguousarray(arr, dtype=np.float64)
        (err,) = self.rt.cudaMemcpy(self.ptrs[i], a.ctypes.data, a.nbytes, self.rt.cudaMemcpyKind.cudaMemcpyHostToDevice)
        assert int(err) == 0

    def get(self, i, shape):
        out = np.empty(shape, dtype=np.float64)
        # a blocking copy on the default stream: ordered behind the two kernels the call enqueued there
        (err,) = self.rt.cudaMemcpy(out.ctypes.data, self.ptrs[i], out.nbytes, self.rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)
        assert int(err) == 0
        return out

    def fill(self, i, byte, nbytes):
        (err,) = self.rt.cudaMemset(self.ptrs[i], byte, nbytes)
        assert int(err) == 0

    def free_bytes(self):
        (err,) = self.rt.cudaDeviceSynchronize()
        assert int(err) == 0
        err, free, _total = self.rt.cudaMemGetInfo()
        assert int(err) == 0
        return int(free)

    def close(self):
        for p in self.ptrs:
            self.rt.cudaFree(p)
        self.ptrs = []


def test_no_alloc_mlpg_solve_matches_mlpg_create_on_two_calls_in_a_row():
    eng = jb.Engine.load([NITECH])
    cases = [  # (seed, states, dim, windows): the second call is smaller and has a wider band than the first
        (3, 260, 35, WIN_133),
        (4, 90, 7, WIN_15),
        (5, 40, 3, WIN_STATIC),
    ]
    inputs, want = [], []
    for seed, S, dim, win in cases:
        dur, mean, vari = stream(seed, S, dim, len(win))
        want.append(eng.mlpg_create(dur, dim, win, mean, vari))  # the existing entry point (state-level input)
        inputs.append((np.repeat(mean, dur, axis=0), np.repeat(vari, dur, axis=0), int(dur.sum()), dim, win))
        assert want[-1].shape == (int(dur.sum()), dim)
        assert int(dur.sum()) * dim <= jb.MLPG_NOALLOC_MAX_CELLS

    cap = max(m.nbytes for m, _, _, _, _ in inputs)
    buf = DeviceBuffers(cap, 3)  # mean, vari, out
    try:
        def run(k):
            m, v, T, dim, win = inputs[k]
            buf.put(0, m)
            buf.put(1, v)
            buf.fill(2, 0xFF, T * dim * 8)  # NaN pattern: every output must be written by the call
            jb.mlpg_solve_noalloc(0, buf.ptrs[0], buf.ptrs[1], T, dim, win, buf.ptrs[2])
            return buf.get(2, (T, dim))

        first = run(0)  # also the per-device set-up (cudaGetSymbolAddress of the scratch array)
        assert np.isfinite(first).all()
        assert rel_err(first, want[0]) <= PARAM_RTOL
        free0 = buf.free_bytes()
        second = run(1)  # straight after, over the first call's leftovers in the scratch array
        assert np.isfinite(second).all()
        assert rel_err(second, want[1]) <= PARAM_RTOL
        third = run(2)
        assert rel_err(third, want[2]) <= PARAM_RTOL
        again = run(0)
        assert np.array_equal(again, first)  # same bits whatever the scratch held before
        assert buf.free_bytes() == free0  # the calls allocated nothing on the device

        # beyond the stated maximum: a clear error, nothing launched, and the entry point keeps working
        m, v, T, dim, win = inputs[0]
        too_many = jb.MLPG_NOALLOC_MAX_CELLS // dim + 1
        with pytest.raises(jb.JbError) as ei:
            jb.mlpg_solve_noalloc(0, buf.ptrs[0], buf.ptrs[1], too_many, dim, win, buf.ptrs[2])
        assert ei.value.kind == "Argument" and "JB_MLPG_NOALLOC_MAX_CELLS" in str(ei.value)
        with pytest.raises(jb.JbError):  # even window width: not a window
            jb.mlpg_solve_noalloc(0, buf.ptrs[0], buf.ptrs[1], T, dim, [[1.0], [0.5, 0.5]], buf.ptrs[2])
        jb.mlpg_solve_noalloc(0, buf.ptrs[0], buf.ptrs[1], 0, dim, win, buf.ptrs[2])  # empty utterance: no-op
        assert np.array_equal(run(1), second)
    finally:
        buf.close()
