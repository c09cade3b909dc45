"""jbonsai_b200 — ctypes binding of the C ABI in ``include/jbonsai_b200.h``.

The product is the shared library ``jbonsai_b200/libjbonsai_b200.so`` (C++ host engine + CUDA
kernels for sm_100a).  This module is only the thin Python face of that C ABI, shaped like
the reference's public API so that tests read like the reference's own
(``Engine::load`` / ``Engine::synthesize`` / ``Engine::generator``, src/engine.rs:247-366;
``SpeechGenerator``, src/speech.rs:25-96).  There is no CPU fallback: when the library is
missing the import raises, and synthesis calls raise ``JbError`` when no GPU is usable.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Iterable, List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("JBONSAI_B200_LIB") or os.path.join(_HERE, "libjbonsai_b200.so")

# every symbol include/jbonsai_b200.h declares (checked by tests/test_capi_symbols.py)
C_ABI_SYMBOLS = [
    "jb_last_error", "jb_version", "jb_device_count", "jb_device_measure_peaks", "jb_engine_load",
    "jb_engine_load_from_bytes", "jb_engine_free", "jb_engine_set_condition",
    "jb_engine_get_condition", "jb_engine_set_interpolation_weight", "jb_engine_set_devices",
    "jb_engine_set_mlsa_variant", "jb_engine_set_mlsa_segmenting", "jb_engine_set_device_gather",
    "jb_engine_set_output_mode", "jb_engine_set_output_fractions", "jb_engine_set_pipeline_chunks", "jb_engine_synthesize", "jb_engine_synthesize_batch", "jb_engine_synthesize_batch_pcm16", "jb_free",
    "jb_engine_debug_batch", "jb_engine_noise", "jb_engine_durations", "jb_engine_tree_index",
    "jb_engine_stream_parameters", "jb_engine_duration_parameters", "jb_engine_gv_parameters",
    "jb_resident_prepare", "jb_resident_prepare_tracks", "jb_resident_rows", "jb_engine_vocoder_synthesize", "jb_resident_run", "jb_resident_fetch", "jb_resident_free",
    "jb_stream_open", "jb_stream_fperiod", "jb_stream_synthesized_frames", "jb_stream_step",
    "jb_stream_close", "jb_resident_device_output", "jb_engine_synthesize_batch_device", "jb_set_host_threads",
    "jb_trim_pools", "jb_device_pitch_period", "jb_engine_mlpg_create", "jb_mlpg_solve_noalloc",
    "jb_engine_set_mlsa_loud_warm", "jb_alloc_pinned", "jb_debug_plan_segments",
]

STATUS_NAMES = {
    0: "OK", -1: "EmptyVoice", -2: "MetadataError", -3: "Io", -4: "ParserError",
    -5: "ParseOptionError", -6: "JLabelParse", -7: "MissingLabel", -8: "FloatParse",
    -9: "LengthMismatch", -10: "WeightError", -11: "IndexNotFound", -12: "Argument",
    -13: "Cuda", -14: "Unsupported",
}


class VocoderConfig(C.Structure):
    """jb_vocoder_config: the arguments of Vocoder::new (src/vocoder/mod.rs:45-70)."""
    _fields_ = [("nmcp", C.c_size_t), ("nlpf", C.c_size_t), ("stage", C.c_size_t), ("use_log_gain", C.c_int),
                ("rate", C.c_size_t), ("alpha", C.c_double), ("beta", C.c_double), ("volume", C.c_double),
                ("fperiod", C.c_size_t), ("coefficient_rows", C.c_int)]


class JbError(RuntimeError):
    """Mirror of EngineError (src/engine.rs:16-28): ``code`` is the jb_status."""

    def __init__(self, code: int, msg: str):
        super().__init__(f"[{STATUS_NAMES.get(code, code)}] {msg}")
        self.code = code
        self.kind = STATUS_NAMES.get(code, str(code))


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C jbonsai_b200/csrc` "
            "(or __graft_entry__.build()). jbonsai_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    lib.jb_last_error.restype = C.c_char_p
    lib.jb_version.restype = C.c_char_p
    lib.jb_stream_fperiod.restype = C.c_size_t
    lib.jb_stream_fperiod.argtypes = [C.c_void_p]
    lib.jb_stream_synthesized_frames.restype = C.c_size_t
    lib.jb_stream_synthesized_frames.argtypes = [C.c_void_p]
    lib.jb_free.argtypes = [C.c_void_p]
    lib.jb_free.restype = None
    lib.jb_engine_free.argtypes = [C.c_void_p]
    lib.jb_engine_free.restype = None
    lib.jb_stream_close.argtypes = [C.c_void_p]
    lib.jb_stream_close.restype = None
    lib.jb_resident_free.argtypes = [C.c_void_p, C.c_void_p]
    lib.jb_resident_free.restype = None
    return lib


lib = _load()


def _check(rc: int) -> None:
    if rc != 0:
        raise JbError(rc, (lib.jb_last_error() or b"").decode("utf-8", "replace"))


def version() -> str:
    return lib.jb_version().decode()


def device_count() -> int:
    return int(lib.jb_device_count())


def measure_peaks(device: int = 0):
    """(FP64 FMA TFLOP/s, HBM copy GB/s) measured live on `device`."""
    t, g = C.c_double(), C.c_double()
    _check(lib.jb_device_measure_peaks(int(device), C.byref(t), C.byref(g)))
    return t.value, g.value


def device_pitch_period(lf0: np.ndarray, rate: float = 48000.0, device: int = 0) -> np.ndarray:
    """rate / exp(clamp(lf0)) as the vocoder's frame-prep kernel computes it (0 for NODATA)."""
    x = np.ascontiguousarray(lf0, dtype=np.float64)
    out = np.empty_like(x)
    D = C.POINTER(C.c_double)
    _check(lib.jb_device_pitch_period(int(device), x.ctypes.data_as(D), C.c_size_t(x.size), C.c_double(rate),
                                      out.ctypes.data_as(D)))
    return out


MLPG_NOALLOC_MAX_CELLS = 131072  # JB_MLPG_NOALLOC_MAX_CELLS (include/jbonsai_b200.h)


def mlpg_solve_noalloc(device: int, d_mean: int, d_vari: int, n_frames: int, dim: int, windows, d_out: int,
                       stream: int = 0) -> None:
    """jb_mlpg_solve_noalloc: MLPG solve of one non-MSD, non-GV utterance with no allocation (the work planes are a
    fixed __device__ array of the library). `d_mean`, `d_vari` ([n_frames][len(windows) * dim]) and `d_out`
    ([n_frames][dim]) are DEVICE addresses as integers, `stream` a cudaStream_t as an integer (0: default stream);
    the kernels are only enqueued. n_frames * dim <= MLPG_NOALLOC_MAX_CELLS. Calls that can overlap on one GPU
    race on the shared scratch."""
    widths = (C.c_size_t * len(windows))(*[len(w) for w in windows])
    coef = np.ascontiguousarray(np.concatenate([np.asarray(w, dtype=np.float64) for w in windows]))
    lib.jb_mlpg_solve_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                          C.POINTER(C.c_size_t), C.POINTER(C.c_double), C.c_void_p, C.c_void_p]
    _check(lib.jb_mlpg_solve_noalloc(int(device), C.c_void_p(int(d_mean)), C.c_void_p(int(d_vari)), int(n_frames),
                                     int(dim), len(windows), widths, coef.ctypes.data_as(C.POINTER(C.c_double)),
                                     C.c_void_p(int(d_out)), C.c_void_p(int(stream))))


class _PinnedOwner:
    def __init__(self, ptr):
        self.ptr = ptr

    def __del__(self):
        try:
            if self.ptr:
                lib.jb_free(C.c_void_p(self.ptr))
                self.ptr = None
        except Exception:
            pass


def pinned_array(shape, dtype=np.float64) -> np.ndarray:
    """A zeroed numpy array in page-locked host memory of the library's pool (jb_alloc_pinned), e.g. the buffer of
    Generator.generate_step; returned to the pool when the array is garbage-collected."""
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = C.c_void_p()
    _check(lib.jb_alloc_pinned(C.c_size_t(max(n, 8)), C.byref(p)))
    owner = _PinnedOwner(p.value)
    raw = (C.c_char * max(n, 8)).from_address(p.value)
    raw._jb_owner = owner  # the ctypes view keeps the allocation alive, numpy keeps the view
    arr = np.frombuffer(raw, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    arr[...] = 0
    return arr


def plan_segments(frames, warm: int = 40, warm_loud: int = 0, threshold: float = 2.0, gain=None, target: int = 7104):
    """jb_debug_plan_segments: [n_segments, 5] rows of (utt, begin, end, warm, prev) - the planner's time segments for
    utterances of `frames` frames each; `gain`: per-frame log gain (concatenated) for the warm-up by loudness."""
    fr = np.ascontiguousarray(frames, dtype=np.int64)
    g = None if gain is None else np.ascontiguousarray(gain, dtype=np.float64)
    gp = None if g is None else g.ctypes.data_as(C.POINTER(C.c_double))
    n = C.c_size_t()
    args = (fr.ctypes.data_as(C.POINTER(C.c_int64)), C.c_size_t(len(fr)), int(warm), int(warm_loud), C.c_double(threshold), gp,
            C.c_int64(int(target)))
    _check(lib.jb_debug_plan_segments(*args, None, C.c_size_t(0), C.byref(n)))
    out = np.zeros((n.value, 5), dtype=np.int32)
    _check(lib.jb_debug_plan_segments(*args, out.ctypes.data_as(C.POINTER(C.c_int32)), C.c_size_t(out.size), C.byref(n)))
    return out


def set_host_threads(n: int) -> None:
    """Host threads of the front-end / batching layer (0 = all cores): cores // ranks when ranks share a box."""
    _check(lib.jb_set_host_threads(int(n)))


def trim_pools() -> int:
    """Frees the cached pinned host buffers of the library; returns the bytes released."""
    lib.jb_trim_pools.restype = C.c_size_t
    return int(lib.jb_trim_pools())


def _lines(lines: Sequence[str]):
    enc = [l.encode("utf-8") if isinstance(l, str) else bytes(l) for l in lines]
    return (C.c_char_p * max(len(enc), 1))(*enc), len(enc)


def _flatten(utts: Sequence[Sequence[str]]):
    flat: List[str] = []
    off = [0]
    for u in utts:
        flat.extend(u)
        off.append(len(flat))
    arr, _ = _lines(flat)
    offs = (C.c_size_t * len(off))(*off)
    return arr, offs, len(utts)


def _take(ptr, n: int) -> np.ndarray:
    """Copy n doubles out of a library-owned buffer and release it."""
    if n:
        a = np.ctypeslib.as_array(ptr, shape=(n,)).copy()
    else:
        a = np.zeros(0, dtype=np.float64)
    lib.jb_free(ptr)
    return a


TIMING_KEYS = ["host_frontend_ms", "pack_ms", "h2d_ms", "kernels_ms", "d2h_ms", "mlpg_ms",
               "excitation_ms", "mlsa_ms", "h2d_bytes", "d2h_bytes", "frames", "samples",
               "voiced_samples", "launches", "mlsa_segments", "mlsa_seg_frames", "mlsa_cut_worst",
               "mlsa_fallback", "mlsa_repaired"]


class Engine:
    """``jbonsai::Engine`` (src/engine.rs:247-366) on B200."""

    def __init__(self, handle: C.c_void_p):
        self._h = handle

    # -- Engine::load / load_from_bytes (src/engine.rs:257-286)
    @classmethod
    def load(cls, paths: Iterable[str]) -> "Engine":
        arr, n = _lines([os.fspath(p) for p in paths])
        h = C.c_void_p()
        _check(lib.jb_engine_load(arr, C.c_size_t(n), C.byref(h)))
        return cls(h)

    @classmethod
    def load_from_bytes(cls, blobs: Sequence[bytes]) -> "Engine":
        n = len(blobs)
        keep = [C.create_string_buffer(bytes(b), len(b)) for b in blobs]
        ptrs = (C.c_void_p * max(n, 1))(*[C.cast(k, C.c_void_p) for k in keep])
        lens = (C.c_size_t * max(n, 1))(*[len(b) for b in blobs])
        h = C.c_void_p()
        _check(lib.jb_engine_load_from_bytes(ptrs, lens, C.c_size_t(n), C.byref(h)))
        return cls(h)

    def close(self) -> None:
        if self._h:
            lib.jb_engine_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- engine.condition (src/engine.rs:32-243)
    def set(self, key: str, value: float, idx: int = 0) -> None:
        _check(lib.jb_engine_set_condition(self._h, key.encode(), C.c_size_t(idx), C.c_double(value)))

    def get(self, key: str, idx: int = 0) -> float:
        v = C.c_double()
        _check(lib.jb_engine_get_condition(self._h, key.encode(), C.c_size_t(idx), C.byref(v)))
        return v.value

    def set_interpolation_weight(self, kind: str, stream: int, w: Sequence[float]) -> None:
        k = {"duration": 0, "parameter": 1, "gv": 2}[kind]
        arr = (C.c_double * len(w))(*w)
        _check(lib.jb_engine_set_interpolation_weight(self._h, k, C.c_size_t(stream), arr,
                                                      C.c_size_t(len(w))))

    def set_devices(self, devices: Sequence[int]) -> None:
        arr = (C.c_int * len(devices))(*devices)
        _check(lib.jb_engine_set_devices(self._h, arr, C.c_size_t(len(devices))))

    def set_mlsa_variant(self, v: int) -> None:
        _check(lib.jb_engine_set_mlsa_variant(self._h, int(v)))

    def set_device_gather(self, enabled: bool) -> None:
        _check(lib.jb_engine_set_device_gather(self._h, int(bool(enabled))))

    def set_mlsa_segmenting(self, warm_frames: int = 40, cut_tolerance: float = 1e-7) -> None:
        _check(lib.jb_engine_set_mlsa_segmenting(self._h, int(warm_frames), C.c_double(cut_tolerance)))

    def set_mlsa_loud_warm(self, loud_frames: int = 24, threshold: float = 2.0) -> None:
        """Warm-up of cuts that lie behind loud frames (0 = every cut gets warm_frames)."""
        _check(lib.jb_engine_set_mlsa_loud_warm(self._h, int(loud_frames), C.c_double(threshold)))

    def set_output_mode(self, mode: int) -> None:
        """1 (default): three filter waves, each copied under the next (10 + w: w equal waves); 2: zero-copy stores; 0: one copy."""
        _check(lib.jb_engine_set_output_mode(self._h, int(mode)))

    def set_output_fractions(self, fractions: Sequence[float]) -> None:
        arr = (C.c_double * len(fractions))(*fractions)
        _check(lib.jb_engine_set_output_fractions(self._h, arr, C.c_size_t(len(fractions))))

    def set_pipeline_chunks(self, chunks: int) -> None:
        _check(lib.jb_engine_set_pipeline_chunks(self._h, int(chunks)))

    # -- Engine::synthesize (src/engine.rs:294-296)
    def synthesize(self, lines: Sequence[str]) -> np.ndarray:
        arr, n = _lines(lines)
        out = C.POINTER(C.c_double)()
        ln = C.c_size_t()
        _check(lib.jb_engine_synthesize(self._h, arr, C.c_size_t(n), C.byref(out), C.byref(ln)))
        return _take(out, ln.value)

    def synthesize_batch(self, utts: Sequence[Sequence[str]], timings: Optional[dict] = None):
        """Returns (samples, out_off): utterance u = samples[out_off[u]:out_off[u+1]]."""
        arr, offs, n = _flatten(utts)
        out = C.POINTER(C.c_double)()
        out_off = (C.c_size_t * (n + 1))()
        tm = (C.c_double * 20)()
        _check(lib.jb_engine_synthesize_batch(self._h, arr, offs, C.c_size_t(n), C.byref(out),
                                              out_off, tm))
        off = np.array(list(out_off), dtype=np.int64)
        if timings is not None:
            timings.update({k: tm[i] for i, k in enumerate(TIMING_KEYS)})
        return _take(out, int(off[-1])), off

    def synthesize_batch_pcm16(self, utts: Sequence[Sequence[str]], timings: Optional[dict] = None):
        """16-bit PCM (clamp, truncate toward zero: examples/genji/main.rs:34-35) -> (int16 samples, out_off)."""
        arr, offs, n = _flatten(utts)
        out = C.POINTER(C.c_int16)()
        out_off = (C.c_size_t * (n + 1))()
        tm = (C.c_double * 20)()
        _check(lib.jb_engine_synthesize_batch_pcm16(self._h, arr, offs, C.c_size_t(n), C.byref(out), out_off, tm))
        off = np.array(list(out_off), dtype=np.int64)
        if timings is not None:
            timings.update({k: tm[i] for i, k in enumerate(TIMING_KEYS)})
        total = int(off[-1])
        res = np.ctypeslib.as_array(out, shape=(total,)).copy() if total else np.zeros(0, dtype=np.int16)
        lib.jb_free(out)
        return res, off

    def synthesize_batch_raw(self, flat, pcm16: bool = False):
        """Timed e2e call for bench.py: pre-flattened (arr, offs, n); returns the library-owned
        pinned host buffer (release with lib.jb_free) without copying it again."""
        arr, offs, n = flat
        out = C.POINTER(C.c_int16)() if pcm16 else C.POINTER(C.c_double)()
        out_off = (C.c_size_t * (n + 1))()
        tm = (C.c_double * 20)()
        fn = lib.jb_engine_synthesize_batch_pcm16 if pcm16 else lib.jb_engine_synthesize_batch
        _check(fn(self._h, arr, offs, C.c_size_t(n), C.byref(out), out_off, tm))
        return out, out_off, {k: tm[i] for i, k in enumerate(TIMING_KEYS)}

    # -- parity taps
    def debug_batch(self, utts: Sequence[Sequence[str]]) -> dict:
        arr, offs, n = _flatten(utts)
        P = C.POINTER(C.c_double)
        mcp, lf0, lpf, pulse, exc = P(), P(), P(), P(), P()
        foff = (C.c_int64 * (n + 1))()
        _check(lib.jb_engine_debug_batch(self._h, arr, offs, C.c_size_t(n), C.byref(mcp), C.byref(lf0),
                                         C.byref(lpf), C.byref(pulse), C.byref(exc), foff))
        fo = np.array(list(foff), dtype=np.int64)
        TF = int(fo[-1])
        fp = int(self.get("fperiod"))
        return {
            "frame_off": fo,
            "mcp": _take(mcp, TF * 35).reshape(TF, 35),
            "lf0": _take(lf0, TF),
            "lpf": _take(lpf, TF * 31).reshape(TF, 31),
            "pulse": _take(pulse, TF * fp),
            "exc": _take(exc, TF * fp),
        }

    def noise(self, n: int) -> np.ndarray:
        out = np.zeros(n, dtype=np.float64)
        _check(lib.jb_engine_noise(self._h, C.c_size_t(n), out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    # -- host front-end taps
    def durations(self, lines: Sequence[str]) -> np.ndarray:
        arr, n = _lines(lines)
        cap = max(n * 16, 16)
        d = np.zeros(cap, dtype=np.int32)
        ns = C.c_size_t()
        _check(lib.jb_engine_durations(self._h, arr, C.c_size_t(n), d.ctypes.data_as(C.POINTER(C.c_int32)),
                                       C.c_size_t(cap), C.byref(ns)))
        return d[: ns.value].copy()

    def tree_index(self, voice: int, model: str, stream: int, state_index: int, label: str):
        m = {"duration": 0, "stream": 1, "gv": 2}[model]
        t, p = C.c_long(), C.c_long()
        _check(lib.jb_engine_tree_index(self._h, C.c_size_t(voice), m, C.c_size_t(stream),
                                        C.c_size_t(state_index), label.encode(), C.byref(t), C.byref(p)))
        return t.value, p.value

    def stream_parameters(self, lines: Sequence[str], stream: int):
        arr, n = _lines(lines)
        cap = max(n, 1) * 16 * 128
        mean = np.zeros(cap)
        vari = np.zeros(cap)
        msd = np.zeros(cap)
        ns, w = C.c_size_t(), C.c_size_t()
        D = C.POINTER(C.c_double)
        _check(lib.jb_engine_stream_parameters(self._h, arr, C.c_size_t(n), C.c_size_t(stream),
                                               mean.ctypes.data_as(D), vari.ctypes.data_as(D),
                                               msd.ctypes.data_as(D), C.c_size_t(cap), C.byref(ns), C.byref(w)))
        S, W = ns.value, w.value
        return mean[: S * W].reshape(S, W).copy(), vari[: S * W].reshape(S, W).copy(), msd[:S].copy()

    def duration_parameters(self, lines: Sequence[str]):
        arr, n = _lines(lines)
        cap = max(n, 1) * 16
        mean, vari = np.zeros(cap), np.zeros(cap)
        ns = C.c_size_t()
        D = C.POINTER(C.c_double)
        _check(lib.jb_engine_duration_parameters(self._h, arr, C.c_size_t(n), mean.ctypes.data_as(D),
                                                 vari.ctypes.data_as(D), C.c_size_t(cap), C.byref(ns)))
        return mean[: ns.value].copy(), vari[: ns.value].copy()

    def gv_parameters(self, lines: Sequence[str], stream: int):
        arr, n = _lines(lines)
        cap = max(n * 16, 64)
        mean, vari = np.zeros(cap), np.zeros(cap)
        sw = np.zeros(cap, dtype=np.uint8)
        dim, ns, has = C.c_size_t(), C.c_size_t(), C.c_int()
        D = C.POINTER(C.c_double)
        _check(lib.jb_engine_gv_parameters(self._h, arr, C.c_size_t(n), C.c_size_t(stream),
                                           mean.ctypes.data_as(D), vari.ctypes.data_as(D),
                                           sw.ctypes.data_as(C.POINTER(C.c_uint8)), C.c_size_t(cap),
                                           C.byref(dim), C.byref(ns), C.byref(has)))
        if not has.value:
            return None
        return mean[: dim.value].copy(), vari[: dim.value].copy(), sw[: ns.value].astype(bool)

    # -- resident batch (bench "value": inputs already in HBM)
    def prepare(self, utts: Sequence[Sequence[str]], device_index: int = 0) -> "Resident":
        arr, offs, n = _flatten(utts)
        r = C.c_void_p()
        _check(lib.jb_resident_prepare(self._h, arr, offs, C.c_size_t(n), int(device_index), C.byref(r)))
        return Resident(self, r, n)

    def synthesize_batch_device(self, utts: Sequence[Sequence[str]], device_index: int = 0):
        """Engine::synthesize for a batch whose Vec<f64> stays in HBM: (Resident to keep alive, device pointer,
        out_off)."""
        arr, offs, n = _flatten(utts)
        r, ptr = C.c_void_p(), C.c_void_p()
        out_off = (C.c_size_t * (n + 1))()
        _check(lib.jb_engine_synthesize_batch_device(self._h, arr, offs, C.c_size_t(n), int(device_index), C.byref(r),
                                                     C.byref(ptr), out_off, None))
        return Resident(self, r, n), int(ptr.value or 0), np.array(list(out_off), dtype=np.int64)

    # -- the MLPG seam: MlpgAdjust::new + create (src/mlpg_adjust/mod.rs:31-95) for one stream
    def mlpg_create(self, durations, dim, windows, mean, vari, msd=None, msd_threshold=0.5, gv=None, gv_weight=1.0,
                    device_index: int = 0):
        """State-level parameters -> [frames, dim] trajectories. `durations`, `mean`, `vari`, `msd` are one array
        per utterance (or a single utterance's arrays); windows: coefficient lists, static window first;
        gv: (mean [dim], vari [dim], switch [states]) per utterance or None. Returns a list of [T, dim] arrays
        (one array for a single utterance)."""
        single = not isinstance(durations, (list, tuple))
        if single:
            durations, mean, vari = [durations], [mean], [vari]
            msd = None if msd is None else [msd]
            gv = None if gv is None else [gv]
        n = len(durations)
        dur = [np.ascontiguousarray(d, dtype=np.int32).reshape(-1) for d in durations]
        soff = np.zeros(n + 1, dtype=np.uint64)
        soff[1:] = np.cumsum([len(d) for d in dur])
        S, L = int(soff[-1]), dim * len(windows)
        cat = lambda xs, dt=np.float64: np.ascontiguousarray(np.concatenate(
            [np.asarray(x, dtype=dt).reshape(-1) for x in xs])) if xs and S else np.zeros(1, dtype=dt)
        d_all, m_all, v_all = cat(dur, np.int32), cat(mean), cat(vari)
        if S and (m_all.size != S * L or v_all.size != S * L):
            raise ValueError("mean / vari must hold states x (windows x dim) values")
        widths = (C.c_size_t * len(windows))(*[len(w) for w in windows])
        coef = np.ascontiguousarray(np.concatenate([np.asarray(w, dtype=np.float64) for w in windows]))
        D = C.POINTER(C.c_double)
        inp = MlpgInput()
        inp.n_utt, inp.dim = n, dim
        inp.state_off = soff.ctypes.data_as(C.POINTER(C.c_size_t))
        inp.durations = d_all.ctypes.data_as(C.POINTER(C.c_int32))
        inp.n_windows, inp.window_widths = len(windows), widths
        inp.window_coefficients = coef.ctypes.data_as(D)
        inp.mean, inp.vari = m_all.ctypes.data_as(D), v_all.ctypes.data_as(D)
        keep = []
        if msd is not None:
            ms = cat(msd)
            keep.append(ms)
            inp.msd = ms.ctypes.data_as(D)
        inp.msd_threshold, inp.gv_weight = float(msd_threshold), float(gv_weight)
        if gv is not None:
            gm, gvv, gs = cat([g[0] for g in gv]), cat([g[1] for g in gv]), cat([g[2] for g in gv], np.uint8)
            keep += [gm, gvv, gs]
            inp.gv_mean, inp.gv_vari = gm.ctypes.data_as(D), gvv.ctypes.data_as(D)
            inp.gv_switch = gs.ctypes.data_as(C.POINTER(C.c_uint8))
        out = D()
        foff = (C.c_size_t * (n + 1))()
        _check(lib.jb_engine_mlpg_create(self._h, C.byref(inp), int(device_index), C.byref(out), foff))
        fo = list(foff)
        par = _take(out, fo[-1] * dim).reshape(fo[-1], dim)
        res = [par[fo[u]:fo[u + 1]] for u in range(n)]
        return res[0] if single else res

    # -- the vocoder seam: Vocoder::new + SpeechGenerator::generate_all over parameter tracks
    @staticmethod
    def _tracks(spectrum, lf0, lpf, stage, use_log_gain, rate, alpha, beta, volume, fperiod, first_rows=None):
        """-> (config struct, n_utt, frame_off array, the three concatenated tracks)."""
        if isinstance(spectrum, np.ndarray) and spectrum.ndim == 2:
            spectrum, lf0, lpf = [spectrum], [lf0], [lpf]
        spec = [np.ascontiguousarray(s, dtype=np.float64) for s in spectrum]
        nmcp = spec[0].shape[1] if spec else 0
        lf = [np.ascontiguousarray(x, dtype=np.float64).reshape(-1) for x in lf0]
        lp = [np.ascontiguousarray(x, dtype=np.float64) for x in lpf]
        nlpf = max([x.shape[1] for x in lp if x.ndim == 2] + [0])
        lp = [x.reshape(len(s), nlpf) for x, s in zip(lp, spec)]
        for s, a, b in zip(spec, lf, lp):
            if len(a) != len(s) or len(b) != len(s):  # speech.rs:32-34
                raise ValueError("The length of spectrum, lf0, and lpf must be the same.")
        off = np.zeros(len(spec) + 1, dtype=np.uint64)
        off[1:] = np.cumsum([len(s) for s in spec])
        cat = lambda xs, w: np.ascontiguousarray(np.concatenate(xs).reshape(-1)) if xs and w else np.zeros(1)
        cfg = VocoderConfig(nmcp, nlpf, int(stage), int(bool(use_log_gain)), int(rate), float(alpha), float(beta),
                            float(volume), int(fperiod), int(first_rows is not None))
        if first_rows is not None:  # test seam: spectrum = coefficient rows, then the first rows
            spec = spec + [np.ascontiguousarray(first_rows, dtype=np.float64).reshape(len(spec), nmcp)]
        return cfg, len(lf), off, cat(spec, nmcp), cat(lf, 1), cat(lp, nlpf)

    def vocoder_synthesize(self, spectrum, lf0, lpf, stage=0, use_log_gain=False, rate=48000, alpha=0.55,
                           beta=0.0, volume=1.0, fperiod=240):
        """``SpeechGenerator::new(fperiod, Vocoder::new(...), spectrum, lf0, lpf).generate_all()``
        (src/speech.rs:25-96, src/vocoder/mod.rs:45-70) for one track set ([T, nmcp], [T], [T, nlpf]) or a
        list of them; ``stage > 0`` takes LSP frames. Returns one waveform or a list of waveforms."""
        single = isinstance(spectrum, np.ndarray) and spectrum.ndim == 2
        cfg, n, off, sp, lf, lp = self._tracks(spectrum, lf0, lpf, stage, use_log_gain, rate, alpha, beta, volume, fperiod)
        D = C.POINTER(C.c_double)
        out = D()
        out_off = (C.c_size_t * (n + 1))()
        _check(lib.jb_engine_vocoder_synthesize(self._h, C.byref(cfg), C.c_size_t(n),
                                                off.ctypes.data_as(C.POINTER(C.c_size_t)), sp.ctypes.data_as(D),
                                                lf.ctypes.data_as(D), lp.ctypes.data_as(D), C.byref(out), out_off))
        o = list(out_off)
        w = _take(out, int(o[-1]))
        waves = [w[o[i]:o[i + 1]] for i in range(n)]
        return waves[0] if single else waves

    def prepare_tracks(self, spectrum, lf0, lpf, device_index: int = 0, **kw) -> "Resident":
        k = dict(stage=0, use_log_gain=False, rate=48000, alpha=0.55, beta=0.0, volume=1.0, fperiod=240, first_rows=None)
        k.update(kw)
        cfg, n, off, sp, lf, lp = self._tracks(spectrum, lf0, lpf, **k)
        D = C.POINTER(C.c_double)
        r = C.c_void_p()
        _check(lib.jb_resident_prepare_tracks(self._h, C.byref(cfg), C.c_size_t(n),
                                              off.ctypes.data_as(C.POINTER(C.c_size_t)), sp.ctypes.data_as(D),
                                              lf.ctypes.data_as(D), lp.ctypes.data_as(D), int(device_index), C.byref(r)))
        return Resident(self, r, n)

    # -- Engine::generator (src/engine.rs:301-366), batched
    def generator(self, utts: Sequence[Sequence[str]], device: int = 0) -> "SpeechGenerator":
        if utts and isinstance(utts[0], str):
            utts = [utts]  # a single utterance, like the reference's signature
        arr, offs, n = _flatten(utts)
        s = C.c_void_p()
        _check(lib.jb_stream_open(self._h, arr, offs, C.c_size_t(n), int(device), C.byref(s)))
        return SpeechGenerator(self, s, n)


class MlpgInput(C.Structure):
    _fields_ = [("n_utt", C.c_size_t), ("dim", C.c_size_t), ("state_off", C.POINTER(C.c_size_t)),
                ("durations", C.POINTER(C.c_int32)), ("n_windows", C.c_size_t), ("window_widths", C.POINTER(C.c_size_t)),
                ("window_coefficients", C.POINTER(C.c_double)), ("mean", C.POINTER(C.c_double)),
                ("vari", C.POINTER(C.c_double)), ("msd", C.POINTER(C.c_double)), ("msd_threshold", C.c_double),
                ("gv_mean", C.POINTER(C.c_double)), ("gv_vari", C.POINTER(C.c_double)),
                ("gv_switch", C.POINTER(C.c_uint8)), ("gv_weight", C.c_double)]


class Resident:
    def __init__(self, engine: Engine, handle, n_utt: int):
        self._e, self._h, self.n_utt = engine, handle, n_utt

    def run(self) -> dict:
        ms = C.c_double()
        tm = (C.c_double * 20)()
        _check(lib.jb_resident_run(self._e._h, self._h, C.byref(ms), tm))
        d = {k: tm[i] for i, k in enumerate(TIMING_KEYS)}
        d["device_ms"] = ms.value
        return d

    def rows(self, frames: int, nmcp: int):
        """Filter coefficient rows of the last run: ([frames, nmcp], [n_utt, nmcp] first rows)."""
        buf = np.zeros((frames + self.n_utt) * nmcp)
        _check(lib.jb_resident_rows(self._e._h, self._h, buf.ctypes.data_as(C.POINTER(C.c_double)), C.c_size_t(buf.size)))
        return buf[:frames * nmcp].reshape(frames, nmcp), buf[frames * nmcp:].reshape(self.n_utt, nmcp)

    def device_output(self):
        """(device pointer, out_off, CUDA device) of the last run's waveform: it stays in HBM (no D2H copy)."""
        ptr = C.c_void_p()
        out_off = (C.c_size_t * (self.n_utt + 1))()
        dev = C.c_int()
        _check(lib.jb_resident_device_output(self._e._h, self._h, C.byref(ptr), out_off, C.byref(dev)))
        return int(ptr.value or 0), np.array(list(out_off), dtype=np.int64), dev.value

    def fetch(self):
        out = C.POINTER(C.c_double)()
        out_off = (C.c_size_t * (self.n_utt + 1))()
        _check(lib.jb_resident_fetch(self._e._h, self._h, C.byref(out), out_off))
        off = np.array(list(out_off), dtype=np.int64)
        return _take(out, int(off[-1])), off

    def close(self):
        if self._h:
            lib.jb_resident_free(self._e._h, self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class SpeechGenerator:
    """``SpeechGenerator`` (src/speech.rs:25-96) for a batch of utterances."""

    def __init__(self, engine: Engine, handle, n_utt: int):
        self._e, self._h, self.n_utt = engine, handle, n_utt

    def fperiod(self) -> int:
        return int(lib.jb_stream_fperiod(self._h))

    def synthesized_frames(self) -> int:
        return int(lib.jb_stream_synthesized_frames(self._h))

    def generate_step(self, n_frames: int = 1, buf: Optional[np.ndarray] = None):
        """Returns (buf [n_utt, n_frames*fperiod], written [n_utt], device_ms)."""
        fp = self.fperiod()
        if buf is None:
            buf = pinned_array((self.n_utt, n_frames * fp))  # page-locked: the GPU writes it directly
        written = (C.c_size_t * max(self.n_utt, 1))()
        ms = C.c_double()
        _check(lib.jb_stream_step(self._h, C.c_size_t(n_frames), buf.ctypes.data_as(C.POINTER(C.c_double)),
                                  C.c_size_t(buf.size), written, C.byref(ms)))
        return buf, np.array(list(written)[: self.n_utt], dtype=np.int64), ms.value

    def generate_all(self, n_frames: int = 64) -> List[np.ndarray]:
        parts: List[List[np.ndarray]] = [[] for _ in range(self.n_utt)]
        while True:
            buf, written, _ = self.generate_step(n_frames)
            if not written.any():
                break
            for u in range(self.n_utt):
                if written[u]:
                    parts[u].append(buf[u, : written[u]].copy())
        return [np.concatenate(p) if p else np.zeros(0) for p in parts]

    def close(self):
        if self._h:
            lib.jb_stream_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

from . import workload  # noqa: E402,F401  (synthetic label batches for bench/tests)
