"""ctypes face of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY (see oracle/oracle.h)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_L = C.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
_L.orc_last_error.restype = C.c_char_p
_L.orc_free_buf.argtypes = [C.c_void_p]
_L.orc_free.argtypes = [C.c_void_p]
D = C.POINTER(C.c_double)


def _lines(lines):
    enc = [l.encode("utf-8") for l in lines]
    return (C.c_char_p * max(len(enc), 1))(*enc), len(enc)


def _take(p, n):
    a = np.ctypeslib.as_array(p, shape=(n,)).copy() if n else np.zeros(0)
    _L.orc_free_buf(p)
    return a


class OracleError(RuntimeError):
    def __init__(self, code):
        super().__init__((_L.orc_last_error() or b"").decode())
        self.code = code


def _ck(rc):
    if rc != 0:
        raise OracleError(rc)


class Oracle:
    def __init__(self, paths):
        arr, n = _lines(paths)
        self.h = C.c_void_p()
        _ck(_L.orc_load(arr, C.c_size_t(n), C.byref(self.h)))

    def set(self, key, value, idx=0):
        _ck(_L.orc_set_condition(self.h, key.encode(), C.c_size_t(idx), C.c_double(value)))

    def set_weights(self, kind, stream, w):
        k = {"duration": 0, "parameter": 1, "gv": 2}[kind]
        _ck(_L.orc_set_weights(self.h, k, C.c_size_t(stream), (C.c_double * len(w))(*w), C.c_size_t(len(w))))

    def synthesize(self, lines, fma=False):
        arr, n = _lines(lines)
        out, ln = D(), C.c_size_t()
        _ck(_L.orc_synthesize(self.h, arr, C.c_size_t(n), int(fma), C.byref(out), C.byref(ln)))
        return _take(out, ln.value)

    def synthesize_tap(self, lines, fma=False):
        arr, n = _lines(lines)
        out, exc, pulse, gained, ln = D(), D(), D(), D(), C.c_size_t()
        _ck(_L.orc_synthesize_tap(self.h, arr, C.c_size_t(n), int(fma), C.byref(out), C.byref(exc),
                                  C.byref(pulse), C.byref(gained), C.byref(ln)))
        return {"wave": _take(out, ln.value), "exc": _take(exc, ln.value),
                "pulse": _take(pulse, ln.value), "gained": _take(gained, ln.value)}

    def parameters(self, lines):
        arr, n = _lines(lines)
        mcp, lf0, lpf, T = D(), D(), D(), C.c_size_t()
        _ck(_L.orc_parameters(self.h, arr, C.c_size_t(n), C.byref(mcp), C.byref(lf0), C.byref(lpf), C.byref(T)))
        t = T.value
        return _take(mcp, t * 35).reshape(t, 35), _take(lf0, t), _take(lpf, t * 31).reshape(t, 31)

    def synthesize_batch(self, utts, fma=False, nthreads=1):
        flat, off = [], [0]
        for u in utts:
            flat.extend(u)
            off.append(len(flat))
        arr, _ = _lines(flat)
        offs = (C.c_size_t * len(off))(*off)
        out = D()
        out_off = (C.c_size_t * (len(utts) + 1))()
        _ck(_L.orc_synthesize_batch(self.h, arr, offs, C.c_size_t(len(utts)), int(fma), int(nthreads),
                                    C.byref(out), out_off))
        o = np.array(list(out_off), dtype=np.int64)
        return _take(out, int(o[-1])), o

    def close(self):
        if self.h:
            _L.orc_free(self.h)
            self.h = None


def pitch_period(lf0, rate=48000.0):
    x = np.ascontiguousarray(lf0, dtype=np.float64)
    out = np.empty_like(x)
    _L.orc_pitch_period.restype = None
    _L.orc_pitch_period(x.ctypes.data_as(D), C.c_size_t(x.size), C.c_double(rate), out.ctypes.data_as(D))
    return out


def noise(n):
    out = np.zeros(n)
    _L.orc_noise(C.c_size_t(n), out.ctypes.data_as(D))
    return out


def vocoder(mcp, lf0, lpf, rate=48000, fperiod=240, alpha=0.55, beta=0.0, volume=1.0, fma=False, tap=False):
    mcp = np.ascontiguousarray(mcp, dtype=np.float64)
    lf0 = np.ascontiguousarray(lf0, dtype=np.float64)
    lpf = np.ascontiguousarray(lpf, dtype=np.float64)
    T = mcp.shape[0]
    out = np.zeros(T * fperiod)
    exc = np.zeros(T * fperiod) if tap else None
    _ck(_L.orc_vocoder(C.c_size_t(T), C.c_size_t(mcp.shape[1]), C.c_size_t(lpf.shape[1] if lpf.ndim == 2 else 0),
                       mcp.ctypes.data_as(D), lf0.ctypes.data_as(D), lpf.ctypes.data_as(D), C.c_size_t(rate),
                       C.c_size_t(fperiod), C.c_double(alpha), C.c_double(beta), C.c_double(volume), int(fma),
                       out.ctypes.data_as(D), exc.ctypes.data_as(D) if tap else None))
    return (out, exc) if tap else out


def vocoder_ex(spectrum, lf0, lpf, stage=0, use_log_gain=False, rate=48000, fperiod=240, alpha=0.55, beta=0.0,
               volume=1.0, fma=False, tap=False):
    """Vocoder::new(nmcp, nlpf, stage, use_log_gain, ...) + SpeechGenerator::generate_all; stage > 0 takes
    LSP frames (gain in column 0). lpf with zero columns = a voice without LPF stream (excitation.rs:87-99)."""
    spectrum = np.ascontiguousarray(spectrum, dtype=np.float64)
    lf0 = np.ascontiguousarray(lf0, dtype=np.float64)
    lpf = np.ascontiguousarray(lpf, dtype=np.float64)
    T = spectrum.shape[0]
    nlpf = lpf.shape[1] if lpf.ndim == 2 else 0
    out = np.zeros(T * fperiod)
    exc = np.zeros(T * fperiod) if tap else None
    _ck(_L.orc_vocoder_ex(C.c_size_t(T), C.c_size_t(spectrum.shape[1]), C.c_size_t(nlpf), C.c_size_t(stage),
                          int(use_log_gain), spectrum.ctypes.data_as(D), lf0.ctypes.data_as(D),
                          lpf.ctypes.data_as(D) if nlpf else None, C.c_size_t(rate), C.c_size_t(fperiod),
                          C.c_double(alpha), C.c_double(beta), C.c_double(volume), int(fma),
                          out.ctypes.data_as(D), exc.ctypes.data_as(D) if tap else None))
    return (out, exc) if tap else out


def lsp_coefficients(frame, stage, alpha=0.55, use_log_gain=False, beta=0.0, filtered=True):
    """Stage::NonZero frame transform (vocoder/mod.rs:92-107 / 148-155): LSP frame -> filter coefficients."""
    v = np.ascontiguousarray(frame, dtype=np.float64)
    out = np.zeros(len(v))
    _ck(_L.orc_lsp_coefficients(v.ctypes.data_as(D), C.c_size_t(len(v)), C.c_double(alpha), int(use_log_gain),
                                C.c_size_t(stage), C.c_double(beta), int(filtered), out.ctypes.data_as(D)))
    return out


def postfilter(c, alpha=0.55, beta=0.3):
    """MelCepstrum::postfilter_mcp (cepstrum.rs:23-37) -> (c', energy before, energy after)."""
    v = np.array(c, dtype=np.float64)
    e0, e1 = C.c_double(), C.c_double()
    _L.orc_postfilter.restype = None
    _L.orc_postfilter(v.ctypes.data_as(D), C.c_size_t(len(v)), C.c_double(alpha), C.c_double(beta),
                      C.byref(e0), C.byref(e1))
    return v, e0.value, e1.value


def mlpg(durations, dim, windows, mean, vari, msd, msd_threshold=0.5, gv=None, gv_weight=1.0):
    """windows: list of coefficient lists. gv: (mean[dim], vari[dim], switch[S]) or None."""
    durations = np.ascontiguousarray(durations, dtype=np.int32)
    S = len(durations)
    mean = np.ascontiguousarray(mean, dtype=np.float64)
    vari = np.ascontiguousarray(vari, dtype=np.float64)
    msd = np.ascontiguousarray(msd, dtype=np.float64)
    wl = (C.c_size_t * len(windows))(*[len(w) for w in windows])
    wc = np.ascontiguousarray(np.concatenate([np.asarray(w, dtype=np.float64) for w in windows]))
    T = int(durations.sum())
    out = np.zeros(T * dim)
    if gv is not None:
        gm = np.ascontiguousarray(gv[0], dtype=np.float64)
        gvv = np.ascontiguousarray(gv[1], dtype=np.float64)
        gs = np.ascontiguousarray(gv[2], dtype=np.uint8)
    else:
        gm = gvv = np.zeros(dim)
        gs = np.zeros(S, dtype=np.uint8)
    _ck(_L.orc_mlpg(C.c_size_t(S), durations.ctypes.data_as(C.POINTER(C.c_int32)), C.c_size_t(dim),
                    C.c_size_t(len(windows)), wl, wc.ctypes.data_as(D), mean.ctypes.data_as(D),
                    vari.ctypes.data_as(D), msd.ctypes.data_as(D), C.c_double(msd_threshold),
                    int(gv is not None), gm.ctypes.data_as(D), gvv.ctypes.data_as(D),
                    gs.ctypes.data_as(C.POINTER(C.c_uint8)), C.c_double(gv_weight), out.ctypes.data_as(D)))
    return out.reshape(T, dim)


def boundary_distances(mask):
    m = np.ascontiguousarray(mask, dtype=np.uint8)
    n = len(m)
    l = (C.c_size_t * max(n, 1))()
    r = (C.c_size_t * max(n, 1))()
    _L.orc_boundary_distances(m.ctypes.data_as(C.POINTER(C.c_uint8)), C.c_size_t(n), l, r)
    return list(zip(list(l)[:n], list(r)[:n]))
