"""BASELINE.json configs other than the headline one (bench.py measures config 3), one JSON line each.

  config 1  nitech, the 8-label README example, single utterance: latency of Engine.synthesize
  config 2  one long utterance (examples/genji/genji.lab stands in for the Constitution preamble, which
            is not in the reference repo) with nitech and tohoku-f01-neutral, on 1 GPU
  config 4  tohoku-f01-neutral (no GV), 16384 / 8 = 2048 synthetic utterances per GPU: per-GPU shard
  config 5  streaming: 4096 concurrent utterances stepped in chunks of 1 / 4 / 20 frames
            (SpeechGenerator::generate_step for all of them per call): per-chunk latency and throughput

Usage (GPU box): python tools/bench_configs.py [--quick] > gpurun_out/configs.jsonl
Times are host wall-clock around the C-ABI call (host buffers in and out) unless a key says device."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import jbonsai_b200 as jb  # noqa: E402
from jbonsai_b200 import workload as wl  # noqa: E402
from labels import SAMPLE_SENTENCE_1  # noqa: E402

RATE = 48000.0


def emit(**kw):
    print(json.dumps(kw), flush=True)


def timed(fn, reps, warm=2):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t0) * 1e3)
    return ts


def raw_call(e, utts):
    """jb_engine_synthesize_batch from label strings to the library-owned pinned host buffer (the
    same call bench.py times as e2e); returns the timing dict of the last call."""
    from jbonsai_b200 import _flatten
    flat = _flatten(utts)
    box = {}

    def call():
        out, _, tm = e.synthesize_batch_raw(flat)
        jb.lib.jb_free(out)
        box.update(tm)
    return call, box


def config1(reps):
    e = jb.Engine.load([wl.VOICES["nitech"]])
    n = len(e.synthesize(SAMPLE_SENTENCE_1))
    ts = timed(lambda: e.synthesize(SAMPLE_SENTENCE_1), reps)
    emit(config=1, workload="nitech, 8-label README example, 1 utterance", samples=n, audio_s=n / RATE,
         ms_p50=float(np.median(ts)), ms_min=float(min(ts)), audio_s_per_s=n / RATE / (np.median(ts) * 1e-3))


def config2(reps):
    lab = wl.genji_labels()
    for voice in ("nitech", "tohoku-f01-neutral"):
        e = jb.Engine.load([wl.VOICES[voice]])
        call, tm = raw_call(e, [lab])
        ts = timed(call, reps, warm=2)
        tm = dict(tm)
        n = int(tm["samples"])
        e.set_mlsa_variant(2)  # the sequential kernel: what one utterance costs without time segments
        call2, _ = raw_call(e, [lab])
        tseq = timed(call2, 1, warm=1)
        emit(config=2, workload=f"{voice}, genji.lab as one utterance ({len(lab)} labels)", samples=n,
             audio_s=n / RATE, ms_p50=float(np.median(ts)), audio_s_per_s=n / RATE / (np.median(ts) * 1e-3),
             breakdown_ms={k: tm[k] for k in ("host_frontend_ms", "pack_ms", "h2d_ms", "kernels_ms", "mlsa_ms", "d2h_ms") if k in tm},
             mlsa_segments=tm.get("mlsa_segments"), cut_check_worst=tm.get("mlsa_cut_worst"),
             sequential_kernel_ms=float(tseq[0]))


def config4(n_utt, reps):
    voice = "tohoku-f01-neutral"
    e = jb.Engine.load([wl.VOICES[voice]])
    utts = wl.timed_batch(e, n_utt, wl.SEED_TOHOKU)
    r = e.prepare(utts)
    r.run()
    dev = [r.run() for _ in range(reps)]
    r.close()
    call, tm = raw_call(e, utts)
    ts = timed(call, reps, warm=2)
    n = int(dev[0]["samples"])
    dms = float(np.median([d["device_ms"] for d in dev]))
    emit(config=4, workload=f"{voice} (no GV), {n_utt} synthetic utterances (~5 s) = one GPU's shard of 16384 / 8",
         samples=n, audio_s=n / RATE, device_ms=dms, device_audio_s_per_s=n / RATE / (dms * 1e-3),
         device_breakdown_ms={k: float(np.median([d[k] for d in dev])) for k in ("mlpg_ms", "excitation_ms", "mlsa_ms") if k in dev[0]},
         e2e_ms=float(np.median(ts)), e2e_audio_s_per_s=n / RATE / (np.median(ts) * 1e-3),
         e2e_breakdown_ms={k: tm[k] for k in ("host_frontend_ms", "pack_ms", "h2d_ms", "kernels_ms", "d2h_ms") if k in tm})


def config5(n_utt, max_chunks):
    e = jb.Engine.load([wl.VOICES["nitech"]])
    utts = wl.timed_batch(e, n_utt, wl.SEED_NITECH)
    lib = jb.lib
    for k, pinned in ((1, True), (4, True), (20, True), (1, False), (20, False)):
        t0 = time.perf_counter()
        g = e.generator(utts)  # MLPG for every utterance up front, like Engine::generator
        open_ms = (time.perf_counter() - t0) * 1e3
        fp = g.fperiod()
        # the step buffer: page-locked memory of the library (jb_alloc_pinned) or an ordinary pageable array
        buf = jb.pinned_array((n_utt, k * fp)) if pinned else np.zeros((n_utt, k * fp))
        written = (C.c_size_t * n_utt)()
        ms = C.c_double()
        wall, dev = [], []
        # every utterance has at least 800 frames (timed_batch): time only steps in which all streams are live
        n_timed = min(max_chunks, 800 // k - 3)
        for i in range(n_timed + 3):
            t0 = time.perf_counter()
            rc = lib.jb_stream_step(g._h, C.c_size_t(k), buf.ctypes.data_as(C.POINTER(C.c_double)),
                                    C.c_size_t(buf.size), written, C.byref(ms))
            t1 = time.perf_counter()
            assert rc == 0
            if i >= 3:
                wall.append((t1 - t0) * 1e3)
                dev.append(ms.value)
        g.close()
        wall, dev = np.array(wall), np.array(dev)
        chunk_audio = n_utt * k * fp / RATE
        emit(config=5, workload=f"streaming, {n_utt} concurrent nitech utterances, chunk = {k} frame(s) = {k * 5} ms of audio each, "
                                f"{'page-locked' if pinned else 'pageable'} step buffer",
             chunks_timed=len(wall), open_ms=open_ms, chunk_wall_ms_p50=float(np.percentile(wall, 50)),
             chunk_wall_ms_p99=float(np.percentile(wall, 99)), chunk_device_ms_p50=float(np.percentile(dev, 50)),
             audio_s_per_s=chunk_audio / (float(np.mean(wall)) * 1e-3),
             realtime_streams=float(k * fp / RATE / (np.mean(wall) * 1e-3) * n_utt))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--only", type=int, default=0)
    a = ap.parse_args()
    if jb.device_count() < 1:
        raise SystemExit("needs a CUDA device")
    reps = 3 if a.quick else 7
    if a.only in (0, 1):
        config1(20 if a.quick else 100)
    if a.only in (0, 2):
        config2(reps)
    if a.only in (0, 4):
        config4(256 if a.quick else 2048, reps)
    if a.only in (0, 5):
        config5(256 if a.quick else 4096, 30 if a.quick else 200)


if __name__ == "__main__":
    main()
