#!/usr/bin/env python
"""bench.py — audio-seconds synthesised per second (48 kHz, batched) on N B200s.

One "step" = one pass of the hot path (MLPG solve + mel-cepstral vocoder) over one batch of
synthetic full-context label sequences (BASELINE.json config 3: 1024 utterances
of ~5 s, voice nitech m001, per GPU; weak scaling: every rank owns its own 1024 utterances,
no collective on the data path).

  value         kernels only, packed inputs already resident in HBM, CUDA events on the
                launching stream inside the library (jb_resident_run), max over ranks.
  e2e           through the reference-facing call (jb_engine_synthesize_batch = batched
                Engine::synthesize): host label strings in, host waveform out; the host
                front-end, packing, H2D, kernels and D2H are all inside the timed region.
  roofline      MLSA kernel (dominant): algorithmic FP64 flops / measured kernel time against
                the FP64 peak measured live by an FMA micro-kernel; HBM figure for context.
  cpu_baseline  the oracle (CPU restatement of jbonsai, FMA flavour) on all host cores, on a
                bounded sample of the same workload (rank 0, N=1).

  parity        the e2e output of the utterances the cpu_baseline leg synthesises anyway, compared
                with the oracle's (waveform, 16-bit codes) + MLPG parameters, voiced mask and pulse
                positions of a subset through the debug taps (tests/parity_lib.py).

`--config 4` switches the workload to BASELINE config 4 (16384 tohoku-f01 utterances over all GPUs,
strong scaling); `--single-process` drives all N GPUs from one process through the library's own
sharding (jb_engine_set_devices + one jb_engine_synthesize_batch call) instead of one rank per GPU.
`--impl reference` times the oracle alone (the reference is Rust: no toolchain on this image,
so the "port" is the only CPU arm), same metric/config, rank 0 only.
torch is used for torch.distributed plumbing (barrier / max over ranks) only.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "audio-sec/sec (48 kHz, batched)"
UNIT = "audio-s/s"
RATE = 48000.0
FLOP_PER_SAMPLE = 1281.0      # every sample
FLOP_PER_VOICED_EXTRA = 65.0  # + on voiced samples
MLSA_FLOP_PER_SAMPLE = 1190.0 + 36.0 + 16.0 + 36.0  # fir x5 + df1 + df2 combine + coef interp/volume


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--utts", type=int, default=None, help="utterances per GPU (default: 1024; config 4: 16384 / N)")
    ap.add_argument("--voice", default=None, choices=["nitech", "tohoku-f01-neutral"])
    ap.add_argument("--config", type=int, default=3, choices=[3, 4],
                    help="BASELINE.json config: 3 = 1024 nitech utterances per GPU (weak scaling, the headline), "
                         "4 = 16384 tohoku-f01 utterances sharded over the GPUs (strong scaling)")
    ap.add_argument("--single-process", action="store_true",
                    help="one process drives all --gpus devices through jb_engine_set_devices (not under torchrun)")
    ap.add_argument("--parity-stage-utts", type=int, default=64, help="utterances whose parameters / pulses are compared")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--cpu-sample", type=int, default=256, help="utterances in the CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--mlsa-variant", type=int, default=-1)
    ap.add_argument("--mlsa-warm", type=int, default=-1, help="A/B: warm-up frames of the time-segmented filter")
    ap.add_argument("--mlsa-cut-tolerance", type=float, default=5e-7)
    ap.add_argument("--mlsa-loud-warm", type=int, default=-1, help="A/B: warm-up of cuts behind loud frames (0 = rule off)")
    ap.add_argument("--mlsa-loud-threshold", type=float, default=2.0)
    ap.add_argument("--beta", type=float, default=0.0, help="post-filter coefficient (default off, as the reference)")
    ap.add_argument("--output-mode", type=int, default=-1,
                    help="A/B: 0 one D2H copy, 1 two filter waves, 2 zero-copy stores (library default)")
    ap.add_argument("--device-gather", type=int, default=1, help="0: host interpolates and copies vectors (A/B)")
    args = ap.parse_args()
    if args.voice is None:
        args.voice = "tohoku-f01-neutral" if args.config == 4 else "nitech"
    return args


def dist_env():
    from jbonsai_b200.dist import dist_env as _env
    return _env()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for k, nm in enumerate(names):
                    if r[3 + k].lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        sm.sort()
        # median over the samples taken under load (upper half: idle gaps pull clocks down)
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def utts_per_rank(args, world):
    if args.utts is not None:
        return args.utts
    return 16384 // world if args.config == 4 else 1024


def build_workload(args, rank, world, ranks_here=1):
    """Utterances [rank * n, (rank + ranks_here) * n) of the global synthetic batch (n per GPU)."""
    from jbonsai_b200 import workload as wl
    import jbonsai_b200 as jb
    seed = wl.SEED_NITECH if args.voice == "nitech" else wl.SEED_TOHOKU
    eng = jb.Engine.load([wl.VOICES[args.voice]])  # host front-end only: label durations
    n = utts_per_rank(args, world)
    utts = wl.timed_batch(eng, n * ranks_here, seed, first=rank * n)
    return wl.VOICES[args.voice], utts


def config_dict(args, world):
    n = utts_per_rank(args, world)
    if args.config == 4:
        what = (f"BASELINE config 4: batch of {n * world} synthetic utterances (~5 s each), voice {args.voice} "
                f"(no GV), sharded over {world} GPU(s) in contiguous ranges ({n} per GPU), 48 kHz, fperiod 240")
    else:
        what = (f"BASELINE config 3: batch of {n} synthetic label sequences (~5 s each: 4-6 s "
                f"genji.lab slices between sil labels) per GPU, voice {args.voice}, 48 kHz, fperiod 240")
    return {
        "workload": what, "baseline_config": args.config,
        "utterances_per_gpu": n, "voice": args.voice, "n_gpus": world, "beta": args.beta,
        "parallelism": (f"one process, jb_engine_set_devices x{world}" if args.single_process
                        else f"utterance-sharded x{world}, one rank per GPU") + ", no collective",
        "l2": "inputs and work arrays (> 4 GB per step) are far larger than the 126 MB L2",
    }


def host_cpu():
    """(logical cpus, physical cores, model name) of this box."""
    logical = os.cpu_count() or 1
    cores, model = set(), ""
    try:
        phys = core = None
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name") and not model:
                model = line.split(":", 1)[1].strip()
            elif line.startswith("physical id"):
                phys = line.split(":", 1)[1].strip()
            elif line.startswith("core id"):
                core = line.split(":", 1)[1].strip()
            elif not line.strip():
                if phys is not None and core is not None:
                    cores.add((phys, core))
                phys = core = None
    except OSError:
        pass
    return logical, (len(cores) or logical), model


def oracle_for(voice):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import Oracle
    return Oracle([voice])


def cpu_single_thread(o, utts):
    """Msamples/s of the oracle on ONE host thread (two utterances), the anchor next to the reference's
    published 7.65 Msamples/s on one i5-13500 core (/root/reference README.md:84,124)."""
    t0 = time.perf_counter()
    _, off = o.synthesize_batch(utts[:2], fma=True, nthreads=1)
    return int(off[-1]) / (time.perf_counter() - t0) / 1e6


def cpu_baseline_dict(v, cores, sample_desc, o, utts):
    logical, physical, model = host_cpu()
    return {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample_desc,
            "threads": cores, "logical_cpus": logical, "physical_cores": physical, "cpu_model": model,
            "single_thread_msamples_per_s": cpu_single_thread(o, utts),
            "reference_published_single_thread_msamples_per_s": 7.65,
            "reference_published_on": "jbonsai README.md:84,124 (Intel i5-13500, one core, its own Rust build)"}


def mlsa_traffic():
    """dram read + write bytes of one MLSA launch from the last committed `ncu --set full` capture
    (profiles/mlsa_traffic.json, written by tools/ncu_summary.py with the commit and kernel time it belongs to)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "mlsa_traffic.json")))
    except Exception:
        return None


def run_reference(args, rank, world):
    """--impl reference: the CPU path (oracle port of jbonsai; the Rust crate cannot be built
    here) on all host cores, one utterance per thread."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    voice, utts = build_workload(args, 0, world)
    # the whole per-GPU batch when the run stays within a few minutes, else a bounded sample (stated in config)
    est_full_s = len(utts) * 5.0 / (60.0 * cores)  # >= 60 audio-s/s per host thread observed
    n_s = len(utts) if (args.steps + 1) * est_full_s <= 200.0 else max(1, min(len(utts), args.cpu_sample))
    sample = utts[:n_s]
    o = oracle_for(voice)
    for _ in range(min(args.warmup, 1)):
        o.synthesize_batch(sample[: max(1, len(sample) // 8)], fma=True, nthreads=cores)
    t0 = time.perf_counter()
    samples = 0
    for _ in range(args.steps):
        _, off = o.synthesize_batch(sample, fma=True, nthreads=cores)
        samples += int(off[-1])
    dt = time.perf_counter() - t0
    v = samples / RATE / dt
    desc = (f"{'all' if n_s == len(utts) else 'first'} {len(sample)} of the {len(utts)} utterances of one GPU's batch per step "
            f"({samples // max(args.steps, 1)} samples)")
    cfg = config_dict(args, world)
    cfg["cpu_sample_fraction"] = n_s / max(len(utts), 1)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / max(args.steps, 1) * 1e3,
        "higher_is_better": True, "scaling": "strong" if args.config == 4 else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": cfg,
        "cpu_baseline": cpu_baseline_dict(v, cores, desc, o, utts),
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference jbonsai is Rust (no cargo/rustc on this image): timed arm is the C++ oracle "
                "restatement pinned to the reference's goldens, FMA flavour of fir, one utterance per thread",
    }))


def percentile(xs, q):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, int(round(q * (len(xs) - 1))))] if xs else None


def run_b200(args, rank, world, local):
    import numpy as np
    import jbonsai_b200 as jb

    if jb.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; jbonsai_b200 has no CPU fallback")
    from jbonsai_b200.dist import Reducer
    single = bool(args.single_process)
    if single and world > 1:
        raise SystemExit("bench.py: --single-process is not run under torchrun")
    ngpu = args.gpus if single else world          # GPUs of the whole job
    here = args.gpus if single else 1              # GPUs driven by this process
    if single and jb.device_count() < ngpu:
        raise SystemExit(f"bench.py: --single-process --gpus {ngpu} needs {ngpu} visible devices")
    red = Reducer(world, "nccl" if world > 1 else None, local)
    barrier, max_over_ranks, sum_over_ranks = red.barrier, red.max, red.sum

    if world > 1:  # several ranks share the host: cores / ranks front-end threads each
        jb.set_host_threads(max(1, (os.cpu_count() or 1) // world))
    voice, utts = build_workload(args, rank, ngpu, ranks_here=here)
    eng = jb.Engine.load([voice])
    devs = list(range(ngpu)) if single else [local]
    eng.set_devices(devs)
    if args.mlsa_variant >= 0:
        eng.set_mlsa_variant(args.mlsa_variant)
    if args.output_mode >= 0:
        eng.set_output_mode(args.output_mode)
    if args.beta > 0:
        eng.set("beta", args.beta)
    if args.mlsa_warm >= 0:
        eng.set_mlsa_segmenting(args.mlsa_warm, args.mlsa_cut_tolerance)
    if args.mlsa_loud_warm >= 0:
        eng.set_mlsa_loud_warm(args.mlsa_loud_warm, args.mlsa_loud_threshold)
    eng.set_device_gather(bool(args.device_gather))

    fp64_peak, hbm_live = jb.measure_peaks(devs[0])

    # ---------------- value: kernels only, inputs resident in HBM (one resident batch per device)
    n_per = len(utts) // here
    residents = [eng.prepare(utts[i * n_per:(i + 1) * n_per], i) for i in range(here)]

    def run_all():
        if here == 1:
            return [residents[0].run()]
        out = [None] * here
        th = [threading.Thread(target=lambda i=i: out.__setitem__(i, residents[i].run())) for i in range(here)]
        for x in th:
            x.start()
        for x in th:
            x.join()
        return out

    for _ in range(max(args.warmup, 3)):
        run_all()
    sampler = ClockSampler(devs[0])
    if rank == 0:
        sampler.start()
    barrier()
    dev_ms, mlsa_ms, mlpg_ms, exc_ms = 0.0, 0.0, 0.0, 0.0
    launches = 0
    ts = None
    for _ in range(args.steps):
        ts = run_all()  # CUDA events on the launching stream around all kernels of the step, per device
        dev_ms += max(t["device_ms"] for t in ts)
        mlsa_ms += max(t["mlsa_ms"] for t in ts)
        mlpg_ms += max(t["mlpg_ms"] for t in ts)
        exc_ms += max(t["excitation_ms"] for t in ts)
        launches += int(sum(t["launches"] for t in ts))
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    t = ts[0]
    samples = sum(x["samples"] for x in ts)
    frames, voiced = t["frames"], t["voiced_samples"]
    for r in residents:
        r.close()
    step_ms = max_over_ranks(dev_ms / args.steps)
    total_samples = sum_over_ranks(samples)
    value = total_samples / RATE / (step_ms * 1e-3)

    # ---------------- e2e: label strings -> host waveform through the C ABI
    e2e = None
    last = None  # (out pointer, out_off) of the last timed call, kept for the parity check
    if not args.no_e2e:
        from jbonsai_b200 import _flatten
        flat = _flatten(utts)
        # pinned pools, device arenas and host page tables settle over several calls, and at N = 8 the ranks, which
        # all start their 2 GB copies at the same moment after a barrier, drift into an interleaving that shares the
        # host's 121 GB/s better (250 -> 105 ms per call over ~14 calls): warm up until the last three calls agree to
        # 5 %, at least max(W, 5) calls, at most 40
        n_warm, hist = 0, []
        while True:
            t1 = time.perf_counter()
            out, _, tm = eng.synthesize_batch_raw(flat)
            jb.lib.jb_free(out)
            hist.append(max_over_ranks((time.perf_counter() - t1) * 1e3))
            n_warm += 1
            settled = len(hist) >= 3 and max(hist[-3:]) <= 1.05 * min(hist[-3:])
            if n_warm >= 40 or (n_warm >= max(args.warmup, 5) and settled):
                break
        barrier()
        per_step = []
        t0 = time.perf_counter()
        for i in range(args.steps):
            t1 = time.perf_counter()
            out, out_off, tm = eng.synthesize_batch_raw(flat)
            per_step.append((time.perf_counter() - t1) * 1e3)
            if i + 1 < args.steps:
                jb.lib.jb_free(out)
            else:
                last = (out, out_off)
        mean_ms = (time.perf_counter() - t0) / args.steps * 1e3
        print(f"[rank {rank}] e2e ms per step {[round(x, 2) for x in per_step]}; last call: " +
              ", ".join(f"{k}={tm[k]:.4g}" for k in ("host_frontend_ms", "pack_ms", "h2d_ms", "kernels_ms", "d2h_ms", "mlsa_segments",
                                                     "mlsa_cut_worst", "mlsa_fallback", "mlsa_repaired")),
              file=sys.stderr, flush=True)
        p50 = max_over_ranks(percentile(per_step, 0.5))
        mean_ms = max_over_ranks(mean_ms)
        d2h = tm["d2h_bytes"]
        e2e = {"value": total_samples / RATE / (p50 * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(sum_over_ranks(tm["h2d_bytes"])),
               "d2h_bytes_per_step": int(sum_over_ranks(d2h)),
               "ms_per_step": p50, "ms_per_step_mean": mean_ms, "ms_per_step_p99": max_over_ranks(percentile(per_step, 0.99)),
               "ms_per_step_rank0": [round(x, 3) for x in per_step],
               "timing": f"p50 of {args.steps} calls after {n_warm} warm-up calls (until three in a row agree to 5 %), max over ranks",
               "mlsa_repaired_rank0": int(tm["mlsa_repaired"]), "mlsa_fallback_rank0": int(tm["mlsa_fallback"]),
               "breakdown_ms_rank0": {k: round(tm[k], 3) for k in
                                      ("host_frontend_ms", "pack_ms", "h2d_ms", "kernels_ms", "d2h_ms")}}
        # the same call delivering 16-bit PCM (jb_engine_synthesize_batch_pcm16: the clamp the
        # reference's examples apply to the Vec<f64>, done on the device; 2 B per sample over PCIe)
        for _ in range(3):
            o16, _, tm16 = eng.synthesize_batch_raw(flat, pcm16=True)
            jb.lib.jb_free(o16)
        barrier()
        p16 = []
        for _ in range(max(3, min(args.steps, 10))):
            t1 = time.perf_counter()
            o16, _, tm16 = eng.synthesize_batch_raw(flat, pcm16=True)
            p16.append((time.perf_counter() - t1) * 1e3)
            jb.lib.jb_free(o16)
        p16_ms = max_over_ranks(percentile(p16, 0.5))
        e2e["pcm16"] = {"value": total_samples / RATE / (p16_ms * 1e-3), "unit": UNIT, "ms_per_step": p16_ms,
                        "d2h_bytes_per_step": int(sum_over_ranks(tm16["d2h_bytes"])),
                        "note": "extra: not the reference-facing f64 call; e2e.value is"}

    if rank != 0:
        if last:
            jb.lib.jb_free(last[0])
        return
    # ---------------- roofline of the dominant kernel (MLSA): FP64 pipe
    mlsa_s = mlsa_ms / args.steps * 1e-3
    s0 = t["samples"]  # device 0 of this process: its launch is what mlsa_ms times
    mlsa_flops = MLSA_FLOP_PER_SAMPLE * s0
    achieved = mlsa_flops / mlsa_s / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    mlpg_bytes = frames * (35 * 56 + 56 * (voiced / max(s0, 1)) + 31 * 24)
    traffic = mlsa_traffic()
    roofline = {
        "kernel": "k_mlsa_v3" if t.get("mlsa_segments", 0) > 0 else "k_mlsa_v2", "bound": "fp64",
        "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s",
        "frac": achieved / fp64_peak, "traffic": traffic.get("bytes") if traffic else None,
        "traffic_source": traffic,
        "algorithmic_hbm_bytes": s0 * 16.0 + frames * 35 * 16.0,
        "segments": int(t.get("mlsa_segments", 0)),
        "warmup_overhead": (t.get("mlsa_seg_frames", 0) / frames - 1.0) if frames else None,
        "cut_check_worst": t.get("mlsa_cut_worst"), "sequential_fallback": int(t.get("mlsa_fallback", 0)),
        "repaired_segments": int(t.get("mlsa_repaired", 0)),
        "peak_source": "measured live: DFMA micro-kernel (jb_device_measure_peaks); "
                       "MEASURED_PEAKS.json has no FP64 figure",
        "flops_per_sample": MLSA_FLOP_PER_SAMPLE, "msamples_per_s": s0 / mlsa_s / 1e6,
        "share_of_step": mlsa_ms / dev_ms,
        "hbm": {"kernel": "MLPG phase (gather, expand, band build, LDL, GV, frame prep, pulse walks)",
                "achieved": mlpg_bytes / (mlpg_ms / args.steps * 1e-3) / 1e9, "peak": hbm_peak,
                "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                "live_copy_gbs": hbm_live, "unit": "GB/s", "share_of_step": mlpg_ms / dev_ms},
    }
    roofline["hbm"]["frac"] = roofline["hbm"]["achieved"] / hbm_peak

    # ---------------- cpu_baseline (bounded sample, all host threads) and parity of the same utterances
    cpu, parity = None, None
    if not args.no_cpu_baseline:  # at N > 1 the oracle run only serves the parity block (cpu_baseline is an N = 1 figure)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        cores = os.cpu_count() or 1
        n_s = max(1, min(len(utts), args.cpu_sample))
        stride = max(1, len(utts) // n_s)
        idx = list(range(0, len(utts), stride))[:n_s]  # spread over the batch (and over the devices of one process)
        sample = [utts[i] for i in idx]
        o = oracle_for(voice)
        t0 = time.perf_counter()
        ow, ooff = o.synthesize_batch(sample, fma=True, nthreads=cores)
        dt = time.perf_counter() - t0
        ns = int(ooff[-1])
        desc = (f"{len(sample)} utterances of the workload (every {stride}th; {ns} samples, {dt:.2f} s wall)")
        cpu = cpu_baseline_dict(ns / RATE / dt, cores, desc, o, utts) if world == 1 and not single else None
        if last and not args.no_parity:
            import parity_lib
            out, out_off = last
            off = np.array(list(out_off), dtype=np.int64)
            g = np.ctypeslib.as_array(out, shape=(int(off[-1]),))
            gw = np.concatenate([g[off[i]:off[i + 1]] for i in idx]) if idx else np.zeros(0)
            goff = np.concatenate([[0], np.cumsum([off[i + 1] - off[i] for i in idx])])
            parity = parity_lib.wave_parity(gw, goff, ow, ooff)
            parity["against"] = ("oracle (FMA flavour of fir) on the utterances of the cpu_baseline sample; "
                                 "GPU side = the last timed e2e call")
            nst = max(0, min(args.parity_stage_utts, len(sample)))
            if nst:
                single_dev = jb.Engine.load([voice])  # the debug taps run on one device
                single_dev.set_devices([devs[0]])
                parity.update(parity_lib.stage_parity(single_dev, [voice], sample[:nst], cores))
            parity["ok"] = parity_lib.verdict(parity)
    if last:
        jb.lib.jb_free(last[0])
    print(json.dumps({
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": ngpu, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": step_ms, "higher_is_better": True,
        "scaling": "strong" if args.config == 4 else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(args, ngpu), "e2e": e2e, "gpu_launches": launches,
        "roofline": roofline, "cpu_baseline": cpu, "parity": parity, "clocks": clocks,
        "breakdown_ms_rank0": {"mlpg": mlpg_ms / args.steps, "excitation": exc_ms / args.steps,
                               "mlsa": mlsa_ms / args.steps},
        "samples_per_step": int(total_samples),
    }))
    red.close()


def main():
    args = parse_args()
    rank, world, local = dist_env()
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, world, local)


if __name__ == "__main__":
    main()
