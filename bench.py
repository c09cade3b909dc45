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
    ap.add_argument("--utts", type=int, default=1024, help="utterances per GPU")
    ap.add_argument("--voice", default="nitech", choices=["nitech", "tohoku-f01-neutral"])
    ap.add_argument("--cpu-sample", type=int, default=256, help="utterances in the CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--mlsa-variant", type=int, default=-1)
    ap.add_argument("--mlsa-warm", type=int, default=-1, help="A/B: warm-up frames of the time-segmented filter")
    ap.add_argument("--mlsa-cut-tolerance", type=float, default=1e-7)
    ap.add_argument("--beta", type=float, default=0.0, help="post-filter coefficient (default off, as the reference)")
    ap.add_argument("--output-mode", type=int, default=-1,
                    help="A/B: 0 one D2H copy, 1 two filter waves, 2 zero-copy stores (library default)")
    ap.add_argument("--device-gather", type=int, default=1, help="0: host interpolates and copies vectors (A/B)")
    return ap.parse_args()


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


def build_workload(args, rank):
    from jbonsai_b200 import workload as wl
    import jbonsai_b200 as jb
    seed = wl.SEED_NITECH if args.voice == "nitech" else wl.SEED_TOHOKU
    eng = jb.Engine.load([wl.VOICES[args.voice]])  # host front-end only: label durations
    from jbonsai_b200.dist import rank_utterance_range
    utts = wl.timed_batch(eng, args.utts, seed, first=rank_utterance_range(args.utts, rank)[0])
    return wl.VOICES[args.voice], utts


def config_dict(args, world):
    return {
        "workload": f"BASELINE config 3: batch of {args.utts} synthetic label sequences (~5 s each: 4-6 s "
                    f"genji.lab slices between sil labels) per GPU, voice {args.voice}, 48 kHz, fperiod 240",
        "utterances_per_gpu": args.utts, "voice": args.voice, "n_gpus": world, "beta": args.beta,
        "parallelism": f"utterance-sharded x{world}, no collective",
        "l2": "inputs and work arrays (> 4 GB per step) are far larger than the 126 MB L2",
    }


def cpu_oracle_rate(voice, utts, nthreads, repeats=1):
    """audio-s/s of the oracle (FMA flavour of fir = jbonsai 'native') on nthreads host cores."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import Oracle
    o = Oracle([voice])
    best, samples = None, 0
    for _ in range(repeats):
        t0 = time.perf_counter()
        w, off = o.synthesize_batch(utts, fma=True, nthreads=nthreads)
        dt = time.perf_counter() - t0
        samples = int(off[-1])
        best = dt if best is None else min(best, dt)
    return samples / RATE / best, samples, best


def ncu_traffic_bytes():
    """dram__bytes_read.sum + dram__bytes_write.sum of one MLSA launch, from the committed ncu capture."""
    try:
        tot = 0.0
        for line in open(os.path.join(ROOT, "profiles", "r1_mlsa_v3_full_summary.txt")):
            if line.startswith("dram__bytes_read.sum [") or line.startswith("dram__bytes_write.sum ["):
                unit = line.split("[")[1].split("]")[0]
                val = float(line.split("=")[1])
                tot += val * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[unit]
        return tot or None
    except Exception:
        return None


def run_reference(args, rank, world):
    """--impl reference: the CPU path (oracle port of jbonsai; the Rust crate cannot be built
    here) on all host cores, one utterance per thread, bounded sample per step."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    voice, utts = build_workload(args, 0)
    sample = utts[: max(1, min(len(utts), args.cpu_sample))]
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import Oracle
    o = Oracle([voice])
    for _ in range(min(args.warmup, 1)):
        o.synthesize_batch(sample[: max(1, len(sample) // 8)], fma=True, nthreads=cores)
    t0 = time.perf_counter()
    samples = 0
    for _ in range(args.steps):
        _, off = o.synthesize_batch(sample, fma=True, nthreads=cores)
        samples += int(off[-1])
    dt = time.perf_counter() - t0
    v = samples / RATE / dt
    desc = f"first {len(sample)} utterances of the workload per step ({samples // max(args.steps,1)} samples)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / max(args.steps, 1) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": config_dict(args, world),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference jbonsai is Rust (no cargo/rustc on this image): timed arm is the C++ oracle "
                "restatement pinned to the reference's goldens, FMA flavour of fir, one utterance per thread",
    }))


def run_b200(args, rank, world, local):
    import jbonsai_b200 as jb

    if jb.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; jbonsai_b200 has no CPU fallback")
    from jbonsai_b200.dist import Reducer
    red = Reducer(world, "nccl" if world > 1 else None, local)
    barrier, max_over_ranks, sum_over_ranks = red.barrier, red.max, red.sum

    voice, utts = build_workload(args, rank)
    eng = jb.Engine.load([voice])
    eng.set_devices([local])
    if args.mlsa_variant >= 0:
        eng.set_mlsa_variant(args.mlsa_variant)
    if args.output_mode >= 0:
        eng.set_output_mode(args.output_mode)
    if args.beta > 0:
        eng.set("beta", args.beta)
    if args.mlsa_warm >= 0:
        eng.set_mlsa_segmenting(args.mlsa_warm, args.mlsa_cut_tolerance)
    eng.set_device_gather(bool(args.device_gather))

    fp64_peak, hbm_live = jb.measure_peaks(local)

    # ---------------- value: kernels only, inputs resident in HBM
    res = eng.prepare(utts, 0)
    for _ in range(max(args.warmup, 3)):
        res.run()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    dev_ms, mlsa_ms, mlpg_ms, exc_ms = 0.0, 0.0, 0.0, 0.0
    launches = 0
    t = None
    for _ in range(args.steps):
        t = res.run()  # CUDA events on the launching stream around all kernels of the step
        dev_ms += t["device_ms"]
        mlsa_ms += t["mlsa_ms"]
        mlpg_ms += t["mlpg_ms"]
        exc_ms += t["excitation_ms"]
        launches += int(t["launches"])
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    samples, frames, voiced = t["samples"], t["frames"], t["voiced_samples"]
    res.close()
    step_ms = max_over_ranks(dev_ms / args.steps)
    total_samples = sum_over_ranks(samples)
    value = total_samples / RATE / (step_ms * 1e-3)

    # ---------------- e2e: label strings -> host waveform through the C ABI
    e2e = None
    if not args.no_e2e:
        from jbonsai_b200 import _flatten
        flat = _flatten(utts)
        for _ in range(2):
            out, _, tm = eng.synthesize_batch_raw(flat)
            jb.lib.jb_free(out)
        barrier()
        n_e2e = max(2, min(args.steps, 5))
        t0 = time.perf_counter()
        per_step = []
        for _ in range(n_e2e):
            t1 = time.perf_counter()
            out, _, tm = eng.synthesize_batch_raw(flat)
            jb.lib.jb_free(out)
            per_step.append(round((time.perf_counter() - t1) * 1e3, 2))
        e2e_ms = (time.perf_counter() - t0) / n_e2e * 1e3
        print(f"[rank {rank}] e2e ms per step {per_step}; last call: " +
              ", ".join(f"{k}={tm[k]:.4g}" for k in ("host_frontend_ms", "pack_ms", "h2d_ms", "kernels_ms", "d2h_ms", "mlsa_segments",
                                                     "mlsa_cut_worst", "mlsa_fallback", "mlsa_repaired")),
              file=sys.stderr, flush=True)
        e2e_ms = max_over_ranks(e2e_ms)
        e2e = {"value": total_samples / RATE / (e2e_ms * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(sum_over_ranks(tm["h2d_bytes"])),
               "d2h_bytes_per_step": int(sum_over_ranks(tm["d2h_bytes"])),
               "ms_per_step": e2e_ms,
               "breakdown_ms_rank0": {k: round(tm[k], 3) for k in
                                      ("host_frontend_ms", "pack_ms", "h2d_ms", "kernels_ms", "d2h_ms")}}
        # the same call delivering 16-bit PCM (jb_engine_synthesize_batch_pcm16: the clamp the
        # reference's examples apply to the Vec<f64>, done on the device; 2 B per sample over PCIe)
        for _ in range(2):
            out, _, tm16 = eng.synthesize_batch_raw(flat, pcm16=True)
            jb.lib.jb_free(out)
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            out, _, tm16 = eng.synthesize_batch_raw(flat, pcm16=True)
            jb.lib.jb_free(out)
        p16_ms = max_over_ranks((time.perf_counter() - t0) / n_e2e * 1e3)
        e2e["pcm16"] = {"value": total_samples / RATE / (p16_ms * 1e-3), "unit": UNIT, "ms_per_step": p16_ms,
                        "d2h_bytes_per_step": int(sum_over_ranks(tm16["d2h_bytes"])),
                        "note": "extra: not the reference-facing f64 call; e2e.value is"}

    if rank != 0:
        return
    # ---------------- roofline of the dominant kernel (MLSA): FP64 pipe
    mlsa_s = mlsa_ms / args.steps * 1e-3
    mlsa_flops = MLSA_FLOP_PER_SAMPLE * samples
    achieved = mlsa_flops / mlsa_s / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    mlpg_bytes = frames * (35 * 56 + 56 * (voiced / max(samples, 1)) + 31 * 24)
    roofline = {
        "kernel": "k_mlsa_v3" if t.get("mlsa_segments", 0) > 0 else "k_mlsa_v2", "bound": "fp64",
        "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s",
        "frac": achieved / fp64_peak, "traffic": ncu_traffic_bytes(),
        "traffic_source": "profiles/r1_mlsa_v3_full_summary.txt (dram read + write of one launch, ncu --set full)",
        "algorithmic_hbm_bytes": samples * 16.0 + frames * 35 * 8.0,
        "segments": int(t.get("mlsa_segments", 0)),
        "warmup_overhead": (t.get("mlsa_seg_frames", 0) / frames - 1.0) if frames else None,
        "cut_check_worst": t.get("mlsa_cut_worst"), "sequential_fallback": int(t.get("mlsa_fallback", 0)),
        "repaired_segments": int(t.get("mlsa_repaired", 0)),
        "peak_source": "measured live: DFMA micro-kernel (jb_device_measure_peaks); "
                       "MEASURED_PEAKS.json has no FP64 figure",
        "flops_per_sample": MLSA_FLOP_PER_SAMPLE, "msamples_per_s": samples / mlsa_s / 1e6,
        "share_of_step": mlsa_ms / dev_ms,
        "hbm": {"kernel": "k_mlpg (all streams + expand + frame_prep)",
                "achieved": mlpg_bytes / (mlpg_ms / args.steps * 1e-3) / 1e9, "peak": hbm_peak,
                "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                "live_copy_gbs": hbm_live, "unit": "GB/s", "share_of_step": mlpg_ms / dev_ms},
    }
    roofline["hbm"]["frac"] = roofline["hbm"]["achieved"] / hbm_peak
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        sample = utts[: max(1, min(len(utts), args.cpu_sample))]
        v, ns, dt = cpu_oracle_rate(voice, sample, cores)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"first {len(sample)} utterances of the workload ({ns} samples, {dt:.2f} s wall)"}
    print(json.dumps({
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": step_ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(args, world), "e2e": e2e, "gpu_launches": launches,
        "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
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
