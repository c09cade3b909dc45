#!/bin/bash
# GPU-box visit for the MLPG-phase kernels: parity tests, a short bench, ncu --set full of the small kernels.
# Usage: bash tools/gpu_small.sh [tag] [kernel-regex]
TAG=${1:-r2s}
RX=${2:-'k_mlpg_gv|k_ldl|k_pulse_runs|k_coef_rows|k_band_build'}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 5 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$RX" -c 12 -o $O/${TAG}_small_full -f \
  python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --no-parity > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 $O/${TAG}_pytest_gpu.log; cut -c1-400 $O/${TAG}_bench.json; python -c "
import json;d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1]);print(d['breakdown_ms_rank0'], d['parity']['ok'], d['roofline']['frac'])"
