#!/bin/bash
# Multi-GPU visit (charged N x): the scaling line of config 3 under torchrun, config 4 under torchrun and in one process.
# Usage: bash tools/gpu_roundN.sh N [tag]
N=${1:-2}
TAG=${2:-r2}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --gpus $N --steps 8 --warmup 3 > $O/${TAG}_bench_n$N.json 2> $O/${TAG}_bench_n$N.err; echo "c3 torchrun rc=$?"
if [ "$N" = "8" ]; then
timeout 600 $TR bench.py --gpus $N --config 4 --steps 5 --warmup 3 > $O/${TAG}_bench_c4_n$N.json 2> $O/${TAG}_bench_c4_n$N.err; echo "c4 torchrun rc=$?"
timeout 600 python bench.py --single-process --gpus $N --config 4 --steps 5 --warmup 3 > $O/${TAG}_bench_c4_sp_n$N.json 2> $O/${TAG}_bench_c4_sp_n$N.err; echo "c4 single-process rc=$?"
else
timeout 600 python bench.py --single-process --gpus $N --steps 8 --warmup 3 > $O/${TAG}_bench_sp_n$N.json 2> $O/${TAG}_bench_sp_n$N.err; echo "c3 single-process rc=$?"
fi
for f in $O/${TAG}_bench*_n$N.json; do python - <<PY
import json
try:
    d=json.loads(open("$f").read().strip().splitlines()[-1]); e=d.get("e2e") or {}; r=d.get("roofline") or {}; p=d.get("parity") or {}
    print("$f", d["n_gpus"], round(d["value"]), round(d["ms_per_step"],2), "e2e", round(e.get("value",0)), round(e.get("ms_per_step",0),1), "pcm16", round((e.get("pcm16") or {}).get("value",0)), "frac", round(r.get("frac",0),4), "parity", p.get("ok"))
except Exception as ex:
    print("$f", "unreadable", ex)
PY
done
