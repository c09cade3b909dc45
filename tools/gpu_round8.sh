#!/bin/bash
# One 8-GPU visit (charged 8x): the library's own sharding, BASELINE config 4 as one batch, the scaling line, PCIe probe.
TAG=${1:-r2}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
timeout 300 python -m pytest tests -m gpu -q -k "multi_gpu" 2>&1 | tail -3 > $O/${TAG}_multi_gpu_test_n8.log; cat $O/${TAG}_multi_gpu_test_n8.log
timeout 600 $TR bench.py --gpus 8 --config 4 --steps 5 --warmup 3 > $O/${TAG}_bench_c4_n8.json 2> $O/${TAG}_bench_c4_n8.err; echo "c4 torchrun rc=$?"
timeout 600 python bench.py --single-process --gpus 8 --config 4 --steps 5 --warmup 3 > $O/${TAG}_bench_c4_sp_n8.json 2> $O/${TAG}_bench_c4_sp_n8.err; echo "c4 single-process rc=$?"
timeout 600 $TR bench.py --gpus 8 --steps 5 --warmup 3 > $O/${TAG}_bench_n8.json 2> $O/${TAG}_bench_n8.err; echo "c3 torchrun rc=$?"
timeout 200 python tools/pcie_probe.py > $O/${TAG}_pcie_probe_n8.txt 2>&1
grep "e2e ms" $O/${TAG}_bench_c4_n8.err $O/${TAG}_bench_n8.err | cut -c1-260
