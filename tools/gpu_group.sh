#!/bin/bash
# Multi-GPU call (gpurun --gpus N): the group API tests, the torchrun tests, then the BASELINE graph configs through
# shb_group_* (one process, N devices) and, for comparison, through the torch.distributed drivers.
N=${N:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_x$N.txt 2>&1
if [ -z "$QUICK" ] && [ -x tools/microbench/gather_bench ]; then echo "== gather microbench"; timeout 60 tools/microbench/gather_bench | tee gpurun_out/gather_bench.jsonl; fi
[ -z "$QUICK" ] && echo "== pytest group + multi" && timeout 300 python -m pytest tests/test_gpu_group.py tests/test_gpu_multi.py -m gpu -x -q --timeout 100 2>&1 | tail -8
[ -z "$QUICK" ] && echo "== group bench small" && timeout 120 python tools/group_bench.py --gpus $N --small 2>&1 | tail -3 | cut -c1-700
echo "== group bench full"; timeout 300 python tools/group_bench.py --gpus $N > gpurun_out/group_bench_x$N.jsonl 2> gpurun_out/group_bench_x$N.err; cut -c1-900 gpurun_out/group_bench_x$N.jsonl; tail -3 gpurun_out/group_bench_x$N.err
echo "== host pr_harness --gpus $N"; timeout 120 host/bin/pr_harness -e x -n box -f rmat -k tests/golden/kernels/spmv/awrg-alcl-fdp-chunk-rsa-8.json -r tests/golden/kernels/runfile.csv -i 3 -t 2000 --mode canonical --synthetic rmat-pr:22:16 --max_iterations 20 -c 0 --gpus $N 2>/dev/null | grep -E "PageRank:|INSERT" | cut -c1-300
[ -z "$QUICK" ] && echo "== pcie" && timeout 120 python tools/pcie_bench.py | tee gpurun_out/pcie_bench.json | cut -c1-600
echo "== torchrun bench"; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_x$N.json 2> gpurun_out/bench_x$N.err; python - <<PY
import json
try:
    l=json.load(open("gpurun_out/bench_x$N.json"))
    print({k:l[k] for k in ("value","ms_per_step","n_gpus")}, l.get("scaling_curve"))
    for k,v in l.get("graph_gteps",{}).items():
        print(k, {kk:v.get(kk) for kk in ("driver","ms","gteps","parity_ok","error","partitioned_ms","single_device_ms")} if isinstance(v,dict) else v)
except Exception as e:
    print("bench parse failed",e); print(open("gpurun_out/bench_x$N.err").read()[-1500:])
PY
