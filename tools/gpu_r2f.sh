#!/bin/bash
# round 2, multi-GPU call: tests on 2+ devices, bench at --gpus N (N = device count)
mkdir -p gpurun_out; O=gpurun_out
N=$(python -c "import torch; print(torch.cuda.device_count())")
timeout 1200 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $O/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -15 $O/pytest_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench N=$N rc=$?"
tail -5 $O/bench_n$N.err | cut -c1-300
python - <<PY
import json
d=[json.loads(l) for l in open("$O/bench_n$N.json") if l.startswith("{")][-1]
print({k:d[k] for k in ("value","n_gpus","ms_per_step")}, d["e2e"]["value"], d["extra"].get("hessian_kernel_ms"), d["extra"].get("linear_solver"))
print(json.dumps(d["extra"]["configs4_strong_scaling"]))
PY
