#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
for v in base s3 s6; do
  for a in "32 8" "256 0"; do
    echo -n "$v $a: "; M3R_LIB_PATH=$PWD/mast3r_slam_ysh_b200/variants/libm3r_$v.so timeout 300 python tools/time_phases.py $a | python -c "import json,sys; d=json.load(sys.stdin); print('hessian=%.1f us iter=%.1f us begin=%.1f us recs=%d' % (d['hessian_ms']*1e3, d['iter_ms']*1e3, d['begin_ms']*1e3, d['records']))"
  done
done 2>&1 | tee $O/variants_h7c.txt
