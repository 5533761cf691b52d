#!/bin/bash
# round 2, call C: Hessian v7 (self-contained records + TMA bulk ring): tests, variants, ncu
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.log
for v in s4_128x3 s3_128x3 s6_128x3 s4_256x1 s6_96x4; do
  for a in "32 8" "256 0"; do
    echo -n "$v $a: "; M3R_LIB_PATH=$PWD/mast3r_slam_ysh_b200/variants/libm3r_$v.so timeout 300 python tools/time_phases.py $a | python -c "import json,sys; d=json.load(sys.stdin); print('hessian=%.1f us iter=%.1f us begin=%.1f us recs=%d' % (d['hessian_ms']*1e3, d['iter_ms']*1e3, d['begin_ms']*1e3, d['records']))"
  done
done 2>&1 | tee $O/variants_h7b.txt
timeout 300 python tools/prof_gn.py 32 8 2 > $O/prof_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:^hessian -s 1 -c 1 -o $O/hessian_v7b_full -f python tools/prof_gn.py 32 8 2 > $O/ncu_h.log 2>&1
tail -2 $O/ncu_h.log
