#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
echo "== parity with wr4"; M3R_LIB_PATH=$PWD/mast3r_slam_ysh_b200/variants/libm3r_wr4.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_vs_ref_cuda.py -m gpu -x -q -k "gauss or gn or edge_blocks" 2>&1 | tail -3
for v in base wr3 wr4 wr6 wr8; do
  for a in "32 8" "256 0"; do
    echo -n "$v $a: "; M3R_LIB_PATH=$PWD/mast3r_slam_ysh_b200/variants/libm3r_$v.so timeout 300 python tools/time_phases.py $a | python -c "import json,sys; d=json.load(sys.stdin); print('hessian=%.1f us iter=%.1f us recs=%d' % (d['hessian_ms']*1e3, d['iter_ms']*1e3, d['records']))"
  done
done 2>&1 | tee $O/variants_h8.txt
