#!/bin/bash
# round 2, call A: full GPU test suite (incl. full-size parity vs the reference CUDA), phase timings, bench
mkdir -p gpurun_out; O=gpurun_out; : > $O/rc.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
timeout 300 python tools/time_phases.py 32 8 > $O/phases_32.json 2>&1
timeout 300 python tools/time_phases.py 10 3 > $O/phases_10.json 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
cat $O/rc.txt; tail -15 $O/pytest_gpu.log; cat $O/phases_*.json; cut -c1-1500 $O/bench.json; cat $O/parity_full_size.json
