#!/bin/bash
# round 2, call E: bench (ours + reference arm) at N = 1
mkdir -p gpurun_out; O=gpurun_out
( time timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err ) 2>&1 | grep real; echo "bench rc=$?"
tail -5 $O/bench.err
( time timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err ) 2>&1 | grep real
tail -3 $O/bench_ref.err
cut -c1-600 $O/bench_ref.json
