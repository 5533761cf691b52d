#!/bin/bash
# round 2: GPU tests, bench (both arms), ncu full capture of the Hessian kernel, ncu launch list of the bench command
mkdir -p gpurun_out; O=gpurun_out; : > $O/rc.txt
KREG='regex:hessian_kernel|pack_|conf_scan|finalize_edges|gn_fused|gn_small|edge_|incidence|iter_proj|refine_matches|repack|finish_kernel|init_state|zero_f32|track_kernel|match_|ray_|desc_|index_prep|skyline|nd_plan|update_pointmap|constrain'
# usage: gpu_round2.sh [a|b]   a = tests, both bench arms, ncu capture of the Hessian kernel; b = launch list of the bench
if [ "${1:-a}" = "a" ]; then
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -3 $O/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?" | tee -a $O/rc.txt
timeout 300 python tools/prof_gn.py 32 8 2 > $O/prof_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:^hessian -s 1 -c 1 -o $O/hessian_v8_full -f python tools/prof_gn.py 32 8 2 > $O/ncu_h.log 2>&1
else
timeout 900 python bench.py --steps 2 --warmup 1 > $O/bench_short.json 2> $O/bench_short.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREG" -c 700 --csv --log-file $O/launches_bench.csv python bench.py --steps 2 --warmup 1 > $O/ncu_list_bench.log 2>&1
echo "ncu list rc=$?" | tee -a $O/rc.txt
fi
cat $O/rc.txt; tail -2 $O/ncu_h.log 2>/dev/null
