#!/bin/bash
# One gpurun call: GPU tests, bench, phase timings, launch list, ncu full captures (each ncu pass only after the
# plain run exited 0).  Outputs under gpurun_out/.   usage: tools/gpu_round.sh [quick]
mkdir -p gpurun_out; O=gpurun_out; : > $O/rc.txt
KREG='regex:hessian_kernel|pack_|finalize_edges|gn_fused|gn_small|edge_|uniq_|incidence|iter_proj|refine_matches|repack|finish_kernel|init_state|concat_ids|zero_f32|track_kernel|match_|ray_|desc_|index_prep'
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt
timeout 300 python tools/time_phases.py 32 8 > $O/phases_32.json 2>&1
timeout 300 python tools/time_phases.py 256 0 > $O/phases_256.json 2>&1
timeout 300 python tools/time_phases.py 10 3 > $O/phases_10.json 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt
timeout 200 python tools/time_track.py > $O/track.json 2>&1
for a in "37 3" "256 8" "256 8 0" "512 8" "256 0"; do M3R_TL=1 timeout 120 python tools/time_solver.py $a 2>&1 | tail -1; done > $O/solver_modes.jsonl
if [ "$1" != "quick" ]; then
timeout 300 python tools/prof_gn.py 32 8 2 > $O/prof_plain.log 2>&1; rc=$?; echo "prof rc=$rc" | tee -a $O/rc.txt
if [ $rc -eq 0 ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREG" -c 400 --csv --log-file $O/launches.csv python tools/prof_gn.py 32 8 2 > $O/ncu_list.log 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:^hessian -s 1 -c 1 -o $O/hessian_full -f python tools/prof_gn.py 32 8 2 > $O/ncu_h.log 2>&1
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREG" -c 600 --csv --log-file $O/launches_bench.csv python bench.py --steps 2 --warmup 1 > $O/ncu_list_bench.log 2>&1
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:track_kernel -s 2 -c 1 -o $O/track_full -f python tools/time_track.py > $O/ncu_t.log 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:pack_edges -s 1 -c 1 -o $O/pack_full -f python tools/prof_gn.py 32 8 2 > $O/ncu_p.log 2>&1
  # envelope solver: launch 69 of tools/time_solver.py is the first FACTOR|BACKSOLVE launch in envelope mode (68 dense-mode launches before it)
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:gn_fused -s 69 -c 1 -o $O/sky_full -f python tools/time_solver.py 256 8 > $O/ncu_sky.log 2>&1
fi
fi
cat $O/rc.txt; tail -3 $O/pytest_gpu.log; cat $O/phases_*.json; cat $O/track.json; cut -c1-400 $O/solver_modes.jsonl; cat $O/bench.json
