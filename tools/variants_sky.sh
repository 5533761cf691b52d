#!/bin/bash
# Build fused-solver variants on the GPU box and time the envelope solver (dev tool).  usage: tools/variants_sky.sh "<flags1>" ...
cd mast3r_slam_ysh_b200/csrc
for f in "$@"; do
  rm -f _obj/gn_fused.o; make -s EXTRA="$f" > /dev/null 2>&1 || { echo "build failed: $f"; continue; }
  echo "== variant: [$f]"
  (cd ../..; M3R_TL=1 python tools/time_solver.py 256 8 | python -c "import json,sys; d=json.load(sys.stdin); print('   n=%d dense=%.1f us envelope=%.1f us diff=%.1e' % (d['n'], d['dense_solve_ms']*1e3, d['envelope_solve_ms']*1e3, d['pose_rel_diff'])); print('   ', d.get('timeline_us'))"
   python tools/time_phases.py 32 8 | python -c "import json,sys; d=json.load(sys.stdin); print('   n=%d iter=%.1f us solve=%.1f us accumulate=%.1f us' % (d['n'], d['iter_ms']*1e3, d['solve_ms']*1e3, d['accumulate_ms']*1e3))")
done
rm -f _obj/gn_fused.o; make -s > /dev/null 2>&1
