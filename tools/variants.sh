#!/bin/bash
# Build solver variants on the GPU box and time them (dev tool).  usage: tools/variants.sh "<flags1>" "<flags2>" ...
cd mast3r_slam_ysh_b200/csrc
for f in "$@"; do
  rm -f _obj/gn_fused.o; make -s EXTRA="$f" > /dev/null 2>&1 || { echo "build failed: $f"; continue; }
  echo "== variant: [$f]"
  (cd ../..; for i in 1 2; do python tools/time_phases.py 32 8 | python -c "import json,sys; d=json.load(sys.stdin); print('   n=%d iter=%.1f us solve=%.1f us accumulate=%.1f us' % (d['n'], d['iter_ms']*1e3, d['solve_ms']*1e3, d['accumulate_ms']*1e3))"; done)
done
rm -f _obj/gn_fused.o; make -s > /dev/null 2>&1
