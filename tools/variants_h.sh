#!/bin/bash
# Build Hessian-kernel variants on the GPU box and time them (dev tool; full rebuild per variant because the thread
# count enters shared headers).  usage: tools/variants_h.sh "<flags1>" "<flags2>" ...
cd mast3r_slam_ysh_b200/csrc
for f in "$@"; do
  rm -f _obj/*.o; make -s -j8 EXTRA="$f" > /dev/null 2>&1 || { echo "build failed: $f"; continue; }
  echo "== variant: [$f]"
  (cd ../..; for a in "32 8" "32 8" "256 0"; do python tools/time_phases.py $a | python -c "import json,sys; d=json.load(sys.stdin); print('   nkf=%d hessian=%.1f us iter=%.1f us' % (d['nkf'], d['hessian_ms']*1e3, d['iter_ms']*1e3))"; done)
done
rm -f _obj/*.o; make -s -j8 > /dev/null 2>&1
