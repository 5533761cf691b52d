#!/bin/bash
# Opcode census of the built C-ABI library (run here, no GPU needed): which Blackwell-native instructions are in the SASS.
# usage: tools/sass_census.sh > profiles/r2_sass_census.txt
SO=mast3r_slam_ysh_b200/libm3r_backends_b200.so
echo "# cuobjdump -sass $SO  (sm_100a only: $(cuobjdump -lelf $SO | grep -c sm_100a) cubins, $(cuobjdump -lelf $SO | grep -vc sm_100a) others)"
echo "# mnemonic guide (B200_PROFILING.md): UTMALDG/UTMASTG = cp.async.bulk.tensor (TMA tensor copies), UBLKCP = cp.async.bulk (TMA bulk),"
echo "# SYNCS = mbarrier, LDGSTS = cp.async, FFMA2/FMUL2/FADD2 = packed f32x2, DFMA = fp64; UTC*MMA/LDTM/STTM (tcgen05) are not used:"
echo "# the path has no GEMM-shaped contraction (DESIGN.md 3.1)"
cuobjdump -sass $SO | grep -E '^\s+/\*[0-9a-f]{4}\*/' | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+(@!?U?P[0-9T]+ )?//' | awk '{print $1}' | sed 's/\..*//; s/;//' | sort | uniq -c | sort -rn | awk '$2 ~ /^(UTMALDG|UTMASTG|UBLKCP|SYNCS|LDGSTS|FFMA2|FMUL2|FADD2|FFMA|DFMA|HMMA|UTCHMMA|LDTM|STTM|HADD2|HMUL2|MUFU|REDUX|SHFL|LDG|STG|LDS|STS|BAR)$/'
echo "# kernels containing UTMALDG / UBLKCP:"
cuobjdump -sass $SO | awk '/Function :/ {f=$3} /UTMALDG/ {a[f]="UTMALDG"} /UBLKCP/ {b[f]="UBLKCP"} END {for (k in a) print a[k], k; for (k in b) print b[k], k}' | c++filt | cut -c1-160 | sort
