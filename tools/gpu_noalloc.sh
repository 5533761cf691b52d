#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
echo "== alone, fresh process"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -4
echo "== after the other tests"; timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_gpu.log
