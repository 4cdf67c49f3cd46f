#!/bin/bash
# quick GPU trip: selected tests (pytest -k "$1"), optionally then the full GPU suite ($2 = all)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x -k "$1" > gpurun_out/test_quick.log 2>&1
rc=$?; echo "== quick tests exit $rc"; tail -n 40 gpurun_out/test_quick.log
if [ "$2" == "all" ]; then
  timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/test_gpu.log 2>&1; rc=$?
  echo "== pytest exit $rc"; tail -n 15 gpurun_out/test_gpu.log
fi
exit $rc
