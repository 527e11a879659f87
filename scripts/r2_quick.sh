#!/bin/bash
# usage: r2_quick.sh <selftest mask> [pytest -k expression]
mkdir -p gpurun_out
O=gpurun_out
cd future-object-detection_b200
timeout 180 ./selftest_fod $1 > ../$O/selftest.log 2>&1; echo "selftest exit $?" >> ../$O/selftest.log
cat ../$O/selftest.log | cut -c1-260
cd ..
if [ -n "$2" ]; then
  FOD_TEST_DUMP=$O timeout 900 python -m pytest tests -m gpu -x -q -k "$2" > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
  tail -25 $O/pytest_gpu.log
fi
