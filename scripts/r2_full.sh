#!/bin/bash
# Full GPU test tier + smoke (what the driver runs at round end)
mkdir -p gpurun_out
O=gpurun_out
FOD_TEST_DUMP=$O timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -12 $O/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke exit $?" >> $O/smoke.log; tail -3 $O/smoke.log | cut -c1-600
