#!/bin/bash
# Quick GPU visit: self-test, kernel + model parity tests, short bench. Outputs -> gpurun_out/
mkdir -p gpurun_out
(cd future-object-detection_b200 && timeout 120 ./selftest_fod > ../gpurun_out/selftest.log 2>&1; echo "selftest exit $?" >> ../gpurun_out/selftest.log)
tail -25 gpurun_out/selftest.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "e2e", "gpu_launches")})
    print(json.dumps(d["kernels"], indent=0)[:3000])
except Exception as e:
    print("bench parse failed", e); print(open("gpurun_out/bench.err").read()[-3000:])
PY
