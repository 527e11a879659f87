#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/profile_e2e.py > gpurun_out/profile_e2e.log 2> gpurun_out/profile_e2e.err; echo "exit $?"; cat gpurun_out/profile_e2e.log | cut -c1-200; tail -3 gpurun_out/profile_e2e.err
