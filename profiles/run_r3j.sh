#!/bin/bash
# head of round 2 (run under gpurun): launch list of the steady-state frames of one sequence (same target as run_r2w.sh)
mkdir -p gpurun_out
python profiles/ncu_target.py 20 > gpurun_out/r3j_target_plain.log 2>&1 || { tail -5 gpurun_out/r3j_target_plain.log; exit 1; }
tail -2 gpurun_out/r3j_target_plain.log | cut -c1-300
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r3j_launches.csv python profiles/ncu_target.py 20 > gpurun_out/r3j_ncu_launch.log 2>&1; tail -1 gpurun_out/r3j_ncu_launch.log | cut -c1-300
ls -la gpurun_out | grep r3j
