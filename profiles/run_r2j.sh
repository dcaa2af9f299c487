#!/bin/bash
# ncu evidence for round 2 (run under gpurun): launch list of the steady-state frames, full captures of the top kernels
mkdir -p gpurun_out
python profiles/ncu_target.py 20 > gpurun_out/r2j_target_plain.log 2>&1 || exit 1
tail -1 gpurun_out/r2j_target_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2j_launches.csv python profiles/ncu_target.py 20 > gpurun_out/r2j_ncu_launch.log 2>&1; tail -1 gpurun_out/r2j_ncu_launch.log
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"k34_track_cluster" -c 2 -o gpurun_out/r2j_k34 python profiles/ncu_target.py 4 > gpurun_out/r2j_ncu_k34.log 2>&1; tail -1 gpurun_out/r2j_ncu_k34.log
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"n2_select|k7_linearize|k9_sc_kernel|k8_top|k5_trace|k1_pyramid|k6_activate" -c 14 -o gpurun_out/r2j_kf python profiles/ncu_target.py 8 > gpurun_out/r2j_ncu_kf.log 2>&1; tail -1 gpurun_out/r2j_ncu_kf.log
ls -la gpurun_out | grep r2j
