#!/bin/bash
# round 2, final build (run under gpurun): GPU tests, bench (both arms), launch list of the steady-state frames of one sequence,
# full captures of the kernels that carry the step
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2w_gputests.log 2>&1; tail -2 gpurun_out/r2w_gputests.log
python bench.py --impl reference > gpurun_out/r2w_reference.json 2> gpurun_out/r2w_reference.err; tail -c 300 gpurun_out/r2w_reference.json; echo
python bench.py > gpurun_out/r2w_bench.json 2> gpurun_out/r2w_bench.err; tail -c 300 gpurun_out/r2w_bench.json; echo
python profiles/ncu_target.py 20 > gpurun_out/r2w_target_plain.log 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2w_launches.csv python profiles/ncu_target.py 20 > gpurun_out/r2w_ncu_launch.log 2>&1; tail -1 gpurun_out/r2w_ncu_launch.log
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"k34_track_cluster" -c 2 -o gpurun_out/r2w_k34 python profiles/ncu_target.py 4 > gpurun_out/r2w_ncu_k34.log 2>&1; tail -1 gpurun_out/r2w_ncu_k34.log
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"n2_walk|n2_classify|k7_linearize|k9_sc_kernel|k8_top|k5_trace|k1_pyramid|k6_activate|n4_undistort" -c 16 -o gpurun_out/r2w_kf python profiles/ncu_target.py 12 > gpurun_out/r2w_ncu_kf.log 2>&1; tail -1 gpurun_out/r2w_ncu_kf.log
ls -la gpurun_out | grep r2w
