#!/bin/bash
# round 2: full captures of the keyframe kernels of the final build (the r2w capture filled up with per-frame kernels)
mkdir -p gpurun_out
python -m pytest tests/test_selector.py tests/test_system.py -m gpu -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke(); print(\"smoke ok\")" 2>&1 | tail -2
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"n2_walk|n2_classify|k7_linearize|k9_sc_kernel|k8_top|k6_activate|k2_emit|k2_count|a2_final" -c 18 -o gpurun_out/r2x_kf python profiles/ncu_target.py 12 > gpurun_out/r2x_ncu_kf.log 2>&1; tail -1 gpurun_out/r2x_ncu_kf.log
ls -la gpurun_out | grep r2x
