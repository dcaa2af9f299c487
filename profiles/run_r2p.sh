#!/bin/bash
# round 2, call p: (1) distance-map tests with the exact-division stamp, (2) is the S >= 16 plateau PCIe? (images resident in HBM),
# (3) ncu: source-level capture of the tracker cluster kernel and of the activation walk
mkdir -p gpurun_out
python -m pytest tests/test_distmap.py -m gpu -x -q 2>&1 | tail -2
SEQ_DEVICE_IMAGES=1 python profiles/seq_scaling.py 8,16,24 > gpurun_out/r2p_seq_dev.log 2>&1; tail -3 gpurun_out/r2p_seq_dev.log
python profiles/ncu_target.py 20 > gpurun_out/r2p_target_plain.log 2>&1 || exit 1
tail -2 gpurun_out/r2p_target_plain.log
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"k34_track_cluster" -c 1 -o gpurun_out/r2p_k34 python profiles/ncu_target.py 4 > gpurun_out/r2p_ncu_k34.log 2>&1; tail -1 gpurun_out/r2p_ncu_k34.log
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"n2_walk|n2_classify" -c 2 -o gpurun_out/r2p_n2 python profiles/ncu_target.py 12 > gpurun_out/r2p_ncu_n2.log 2>&1; tail -1 gpurun_out/r2p_ncu_n2.log
ls -la gpurun_out | grep r2p
