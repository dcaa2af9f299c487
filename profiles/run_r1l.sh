set -x
# 1. the bench line itself (never under a profiler)
timeout 900 python bench.py > gpurun_out/bench_r1l.json 2> gpurun_out/bench_r1l.err; tail -c 600 gpurun_out/bench_r1l.json
# 2. launch list of the same command (short run)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_r1l.csv python bench.py --steps 30 --warmup 30 --cpu-steps 5 --no-window-roofline > gpurun_out/ncu_launch.log 2>&1; tail -2 gpurun_out/ncu_launch.log | cut -c1-300
# 3. batched-window kernels: event-timed table, then one full capture each
timeout 600 python profiles/kernel_bench.py --json gpurun_out/kb_r1l.json > gpurun_out/kb_r1l.log 2>&1; tail -15 gpurun_out/kb_r1l.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k7_linearize|k8_top|k9_sc_kernel|k9_point|k10_resub" -c 5 -o gpurun_out/prof_r1l_window64 python profiles/kernel_bench.py --batches 64 --reps 1 > gpurun_out/ncu_kb.log 2>&1; tail -3 gpurun_out/ncu_kb.log
# 4. the frame kernels (tracker, pyramid, traceOn) out of the closed loop
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k34_track_level|k1_pyramid|k5_trace" -s 150 -c 6 -o gpurun_out/prof_r1l_frame python bench.py --steps 6 --warmup 30 --cpu-steps 5 --no-window-roofline > gpurun_out/ncu_fr.log 2>&1; tail -3 gpurun_out/ncu_fr.log | cut -c1-300
ls -la gpurun_out
