set -x
timeout 900 python bench.py > gpurun_out/bench_r1l.json 2> gpurun_out/bench_r1l.err; tail -c 1500 gpurun_out/bench_r1l.json
timeout 600 python profiles/kernel_bench.py --json gpurun_out/kb_r1l.json > gpurun_out/kb_r1l.log 2>&1; tail -15 gpurun_out/kb_r1l.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k8_top" -c 1 -o gpurun_out/prof_r1l_k8 python profiles/kernel_bench.py --batches 64 --reps 1 > gpurun_out/ncu_k8.log 2>&1; tail -3 gpurun_out/ncu_k8.log
ls -la gpurun_out
