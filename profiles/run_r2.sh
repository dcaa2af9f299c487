#!/bin/bash
# round-2 measurement recipe (run under gpurun from the repo root): GPU tests, the bench line, an ncu launch list
tag=${1:-r2}
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_gpu_tests.log
python bench.py --steps 20 --warmup 5 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?" >> gpurun_out/${tag}_bench.err
tail -3 gpurun_out/${tag}_gpu_tests.log; tail -2 gpurun_out/${tag}_bench.err; cut -c1-200 gpurun_out/${tag}_bench.json
