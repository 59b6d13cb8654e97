#!/bin/bash
out=gpurun_out
export PYTHONPATH=$PWD:$PYTHONPATH
python -m pytest tests -x -q -m gpu > $out/gpu_tests.log 2>&1; echo "suite rc=$?" >> $out/gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1; echo "smoke rc=$?" >> $out/smoke.log
timeout 300 python bench.py --workload heat512 --no-cpu-baseline > $out/bench_heat512_q.json 2> $out/bench_q.err
timeout 600 python bench.py --population distinct --no-cpu-baseline > $out/bench_heat512_distinct_r02_v7.json 2>> $out/bench_q.err
tail -n 2 $out/gpu_tests.log; tail -n 2 $out/smoke.log
