#!/bin/bash
out=gpurun_out
export PYTHONPATH=$PWD:$PYTHONPATH
python -m pytest tests -x -q -m gpu > $out/gpu_tests.log 2>&1; echo "suite rc=$?" >> $out/gpu_tests.log
python scratch/latency.py > $out/latency_r02.txt 2>&1
for wl in heat512 heat256 wave1024; do timeout 300 python bench.py --workload $wl --no-cpu-baseline > $out/bench_${wl}_q.json 2> $out/bench_q.err; done
tail -n 3 $out/gpu_tests.log; grep "us per call" $out/latency_r02.txt
