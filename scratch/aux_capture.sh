#!/bin/bash
# on the GPU box: GPU test suite, single-call latency, producer timings, ncu DRAM numbers of the producer kernels
out=gpurun_out
export PYTHONPATH=$PWD:$PYTHONPATH
python -m pytest tests -x -q -m gpu > $out/gpu_tests.log 2>&1; echo "suite rc=$?" >> $out/gpu_tests.log
python scratch/latency.py > $out/latency_r02.txt 2>&1
python scratch/producers_run.py > $out/producers_r02.txt 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed \
   --clock-control none -k regex:"fd_gradient|poly_derivative|crop_interior" -c 40 --csv --log-file $out/producers_ncu_r02.csv \
   python scratch/producers_run.py > $out/producers_ncu_r02.log 2>&1
timeout 300 python bench.py --no-cpu-baseline > $out/bench_heat512_q.json 2> $out/bench_q.err
timeout 300 python bench.py --workload wave1024 --no-cpu-baseline > $out/bench_wave1024_q.json 2>> $out/bench_q.err
tail -n 3 $out/gpu_tests.log; cat $out/latency_r02.txt $out/producers_r02.txt
