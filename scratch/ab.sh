#!/bin/bash
# A/B of a compile-time switch on the GPU box: bench with the shipped .so, rebuild with the flag flipped, bench again
out=gpurun_out
python -m pytest tests -x -q -m gpu > $out/gpu_tests.log 2>&1; echo "suite rc=$?" >> $out/gpu_tests.log
for wl in heat512 heat256; do timeout 300 python bench.py --workload $wl --no-cpu-baseline > $out/bench_${wl}_A.json 2>$out/bench_A.err; done
timeout 300 python bench.py --no-cpu-baseline --norms both > $out/bench_heat512_both_A.json 2>>$out/bench_A.err
cp epde_b200/libepde_b200.so /tmp/lib_A.so
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared --split-compile=0 $1 epde_b200/csrc/epde_b200.cu -o epde_b200/libepde_b200.so > $out/build_B.log 2>&1
for wl in heat512 heat256; do timeout 300 python bench.py --workload $wl --no-cpu-baseline > $out/bench_${wl}_B.json 2>$out/bench_B.err; done
tail -n 3 $out/gpu_tests.log
