#!/bin/bash
# A/B/C of GRAM_PROD_WARPS on the GPU box: shipped .so (10), then 12, then 8
out=gpurun_out
export PYTHONPATH=$PWD:$PYTHONPATH
run() { for wl in heat512 heat256; do timeout 300 python bench.py --workload $wl --no-cpu-baseline > $out/bench_${wl}_$1.json 2>$out/bench_$1.err; done; }
run P10
for n in 12 8; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared --split-compile=0 -DGRAM_PROD_WARPS=$n epde_b200/csrc/epde_b200.cu -o epde_b200/libepde_b200.so > $out/build_P$n.log 2>&1
  run P$n
done
