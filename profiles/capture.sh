#!/bin/bash
# Run ON the GPU box (through gpurun): bench lines, ncu launch list, one ncu --set full capture of the two
# streaming kernels; leaves only small text files under gpurun_out/ (the .ncu-rep itself exceeds the 64 MiB
# copy-back limit, so it is summarised here and deleted).   usage: bash profiles/capture.sh <tag>
tag=${1:-vX}
out=gpurun_out
mkdir -p $out
timeout 400 python bench.py > $out/bench_heat512_$tag.json 2> $out/bench_heat512_$tag.err
timeout 200 python bench.py --norms both --no-cpu-baseline > $out/bench_heat512_both_$tag.json 2>/dev/null
timeout 200 python bench.py --workload heat256 --no-cpu-baseline > $out/bench_heat256_$tag.json 2>/dev/null
timeout 200 python bench.py --workload wave1024 --no-cpu-baseline > $out/bench_wave1024_$tag.json 2>/dev/null
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/launches_heat512_$tag.csv \
    python bench.py --no-cpu-baseline --steps 2 --warmup 1 > $out/ncu_launch_$tag.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"gram_tile|residual_tile" -c 2 -o /tmp/full_$tag -f \
    python bench.py --no-cpu-baseline --steps 1 --warmup 1 > $out/ncu_full_$tag.log 2>&1
ncu -i /tmp/full_$tag.ncu-rep --page raw --csv > $out/full_raw_$tag.csv 2>/dev/null
ncu -i /tmp/full_$tag.ncu-rep --page source --csv --print-source cuda,sass > /tmp/src_$tag.csv 2>/dev/null
python profiles/source_hotspots.py /tmp/src_$tag.csv 25 > $out/hotspots_$tag.txt 2>&1
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum \
    --clock-control none -k regex:"residual_tile" -c 1 --csv --log-file $out/resid_metrics_$tag.csv \
    python bench.py --no-cpu-baseline --steps 1 --warmup 1 > /dev/null 2>&1
ls -la $out
