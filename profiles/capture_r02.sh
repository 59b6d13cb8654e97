#!/bin/bash
# Run ON the GPU box (through gpurun).  Round-2 evidence set in one call, small text files under gpurun_out/:
#   1. bench lines WITHOUT a profiler (heat512 default with cpu_baseline, heat512 --norms both, heat512 distinct
#      population, heat256, wave1024, wide16 / wide30 populations)
#   2. ncu launch lists (gpu__time_duration.sum, --clock-control none) of the heat512 and heat256 commands
#   3. one ncu --set full capture of the two streaming kernels (pass A DMMA Gram, pass B residual) + source hot spots
#   4. the measured FP64 peaks (epde_measure_fp64_peak)
# Numbers printed by a run under ncu are never bench values: only (1) is.
# usage: bash profiles/capture_r02.sh <tag>
tag=${1:-r02}
out=gpurun_out
mkdir -p $out
run() { name=$1; shift; timeout 900 python bench.py "$@" > $out/bench_${name}_$tag.json 2> $out/bench_${name}_$tag.err; echo "$name rc=$?"; }
run heat512 --workload heat512
run heat512_both --workload heat512 --norms both --no-cpu-baseline
run heat512_distinct --workload heat512 --population distinct --no-cpu-baseline
run heat256 --workload heat256 --no-cpu-baseline
run wave1024 --workload wave1024 --no-cpu-baseline
run heat512_wide16 --workload heat512 --population wide16 --no-cpu-baseline --steps 2 --warmup 3
run heat512_wide30 --workload heat512 --population wide30 --no-cpu-baseline --steps 2 --warmup 3
python - > $out/fp64_peak_$tag.txt 2>&1 <<'EOF'
import epde_b200
print('epde_measure_fp64_peak (dfma, dmma TFLOP/s):', epde_b200.Engine.fp64_peak(0, 2.0))
EOF
for wl in heat512 heat256; do
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/launches_${wl}_$tag.csv \
      python bench.py --workload $wl --no-cpu-baseline --steps 2 --warmup 1 > $out/ncu_launches_${wl}_$tag.log 2>&1
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gram_dmma|residual_cols" -c 2 -o /tmp/full_$tag -f \
    python bench.py --workload heat512 --no-cpu-baseline --steps 1 --warmup 1 > $out/ncu_full_$tag.log 2>&1
ncu -i /tmp/full_$tag.ncu-rep --page raw --csv > $out/full_raw_$tag.csv 2>/dev/null
ncu -i /tmp/full_$tag.ncu-rep --page source --csv --print-source cuda,sass > /tmp/src_$tag.csv 2>/dev/null
python profiles/source_hotspots.py /tmp/src_$tag.csv 30 > $out/hotspots_$tag.txt 2>&1
python profiles/summarize_ncu.py "$tag heat512" $out/launches_heat512_$tag.csv /tmp/full_$tag.ncu-rep > $out/summary_$tag.txt 2>&1
python profiles/brief_ncu.py $out/full_raw_$tag.csv > $out/brief_$tag.txt 2>&1
ls -la $out
