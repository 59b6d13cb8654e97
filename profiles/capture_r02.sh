#!/bin/bash
# Run ON the GPU box (through gpurun): one ncu --set full capture of the two streaming kernels of the round-2 engine
# (column-dictionary DMMA Gram, sparse residual); leaves small text files under gpurun_out/.
# usage: bash profiles/capture_r02.sh <tag> [bench args]
tag=${1:-r02}
shift
out=gpurun_out
mkdir -p $out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"gram_dmma|residual_cols" -c 2 -o /tmp/full_$tag -f \
    python bench.py --no-cpu-baseline --steps 1 --warmup 1 "$@" > $out/ncu_full_$tag.log 2>&1
ncu -i /tmp/full_$tag.ncu-rep --page raw --csv > $out/full_raw_$tag.csv 2>/dev/null
ncu -i /tmp/full_$tag.ncu-rep --page source --csv --print-source cuda,sass > /tmp/src_$tag.csv 2>/dev/null
python profiles/source_hotspots.py /tmp/src_$tag.csv 30 > $out/hotspots_$tag.txt 2>&1
python profiles/summarize_ncu.py "$tag" - /tmp/full_$tag.ncu-rep > $out/summary_$tag.txt 2>&1
ls -la $out
