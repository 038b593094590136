#!/bin/bash
# usage: tools/ncu_capture.sh <workload> <kernel regex> <count> <tag> [extra bench args]
# Runs the bench once WITHOUT ncu (must exit 0), then once under `ncu --set full` for the kernels matching the regex, and leaves
# text exports (details page, raw metrics csv) under gpurun_out/ -- the .ncu-rep itself is too large to travel.
set -e
wl=$1; rx=$2; cnt=$3; tag=$4; shift 4
python bench.py --workload $wl --steps 2 --warmup 1 --no-cpu --no-e2e --no-verify --no-sweep "$@" > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err
ncu --set full --clock-control none --import-source on -k regex:"$rx" -c $cnt -o /tmp/$tag \
    python bench.py --workload $wl --steps 1 --warmup 1 --no-cpu --no-e2e --no-verify --no-sweep "$@" > gpurun_out/${tag}_ncu.log 2>&1
ncu -i /tmp/$tag.ncu-rep --page details > gpurun_out/${tag}_details.txt
ncu -i /tmp/$tag.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv
grep -E "^  [a-z_:<>A-Za-z0-9, ()]+\(|Duration|DRAM Throughput|Registers Per|Achieved Occupancy|Executed Ipc Active|Issue Slots Busy|No Eligible" gpurun_out/${tag}_details.txt | head -60
