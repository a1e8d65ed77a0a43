#!/bin/bash
# ncu launch list + full capture of the acquisition inverse kernel; usage: tools/gpu_ncu_acq.sh <tag>
tag=${1:-rXX}
mkdir -p gpurun_out
python tools/prof_acq.py 25000 3 > gpurun_out/${tag}_prof_acq.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_acq_launches.csv python tools/prof_acq.py 25000 3 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:acq_inv -s 1 -c 1 -o gpurun_out/${tag}_acq_full -f python tools/prof_acq.py 25000 3 > /dev/null 2>&1
ncu -i gpurun_out/${tag}_acq_full.ncu-rep --page details --csv > gpurun_out/${tag}_acq_details.csv 2>/dev/null
ncu -i gpurun_out/${tag}_acq_full.ncu-rep --page source --csv > gpurun_out/${tag}_acq_source.csv 2>/dev/null
cat gpurun_out/${tag}_prof_acq.log
