#!/bin/bash
# One gpurun call: GPU parity tests, bench (both arms), ncu launch lists and one --set full capture per dominant kernel.
# usage: tools/gpu_round.sh <tag>   (outputs under gpurun_out/<tag>_*)
tag=${1:-rXX}
mkdir -p gpurun_out
nproc > gpurun_out/${tag}_host.txt; lscpu | grep -i "model name" >> gpurun_out/${tag}_host.txt
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?" >> gpurun_out/${tag}_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2>> gpurun_out/${tag}_bench.err
timeout 300 python tools/prof_trk.py 10240 3 > gpurun_out/${tag}_prof_trk.log 2>&1 && {
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_trk_launches.csv python tools/prof_trk.py 10240 3 > gpurun_out/${tag}_ncu_trk.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:trk_ -s 1 -c 1 -o gpurun_out/${tag}_trk_full -f python tools/prof_trk.py 10240 3 >> gpurun_out/${tag}_ncu_trk.log 2>&1
ncu -i gpurun_out/${tag}_trk_full.ncu-rep --page details --csv > gpurun_out/${tag}_trk_details.csv 2>/dev/null
}
timeout 300 python tools/prof_acq.py > gpurun_out/${tag}_prof_acq.log 2>&1 && {
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_acq_launches.csv python tools/prof_acq.py > gpurun_out/${tag}_ncu_acq.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:acq_inv -s 1 -c 1 -o gpurun_out/${tag}_acq_full -f python tools/prof_acq.py >> gpurun_out/${tag}_ncu_acq.log 2>&1
ncu -i gpurun_out/${tag}_acq_full.ncu-rep --page details --csv > gpurun_out/${tag}_acq_details.csv 2>/dev/null
}
tail -3 gpurun_out/${tag}_pytest.log; cat gpurun_out/${tag}_bench.json; cat gpurun_out/${tag}_prof_trk.log gpurun_out/${tag}_prof_acq.log
