#!/bin/bash
# One gpurun call: ncu evidence for profiles/ (launch lists of the bench command, one --set full capture per dominant kernel, pipe metrics).
# usage: tools/gpu_profiles.sh <tag>   (outputs under gpurun_out/<tag>_*)
tag=${1:-rXX}
mkdir -p gpurun_out
PIPE=sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_fmalite.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_fp64.sum,smsp__inst_executed.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_xu_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
timeout 600 python bench.py --steps 5 --warmup 3 --quick > gpurun_out/${tag}_bench_quick.json 2> gpurun_out/${tag}_bench_quick.err || exit 1
# launch list of the same command (cold-cache, serialised: shares of the step, not absolute times)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_bench_launches.csv python bench.py --steps 5 --warmup 3 --quick > gpurun_out/${tag}_ncu_bench.log 2>&1
timeout 300 python tools/prof_trk.py 10240 3 > gpurun_out/${tag}_prof.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:trk_run -s 1 -c 1 -o gpurun_out/${tag}_trk_full -f python tools/prof_trk.py 10240 3 >> gpurun_out/${tag}_ncu.log 2>&1
timeout 600 ncu --metrics $PIPE --clock-control none -k regex:trk_run -s 1 -c 1 --csv --log-file gpurun_out/${tag}_trk_pipes.csv python tools/prof_trk.py 10240 3 >> gpurun_out/${tag}_ncu.log 2>&1
timeout 300 python tools/prof_acq.py >> gpurun_out/${tag}_prof.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:acq_inv -s 1 -c 1 -o gpurun_out/${tag}_acq_full -f python tools/prof_acq.py >> gpurun_out/${tag}_ncu.log 2>&1
timeout 600 ncu --metrics $PIPE --clock-control none -k regex:acq_inv -s 1 -c 1 --csv --log-file gpurun_out/${tag}_acq_pipes.csv python tools/prof_acq.py >> gpurun_out/${tag}_ncu.log 2>&1
timeout 300 python tools/prof_trk.py 4096 3 c4 >> gpurun_out/${tag}_prof.log 2>&1
timeout 600 ncu --metrics $PIPE --clock-control none -k regex:trk_correlate -s 1 -c 1 --csv --log-file gpurun_out/${tag}_trkc4_pipes.csv python tools/prof_trk.py 4096 3 c4 >> gpurun_out/${tag}_ncu.log 2>&1
for r in trk acq; do ncu -i gpurun_out/${tag}_${r}_full.ncu-rep --page details --csv > gpurun_out/${tag}_${r}_details.csv 2>/dev/null; done
cat gpurun_out/${tag}_prof.log; tail -3 gpurun_out/${tag}_ncu.log
