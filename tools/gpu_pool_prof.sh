#!/bin/bash
# Second visit for the job-pool kernel: steady-state (128-wave) timing and stage statistics, launch metrics of variants 2 and 3
# side by side, and a source-counter capture of variant 3.
mkdir -p gpurun_out
OUT=gpurun_out/pool_prof.log; : > $OUT
export CASES=${CASES:-dense512}
echo "== 128 waves/launch: variant 2" | tee -a $OUT
WAVES=128 VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
echo "== 128 waves/launch: variant 3" | tee -a $OUT
WAVES=128 VARIANTS=3 POOL="${POOL:-2,1;3,1;4,1}" timeout 900 python tools/perf_probe.py 2>&1 | tee -a $OUT
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed_op_shared_ld.sum,smsp__inst_executed_op_shared_st.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum
VARIANTS=2,3 timeout 900 ncu --metrics $M --clock-control none -k regex:vpt_render -s 2 -c 2 --csv --log-file gpurun_out/pool_metrics.csv python tools/ncu_target.py dense512 64 1.0 > gpurun_out/pool_metrics.log 2>&1
grep -v "^==" gpurun_out/pool_metrics.csv | cut -d, -f5,13- | tee -a $OUT
VARIANTS=3 bash tools/gpu_ncu_fast.sh dense512 32 1.0 pool_fast | tee -a $OUT
