#!/bin/bash
# Faster ncu capture (selected sections + source counters) of one render-kernel variant. $1 cfg $2 waves $3 scale $4 out
CFG=${1:-dense512}; WAVES=${2:-8}; SCALE=${3:-1.0}; OUT=${4:-prof_fast}
CMD="python tools/ncu_target.py $CFG $WAVES $SCALE"
NV=$(python -c "import os; print(len(os.environ.get('VARIANTS','2').split(',')))")
$CMD > gpurun_out/${OUT}_plain.log 2>&1 &&
ncu --section SpeedOfLight --section LaunchStats --section Occupancy --section WarpStateStats --section SchedulerStats --section SourceCounters --section MemoryWorkloadAnalysis --section InstructionStats \
    --clock-control none --import-source on -k regex:vpt_render -s $NV -c $NV -f -o gpurun_out/$OUT $CMD > gpurun_out/${OUT}_ncu.log 2>&1
tail -2 gpurun_out/${OUT}_plain.log; tail -3 gpurun_out/${OUT}_ncu.log
