#!/bin/bash
# ncu --set full capture of the render kernel variants on one config. $1 = config, $2 = waves, $3 = scale, $4 = out name
CFG=${1:-dense512}; WAVES=${2:-4}; SCALE=${3:-1.0}; OUT=${4:-prof}
CMD="python tools/ncu_target.py $CFG $WAVES $SCALE"
NV=$(python -c "import os; print(len(os.environ.get('VARIANTS','0,1').split(',')))")
$CMD > gpurun_out/${OUT}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:vpt_render -s $NV -c $NV -f -o gpurun_out/$OUT $CMD > gpurun_out/${OUT}_ncu.log 2>&1
tail -3 gpurun_out/${OUT}_plain.log; tail -3 gpurun_out/${OUT}_ncu.log
