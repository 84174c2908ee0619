#!/bin/bash
# One GPU-box visit: parity tests, bench line, ncu launch list + one full capture of the render kernel.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 6 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
NCU_CMD="python bench.py --steps 1 --warmup 3 --waves-per-step 1 --no-cpu-baseline"
$NCU_CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches.csv $NCU_CMD > gpurun_out/ncu_launches.log 2>&1
$NCU_CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:vpt_render_kernel -s 3 -c 1 -f -o gpurun_out/prof $NCU_CMD > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
