#!/bin/bash
# One GPU-box visit. `gpu_round.sh` (default, "bench"): smoke, parity tests, both bench arms, then the ncu launch list of
# the bench command (`gpu_round.sh launches`: only that). `gpu_round.sh full`: one --set full capture of the render kernel. One ncu use per visit, and each
# ncu run only after the same command exited 0 without ncu.
mkdir -p gpurun_out
NCU_CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
launch_list() { # per-launch duration + executed warp/thread instructions of the bench command (one pass each, no replay cost)
  $NCU_CMD > gpurun_out/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv $NCU_CMD > gpurun_out/ncu_launches.log 2>&1
}
if [ "${1:-bench}" = "launches" ]; then launch_list; tail -4 gpurun_out/launches.csv; exit 0; fi
if [ "${1:-bench}" = "full" ]; then
  VARIANTS=2 bash tools/gpu_ncu.sh dense512 ${2:-16} 1.0 r01_dense512_sched_full
  exit 0
fi
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench reference rc=$?"
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
launch_list
ls -la gpurun_out | tail -30
