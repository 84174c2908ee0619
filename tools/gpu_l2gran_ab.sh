#!/bin/bash
# L2 fetch granularity A/B (cudaLimitMaxL2FetchGranularity through VPT_L2_FETCH_GRANULARITY): throughput + DRAM/L2 sector counts of the default kernel.
mkdir -p gpurun_out
OUT=gpurun_out/l2gran_ab.log; : > $OUT
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sectors_srcunit_tex_lookup_miss.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,l1tex__t_sectors_pipe_lsu_mem_global_op_ld_lookup_miss.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum
python -c "
import ctypes
rt = ctypes.CDLL('libcudart.so.12')
v = ctypes.c_size_t(0); rt.cudaFree(0); print('default cudaLimitMaxL2FetchGranularity =', rt.cudaDeviceGetLimit(ctypes.byref(v), 5), v.value)" 2>&1 | tee -a $OUT
for G in ${GRANS:-default 32 128}; do
  unset VPT_L2_FETCH_GRANULARITY
  [ $G != default ] && export VPT_L2_FETCH_GRANULARITY=$G
  echo "== granularity $G" | tee -a $OUT
  WAVES=64 CASES=dense512,wdas_cloud VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
  ONLY_FULL_CLOUD=1 WAVES=32 VARIANTS=2 timeout 900 python tools/perf_probe.py 2>&1 | tee -a $OUT
  VARIANTS=2 timeout 600 ncu --metrics $M --clock-control none -k regex:vpt_render -s 1 -c 1 --csv --log-file gpurun_out/l2gran_$G.csv python tools/ncu_target.py dense512 16 1.0 > gpurun_out/l2gran_${G}_ncu.log 2>&1
  python - <<PY | tee -a $OUT
import csv
rows = [r for r in csv.reader(open('gpurun_out/l2gran_$G.csv')) if len(r) > 5]
h = rows[0]
for r in rows[1:]:
    d = dict(zip(h, r)); print('  %-75s %s %s' % (d['Metric Name'], d['Metric Value'], d['Metric Unit']))
PY
done
