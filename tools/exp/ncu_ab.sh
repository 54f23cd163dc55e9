#!/bin/bash
# A/B of two library builds under ncu: a handful of counters of render_frame (third launch = history-scheduled frame)
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum,l1tex__t_sector_hit_rate.pct,lts__t_sectors_srcunit_tex.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__inst_executed_op_shared_ld.sum,smsp__inst_executed_op_shared_st.sum,smsp__inst_executed_op_local_ld.sum,smsp__inst_executed_op_local_st.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_alu.sum,smsp__inst_executed_op_branch.sum,sm__inst_issued.avg.pct_of_peak_sustained_active
mkdir -p gpurun_out
for lib in "$@"; do
  n=$(basename $lib .so)
  RTB_LIB=$PWD/$lib ncu --metrics $M --clock-control none -k regex:render_frame -s 3 -c 1 --csv --log-file gpurun_out/ncu_ab_$n.csv python bench.py --steps 3 --warmup 3 --no-cpu > /dev/null 2>&1
  echo "== $n"; python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/ncu_ab_$n.csv")) if len(r)>10]
h=rows[0]; 
for r in rows[1:]:
    d=dict(zip(h,r)); print("%-70s %s" % (d["Metric Name"], d["Metric Value"]))
PY
done
