#!/bin/bash
# usage: tools/ncu_summary.sh report.ncu-rep [top-lines]   -> key metrics + per-source-line instruction shares
rep=$1; top=${2:-40}
ncu -i $rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); hdr,units,vals=rows[0],rows[1],rows[2]
want=['gpu__time_duration.sum','launch__registers_per_thread','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','launch__grid_size','launch__block_size','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','smsp__thread_inst_executed_per_inst_executed.ratio','sm__inst_executed_pipe_alu.sum.pct_of_peak_sustained_active','sm__inst_executed_pipe_fma.sum.pct_of_peak_sustained_active','sm__inst_executed_pipe_fmaheavy','sm__inst_executed_pipe_xu.sum.pct_of_peak_sustained_active','sm__inst_executed_pipe_lsu.sum.pct_of_peak_sustained_active','sm__inst_executed_pipe_uniform','sm__inst_executed_pipe_cbu','sm__inst_executed_pipe_adu','dram__bytes_read.sum','dram__bytes_write.sum','smsp__sass_thread_inst_executed_op_ffma_pred_on.sum','smsp__sass_thread_inst_executed_op_fadd_pred_on.sum','smsp__sass_thread_inst_executed_op_fmul_pred_on.sum','sass__inst_executed_local','smsp__pcsamp_warps_issue_stalled','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','sm__throughput.avg.pct_of_peak_sustained_elapsed','lts__t_sectors_op_atom','lts__t_sectors_op_red','smsp__inst_executed_op_shared_atom','l1tex__t_bytes_pipe_lsu_mem_local']
for h,u,v in zip(hdr,units,vals):
    if any(h.startswith(w) for w in want) and not h.endswith('.per_second') and 'pct_of_peak_sustained_elapsed' not in h.replace('sm__throughput.avg.pct_of_peak_sustained_elapsed','') : print(f'{h:80s} {u:12s} {v}')
"
ncu -i $rep --page source --csv --print-source cuda,sass 2>/dev/null > /tmp/_src.csv
python $(dirname $0)/ncu_lines.py /tmp/_src.csv $top
