#!/bin/bash
# ncu --set full of one klm_seq_rg launch (cfg1's model, 3 000 characters); the same command runs plainly first
# usage (under gpurun): bash scripts/gpu_ncu_rg.sh <tag>
OUT=gpurun_out/$1; mkdir -p $OUT
CMD="python bench.py --workload cfg1 --steps 1 --warmup 3 --chars 3000 --no-cpu"
timeout 300 $CMD > $OUT/rg_plain.log 2>&1 || { echo "plain run failed"; tail -3 $OUT/rg_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:klm_seq_rg -s 2 -c 1 -o $OUT/prof_rg $CMD > $OUT/ncu_rg.log 2>&1
tail -3 $OUT/ncu_rg.log
ncu -i $OUT/prof_rg.ncu-rep --page raw --csv > $OUT/rg_raw.csv 2>/dev/null
python - $OUT/rg_raw.csv <<'PY'
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
if len(rows) >= 3:
    head, vals = rows[0], rows[2]
    want = ('Kernel Name', 'gpu__time_duration.sum', 'launch__grid_size', 'launch__cluster_size', 'launch__registers_per_thread', 'launch__block_size',
            'sm__inst_executed_pipe_fma.sum', 'smsp__inst_executed.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
            'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__inst_executed_op_shared_st.sum', 'smsp__inst_executed_op_shared_ld.sum',
            'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__average_warp_latency_issue_stalled_barrier.ratio',
            'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
            'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__issue_active.avg.pct_of_peak_sustained_active')
    for k in want:
        if k in head:
            print('%-80s %s %s' % (k, vals[head.index(k)], rows[1][head.index(k)]))
PY
