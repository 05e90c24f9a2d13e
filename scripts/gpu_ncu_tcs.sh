#!/bin/bash
# ncu capture of one launch of the streamed-weights kernel (cfg5 model, 64 streams x 600 characters)
OUT=gpurun_out/${1:-ncu_tcs}; mkdir -p $OUT
CMD="python scripts/tc_trace.py cfg5 600"
timeout 300 $CMD > $OUT/plain.log 2>&1 || { tail -5 $OUT/plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:klm_seq_tcs -s 1 -c 1 -o $OUT/prof_tcs $CMD > $OUT/ncu_full.log 2>&1
tail -3 $OUT/ncu_full.log
ncu -i $OUT/prof_tcs.ncu-rep --page raw --csv > $OUT/prof_tcs_raw.csv 2>/dev/null
python - <<PY
import csv
rows=list(csv.reader(open('$OUT/prof_tcs_raw.csv')))
hdr, units, vals = rows[0], rows[1], rows[2]
want=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','lts__t_sector_hit_rate.pct','lts__t_bytes.sum','sm__inst_executed_pipe_tensor','sm__pipe_tensor_op','sm__inst_executed_pipe_xu','sm__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__data_pipe_lsu_wavefronts_mem_shared','lts__throughput','smsp__cycles_active.avg','sm__cycles_elapsed.max','launch__registers_per_thread','smsp__inst_executed_pipe_xu','pipe_tmem','tensor']
for h,u,v in zip(hdr,units,vals):
    if any(w in h for w in want): print(h,u,v)
PY
