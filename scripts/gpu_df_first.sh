OUT=gpurun_out/r2u; mkdir -p $OUT
for SB in none 128-147; do
  echo "=== KLM_TC_SLOW_BLOCKS=$SB" >> $OUT/place.txt
  if [ $SB = none ]; then timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --chars 30000 2>/dev/null | cut -c1-120 >> $OUT/place.txt
  else KLM_TC_SLOW_BLOCKS=$SB timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --chars 30000 2>/dev/null | cut -c1-120 >> $OUT/place.txt; fi
done
cat $OUT/place.txt
KLM_TC_SLOW_BLOCKS=128-147 timeout 300 python scripts/tc_trace.py cfg2 3000 > $OUT/timeline_placed.txt 2>&1; cat $OUT/timeline_placed.txt
