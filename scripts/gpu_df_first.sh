OUT=gpurun_out/r2o; mkdir -p $OUT
for PUB in 2 1; do for REL in 1 0; do
  echo "=== KLM_TC_PUB=$PUB KLM_TC_PUB_RELEASE=$REL" >> $OUT/pub_modes.txt
  KLM_TC_PUB=$PUB KLM_TC_PUB_RELEASE=$REL timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --chars 30000 2>/dev/null | cut -c1-120 >> $OUT/pub_modes.txt
done; done
cat $OUT/pub_modes.txt
timeout 300 python scripts/tc_trace.py cfg2 3000 > $OUT/timeline_tc_release.txt 2>&1; cat $OUT/timeline_tc_release.txt
KLM_TC_PUB=1 timeout 300 python scripts/tc_trace.py cfg2 3000 > $OUT/timeline_tc_pub1.txt 2>&1; cat $OUT/timeline_tc_pub1.txt
