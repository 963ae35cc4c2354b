# sweep of the tuning variants under simple-path-tracer_b200/lib/variants (see tools/build_variants.py); ITEMS = sweep.py items
log=gpurun_out/sweep_variants.log; : > $log
for v in cur $VARIANTS; do
  lib=""; [ $v != cur ] && lib="--lib simple-path-tracer_b200/lib/variants/libspt_$v.so"
  python tools/sweep.py ${ITEMS:-c2:32 c5:2 c3:8} --tag $v $lib >> $log 2>&1
done
python tools/show_sweep.py $log 2>/dev/null || cat $log
