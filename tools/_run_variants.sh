L=simple-path-tracer_b200/lib/variants
python tools/sweep.py c2:32 c5:2 c3:8 --tag base > gpurun_out/r1_sweep_v7.log 2>&1
for v in t8 t12 t12i8 t16i12 i8; do
  python tools/sweep.py c2:32 c5:2 c3:8 --tag $v --lib $L/libspt_$v.so >> gpurun_out/r1_sweep_v7.log 2>&1
done
