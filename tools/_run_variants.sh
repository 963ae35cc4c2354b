for v in base minb9 minb10; do
  lib=""; [ $v != base ] && lib="--lib simple-path-tracer_b200/lib/variants/libspt_$v.so"
  python tools/sweep.py c2:32 c5:2 c3:8 --tag $v $lib >> gpurun_out/r1_sweep_v22.log 2>&1
done
