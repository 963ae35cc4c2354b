for v in base tpf0 base; do
  lib=""; [ $v != base ] && lib="--lib simple-path-tracer_b200/lib/variants/libspt_$v.so"
  python tools/sweep.py c2:32 c5:2 c3:8 --tag $v $lib >> gpurun_out/r1_sweep_v20.log 2>&1
done
