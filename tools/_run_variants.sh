python tools/sweep.py c2:32 c5:2 c3:8 --tag host >> gpurun_out/r1_sweep_v24.log 2>&1
for v in base r8 r32 r100; do
  lib=""; [ $v != base ] && lib="--lib simple-path-tracer_b200/lib/variants/libspt_$v.so"
  SPT_BVH_BUILD=device python tools/sweep.py c2:32 c5:2 c3:8 --tag ploc_$v $lib >> gpurun_out/r1_sweep_v24.log 2>&1
done
