for v in base nostash base; do
  lib=""; [ $v != base ] && lib="--lib simple-path-tracer_b200/lib/variants/libspt_$v.so"
  python tools/sweep.py c2:32 c5:2 c3:8 c4:8 --tag $v $lib >> gpurun_out/r1_sweep_v21.log 2>&1
done
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3 >> gpurun_out/r1_sweep_v21.log
