timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r1_pytest_v4.log
L=simple-path-tracer_b200/lib/variants
python tools/sweep.py c2:32 c5:2 --pools 4194304,8388608,16777216 --tag base > gpurun_out/r1_sweep_v4.log 2>&1
for v in inl inl_i2 inl_o1 i2 o1; do
  python tools/sweep.py c2:32 c5:2 --pools 4194304 --tag $v --lib $L/libspt_$v.so >> gpurun_out/r1_sweep_v4.log 2>&1
done
