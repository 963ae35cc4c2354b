timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r1_pytest_v3.log
L=simple-path-tracer_b200/lib/variants
python tools/sweep.py c2:32 c5:2 c3:8 c4:8 --pools 1048576,4194304 --tag base > gpurun_out/r1_sweep_v3.log 2>&1
for v in tri1 tri3 idle1 idle12; do
  python tools/sweep.py c2:32 c5:2 --pools 4194304 --tag $v --lib $L/libspt_$v.so >> gpurun_out/r1_sweep_v3.log 2>&1
done
python tools/prof_render.py c2 8 4194304 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -s 35 -c 7 -o gpurun_out/prof_v3 python tools/prof_render.py c2 8 4194304 > gpurun_out/ncu_v3.log 2>&1
