set -x
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1.json 2> gpurun_out/bench_r1.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r1_ref.json 2> gpurun_out/bench_r1_ref.err
python tools/prof_render.py c2 8 > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2_8spp.csv python tools/prof_render.py c2 8 > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -s 35 -c 7 -o gpurun_out/prof_v4 python tools/prof_render.py c2 8 > gpurun_out/ncu_v4.log 2>&1
