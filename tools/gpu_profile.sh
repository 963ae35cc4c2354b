# ncu evidence for profiles/ (run under gpurun; writes into gpurun_out/): TAG names the output files.
#   1. launch list (gpu__time_duration.sum) of one short bench.py run — per-kernel share of the step
#   2. ncu --set full with source counters of one steady-state k_extend, k_connect and the k_shade launches of an iteration (C2, 8 spp)
TAG=${TAG:-r02}
set -x
python bench.py --spp 64 --steps 1 --warmup 3 --no-parity --no-cpu-baseline --no-secondary > gpurun_out/${TAG}_bench_64spp_plain.json 2>/dev/null || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/${TAG}_launches_bench_64spp.csv \
  python bench.py --spp 64 --steps 1 --warmup 3 --no-parity --no-cpu-baseline --no-secondary > gpurun_out/${TAG}_ncu_launches.log 2>&1
python tools/prof_render.py c2 8 > gpurun_out/${TAG}_prof_plain.log 2>&1 || exit 1
for k in k_extend k_connect; do
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:$k -s 6 -c 1 -o gpurun_out/${TAG}_ncu_$k -f python tools/prof_render.py c2 8 > gpurun_out/${TAG}_ncu_$k.log 2>&1
done
timeout 600 ncu --set full --import-source on --clock-control none -k regex:'k_shade' -s 12 -c 3 -o gpurun_out/${TAG}_ncu_k_shade -f python tools/prof_render.py c2 8 > gpurun_out/${TAG}_ncu_k_shade.log 2>&1
# the k_extend launch of the full-size pool (2^24 slots): DRAM traffic per launch for bench.py's roofline.traffic
timeout 900 ncu --set full --clock-control none -k regex:k_extend -s 10 -c 1 -o gpurun_out/${TAG}_ncu_extend_pool24 -f python tools/prof_render.py c2 64 > gpurun_out/${TAG}_ncu_extend_pool24.log 2>&1
ls -la gpurun_out/
