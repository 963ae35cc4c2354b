# round-end evidence at HEAD: default bench line, reference arm, the other configs, ncu launch list of the same command
set -x
python bench.py > gpurun_out/bench_v9.json 2> gpurun_out/bench_v9.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_v9_reference.json 2> gpurun_out/bench_v9_reference.err
for wl in c1 c3 c4; do python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v9_$wl.json 2> gpurun_out/bench_v9_$wl.err; done
python bench.py --workload c5 --spp 512 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v9_c5_512spp.json 2> gpurun_out/bench_v9_c5.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_bench_v9.csv python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2_8spp_v9.csv python tools/prof_render.py c2 8 > gpurun_out/ncu_prof.log 2>&1
for f in gpurun_out/bench_v9*.json; do python -c "
import json,sys; d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', round(d['value'],1), round(d['e2e']['value'],1), d['config']['workload'][:40], round(d.get('roofline',{}).get('frac',0),3))"; done
