# round-end evidence at HEAD: GPU suite, default bench line, the other configs, ncu launch list of the same command
set -x
(timeout 900 python -m pytest tests -m gpu -x -q) > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_v10.json 2> gpurun_out/bench_v10.err || exit 1
for wl in c1 c3 c4; do python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v10_$wl.json 2> gpurun_out/bench_v10_$wl.err; done
python bench.py --workload c5 --spp 512 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v10_c5_512spp.json 2> gpurun_out/bench_v10_c5.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_bench_v10.csv python bench.py --steps 1 --warmup 1 --spp 64 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
set +x
for f in gpurun_out/bench_v10*.json; do python -c "
import json,sys; d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', round(d['value'],1), round(d['e2e']['value'],1), d['config']['workload'][:30], d['config']['pool_slots'], round(d.get('roofline',{}).get('frac',0),3), d['roofline']['traffic'])"; done
