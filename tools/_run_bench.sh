set -x
for wl in c1 c3 c4; do python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r1_$wl.json 2> gpurun_out/bench_r1_$wl.err; done
python bench.py --workload c5 --spp 512 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r1_c5_512spp.json 2> gpurun_out/bench_r1_c5.err
for wl in c1 c3 c4 c5_512spp; do python -c "
import json,sys; d=json.loads(open('gpurun_out/bench_r1_$wl.json').read().strip().splitlines()[-1]); print('$wl', round(d['value'],1), round(d['e2e']['value'],1), d['config']['workload'], round(d['roofline']['frac'],3))"; done
