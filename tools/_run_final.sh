# round-end evidence at HEAD: GPU suite, smoke, default bench line + reference arm, C1 (the other configs and the ncu launch
# list were taken at v10, same kernels)
set -x
(timeout 900 python -m pytest tests -m gpu -x -q) > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_v13.json 2> gpurun_out/bench_v13.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_v13_reference.json 2> gpurun_out/bench_v13_reference.err
python bench.py --workload c1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v13_c1.json 2> gpurun_out/bench_v13_c1.err
set +x
for f in gpurun_out/bench_v13.json gpurun_out/bench_v13_c1.json gpurun_out/bench_v13_reference.json; do python -c "
import json,sys; d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', round(d['value'],2), round(d['e2e']['value'],2), d['config']['workload'][:30], d['config'].get('pool_slots'), round((d.get('roofline') or {}).get('frac',0),3))"; done
