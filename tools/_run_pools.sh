# full-length bench lines per pool size
for wl in c2 c3; do for p in 12582912 16777216; do
  python bench.py --workload $wl --pool $p --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$wl', $p, round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],2))"
done; done
for p in 12582912 16777216; do
  python bench.py --workload c5 --spp 256 --pool $p --steps 2 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c5@256', $p, round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],2))"
done
python tools/sweep.py c2:32 c2:16 c2:8 c1:256 --pools 1048576,2097152,4194304,6291456,8388608 --tag short > gpurun_out/sweep_pools_short.log 2>&1; python tools/show_sweep.py gpurun_out/sweep_pools_short.log
