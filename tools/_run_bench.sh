set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r1_pytest_v12.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_r1.json 2> gpurun_out/bench_r1.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r1_ref.json 2> gpurun_out/bench_r1_ref.err
python -c "
import json,sys
sys.path.insert(0,'simple-path-tracer_b200')
from spt import scenes
json.dump(scenes.c2(), open('gpurun_out/c2.json','w'))
json.dump(dict(scenes.PT_JSON), open('gpurun_out/pt.json','w'))
"
for i in 1 2; do simple-path-tracer_b200/lib/spt_host --scene gpurun_out/c2.json --base-path assets/scene.json --renderer gpurun_out/pt.json --width 1920 --height 1080 --output gpurun_out/c2_cpp.png >> gpurun_out/cpp_host_c2.log 2>&1; done
python tools/prof_render.py c2 8 > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2_8spp.csv python tools/prof_render.py c2 8 > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -s 40 -c 9 -o gpurun_out/prof_v6 python tools/prof_render.py c2 8 > gpurun_out/ncu_v6.log 2>&1
