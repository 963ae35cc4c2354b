set -x
timeout 600 python -m pytest tests/test_cpp_host.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/cpp_host_2gpu_pytest.log
python -c "
import json,sys
sys.path.insert(0,'simple-path-tracer_b200')
from spt import scenes
json.dump(scenes.c2(), open('gpurun_out/c2.json','w'))
json.dump(dict(scenes.PT_JSON), open('gpurun_out/pt.json','w'))
"
rm -f gpurun_out/cpp_host_c2_g2.log
for g in 1 2; do for i in 1 2; do simple-path-tracer_b200/lib/spt_host --scene gpurun_out/c2.json --base-path assets/scene.json --renderer gpurun_out/pt.json --width 1920 --height 1080 --output gpurun_out/c2_cpp_g$g.png --gpus $g >> gpurun_out/cpp_host_c2_g2.log 2>&1; done; done
cmp gpurun_out/c2_cpp_g1.png gpurun_out/c2_cpp_g2.png && echo IDENTICAL >> gpurun_out/cpp_host_c2_g2.log
cat gpurun_out/cpp_host_2gpu_pytest.log gpurun_out/cpp_host_c2_g2.log
