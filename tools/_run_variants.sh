SPT_BVH_BUILD=device timeout 800 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r1_pytest_ploc.log
SPT_BVH_BUILD=device_lbvh timeout 800 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "ray_batches or device_built" 2>&1 | tail -4 >> gpurun_out/r1_pytest_ploc.log
for m in host device device_lbvh; do SPT_BVH_BUILD=$m python tools/sweep.py c2:32 c5:2 c3:8 --tag bvh_$m >> gpurun_out/r1_sweep_v23.log 2>&1; done
for a in ploc lbvh; do SPT_BVH_ALGO=$a timeout 600 python tools/ray_batch_bench.py c5b --log2n 22 --out gpurun_out/ray_batches_c5b_$a.json > gpurun_out/ray_batches_c5b_$a.log 2>&1; done
