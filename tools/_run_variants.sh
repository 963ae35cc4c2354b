timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r1_pytest_v10.log
python tools/sweep.py c1:64 c2:32 c5:2 c3:8 c4:8 --tag streams > gpurun_out/r1_sweep_v10.log 2>&1
python tools/sweep.py c1:64 c2:32 c5:2 c3:8 c4:8 --tag serial --serial >> gpurun_out/r1_sweep_v10.log 2>&1
python tools/sweep.py c2:32 --pools 1048576,2097152 --tag streams >> gpurun_out/r1_sweep_v10.log 2>&1
