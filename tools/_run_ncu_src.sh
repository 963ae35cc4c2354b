# one ncu --set full capture with source counters of a steady-state k_extend, k_connect and k_shade<0> launch (C2, 8 spp)
set -x
python tools/prof_render.py c2 8 > gpurun_out/prof_plain.log 2>&1 || exit 1
for k in k_extend k_connect; do
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:$k -s 6 -c 1 -o gpurun_out/ncu_$k -f python tools/prof_render.py c2 8 > gpurun_out/ncu_$k.log 2>&1
done
timeout 600 ncu --set full --import-source on --clock-control none -k regex:'k_shade' -s 12 -c 3 -o gpurun_out/ncu_k_shade -f python tools/prof_render.py c2 8 > gpurun_out/ncu_k_shade.log 2>&1
ls -la gpurun_out/
