set -x
timeout 200 python profiles/time_pnn.py > gpurun_out/r02_pnn_plain.log 2>&1 || exit 1
cat gpurun_out/r02_pnn_plain.log | cut -c1-220
timeout 500 ncu --set full --clock-control none --import-source on -k regex:pnn_front_kernel -s 3 -c 1 -o gpurun_out/r02_prof_pnn_front -f python profiles/time_pnn.py > gpurun_out/r02_pnn_ncu.log 2>&1
tail -3 gpurun_out/r02_pnn_ncu.log
