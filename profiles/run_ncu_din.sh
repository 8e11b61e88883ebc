set -x
timeout 200 python profiles/time_din.py > gpurun_out/r02_din_plain.log 2>&1 || exit 1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:din_fwd_kernel -s 3 -c 1 -o gpurun_out/r02_prof_din_fwd -f python profiles/time_din.py > gpurun_out/r02_din_ncu.log 2>&1
tail -3 gpurun_out/r02_din_ncu.log
