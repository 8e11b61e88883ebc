# round-2 evidence: launch list of the bench command and a full capture of the main kernel (each after the plain command exited 0)
set -x
timeout 300 python bench.py --steps 2 --warmup 3 > gpurun_out/r02w_plain.log 2>&1 || exit 1
timeout 800 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02w_launches_bench.csv python bench.py --steps 2 --warmup 3 > gpurun_out/r02w_ncu1.log 2>&1
echo rc1=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ctr_fused_kernel -s 6 -c 1 -o gpurun_out/r02w_prof_fused -f python bench.py --steps 2 --warmup 3 > gpurun_out/r02w_ncu2.log 2>&1
echo rc2=$?
