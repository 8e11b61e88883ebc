set -x
timeout 300 ncu --set full --clock-control none --import-source on -k regex:cross_reg_kernel -s 20 -c 1 -o gpurun_out/r02_prof_cross -f python profiles/interaction_bench.py > gpurun_out/r02_inter_ncu1.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:fm_dense_reg_kernel -s 20 -c 1 -o gpurun_out/r02_prof_fmdense -f python profiles/interaction_bench.py > gpurun_out/r02_inter_ncu2.log 2>&1
tail -2 gpurun_out/r02_inter_ncu2.log
