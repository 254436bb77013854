timeout 300 python -m pytest tests/test_gpu_pair.py -x -q 2>&1 | tail -30 > gpurun_out/r4_pair_tests.log
W=c5_sdd_e8,c5_sdd_nt_e8,c5_dsd_e8
echo "# BSMM_PAIR=1 (2-SM kernels)" > gpurun_out/r4_pair_sdd_ab.txt
timeout 300 python tools/sweep.py --workloads $W --iters 30 >> gpurun_out/r4_pair_sdd_ab.txt 2>&1
echo "# BSMM_PAIR=0 (1-SM kernel)" >> gpurun_out/r4_pair_sdd_ab.txt
BSMM_PAIR=0 timeout 300 python tools/sweep.py --workloads $W --iters 30 >> gpurun_out/r4_pair_sdd_ab.txt 2>&1
