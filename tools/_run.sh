timeout 300 python -m pytest tests/test_gpu_pair.py tests/test_moe.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r8_tests.log
echo "# direct" > gpurun_out/r8_dds.txt
timeout 300 python tools/sweep.py --workloads c3b,c5_dds_tn_e8 --iters 30 >> gpurun_out/r8_dds.txt 2>&1
echo "# BSMM_DDS_DIRECT=0" >> gpurun_out/r8_dds.txt
BSMM_DDS_DIRECT=0 timeout 300 python tools/sweep.py --workloads c3b,c5_dds_tn_e8 --iters 30 >> gpurun_out/r8_dds.txt 2>&1
