timeout 300 python -m pytest tests/test_gpu_pair.py tests/test_moe.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r7_tests.log
timeout 300 python tools/sweep.py --workloads c5_sdd_e8,c5_sdd_nt_e8,m_sdd_8_8_32_1024,m_sdd_4_32_32_4096 --iters 30 > gpurun_out/r7_sdd.txt 2>&1
