timeout 120 tools/umma_rate > gpurun_out/r12_umma_rate.txt 2>&1
