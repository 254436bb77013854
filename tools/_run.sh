set -x
timeout 900 python -m pytest tests/test_gpu_streamk.py -m gpu -x -q 2>&1 | tail -30 > gpurun_out/sk2_tests.log
timeout 900 python -m pytest tests/test_gpu_pair.py tests/test_gpu_scale.py tests/test_gpu_moe_topology.py -m gpu -q 2>&1 | tail -15 > gpurun_out/sk2_tests2.log
for m in 0 1; do
BSMM_SPLITK=$m timeout 600 python tools/sweep.py --workloads c2,c5_dsd_e8,c5_dsd_tn_e8,m_dsd_8_2_2_1024,m_dsd_8_4_4_2048 --iters 30 > gpurun_out/sk2_sweep_$m.txt 2>&1
done
timeout 600 python tools/sweep.py --workloads c2,c5_dsd_e8,c5_dsd_tn_e8,c5_dsd_nt_e8 --iters 30 > gpurun_out/sk2_sweep_auto.txt 2>&1
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/sk2_bench.json 2> gpurun_out/sk2_bench.err
BSMM_SPLITK=0 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/sk2_bench_off.json 2> gpurun_out/sk2_bench_off.err
