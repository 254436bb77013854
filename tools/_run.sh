timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r13_tests.log
W=c1,c2,c3a,c3b,c4_dsd_25,c4_sdd_50,c5_dsd_e8
echo "# programmatic dependent launch (default)" > gpurun_out/r13_pdl.txt
timeout 300 python tools/sweep.py --workloads $W --iters 100 >> gpurun_out/r13_pdl.txt 2>&1
echo "# BSMM_PDL=0" >> gpurun_out/r13_pdl.txt
BSMM_PDL=0 timeout 300 python tools/sweep.py --workloads $W --iters 100 >> gpurun_out/r13_pdl.txt 2>&1
python bench.py --workload c5_layer --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r13_layer.json 2> gpurun_out/r13_layer.err
BSMM_PDL=0 python bench.py --workload c5_layer --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r13_layer_nopdl.json 2> gpurun_out/r13_layer_nopdl.err
