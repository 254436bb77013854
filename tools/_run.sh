python bench.py --steps 500 --warmup 10 > gpurun_out/r14_bench.json 2> gpurun_out/r14_bench.err
