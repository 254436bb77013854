set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r1_tests.log
python bench.py > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r1_bench_ref.json 2> gpurun_out/r1_bench_ref.err
python tools/sweep.py --ref --iters 20 > gpurun_out/r1_sweep_ref.txt 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r1_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name regex:bsmm_gemm --launch-skip 5 -c 1 -o gpurun_out/r1_c2_full -f python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r1_ncu2.log 2>&1
ls -la gpurun_out
