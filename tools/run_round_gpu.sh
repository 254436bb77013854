# One GPU call that refreshes the round's evidence: tests, both bench arms, the legacy-kernel
# comparison sweep, the ncu launch list and one ncu --set full capture of the headline kernel.
set -x
R=${1:-r9}
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/${R}_tests.log
python bench.py > gpurun_out/${R}_bench.json 2> gpurun_out/${R}_bench.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${R}_bench_ref.json 2> gpurun_out/${R}_bench_ref.err
python bench.py --workload c5_layer --no-cpu-baseline > gpurun_out/${R}_layer.json 2> gpurun_out/${R}_layer.err
python tools/sweep.py --ref --iters 20 > gpurun_out/${R}_sweep_ref.txt 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${R}_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${R}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --kernel-name regex:bsmm_pair --launch-skip 5 -c 1 -o gpurun_out/${R}_c2_full -f python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${R}_ncu2.log 2>&1
ls -la gpurun_out | tail -12
