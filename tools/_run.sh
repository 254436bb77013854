for N in 8 4 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2960$N bench.py --gpus $N --workload c5_layer --steps 10 --warmup 3 > gpurun_out/r11_layer_n$N.json 2> gpurun_out/r11_layer_n$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N bench.py --gpus $N --steps 1000 --warmup 20 > gpurun_out/r11_c2_n$N.json 2> gpurun_out/r11_c2_n$N.err
done
python bench.py --workload c5_layer --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r11_layer_n1.json 2> gpurun_out/r11_layer_n1.err
python bench.py --steps 1000 --warmup 20 --no-cpu-baseline > gpurun_out/r11_c2_n1.json 2> gpurun_out/r11_c2_n1.err
python -m pytest tests/test_moe.py -m gpu -q 2>&1 | tail -3 > gpurun_out/r11_moe_tests.log
