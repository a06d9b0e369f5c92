cd /root/repo
timeout 600 python -m pytest tests/test_gpu_ops.py tests/test_gpu_nn.py -x -q -m gpu -k "linear or second_order or force_stress or rehearsal" 2>&1 | tail -15 > gpurun_out/t_lin.log
timeout 900 python bench.py --workload finetune --steps 5 --warmup 3 > gpurun_out/ft_bench.json 2> gpurun_out/ft_bench.err
echo "rc=$?" >> gpurun_out/ft_bench.err
