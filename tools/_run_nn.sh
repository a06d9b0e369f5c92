cd /root/repo
timeout 900 python -m pytest tests/test_gpu_nn.py tests/test_gpu_ops.py -x -q -m gpu 2>&1 | tail -30 > gpurun_out/t_nn.log
timeout 900 python bench.py --workload finetune --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ft_bench.json 2> gpurun_out/ft_bench.err
echo "rc=$?" >> gpurun_out/ft_bench.err
