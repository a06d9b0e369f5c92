cd /root/repo
timeout 1200 python -m pytest tests/test_gpu_model.py tests/test_gpu_parallel.py tests/test_gpu_ops.py -q -m gpu 2>&1 | tail -8 > gpurun_out/t_all.log
timeout 900 python bench.py --no-cpu-baseline --no-finetune > gpurun_out/md_bench.json 2> gpurun_out/md_bench.err
echo "rc=$?" >> gpurun_out/md_bench.err
