cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_dataio.py tests/test_gpu_stats.py -q -m gpu -x 2>&1 | tail -12 > gpurun_out/t_io.log
timeout 600 python bench.py > gpurun_out/r_bench.json 2> gpurun_out/r_bench.err
