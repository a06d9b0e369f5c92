cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 > gpurun_out/t_all.log
timeout 200 python tools/lin_time.py > gpurun_out/lin_time.log 2>&1
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/s_bench.json 2> gpurun_out/s_bench.err
