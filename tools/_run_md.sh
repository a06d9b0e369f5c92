cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_ops.py -q -m gpu -x -k "linear" 2>&1 | tail -5 > gpurun_out/t_all.log
timeout 200 python tools/lin_time.py > gpurun_out/lin_time.log 2>&1
timeout 200 python tools/lin_err.py > gpurun_out/lin_err.log 2>&1
