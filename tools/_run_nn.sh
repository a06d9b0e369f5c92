cd /root/repo
timeout 900 python -m pytest tests/test_gpu_nn.py -x -q -m gpu 2>&1 | tail -30 > gpurun_out/t_nn.log
