cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_nn.py -q -m gpu -x -k "nequip" 2>&1 | tail -25 > gpurun_out/t_nn.log
