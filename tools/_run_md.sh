cd /root/repo
timeout 900 python -m pytest tests/test_gpu_model.py -q -m gpu -k "native or efs" 2>&1 | tail -12 > gpurun_out/t_all.log
