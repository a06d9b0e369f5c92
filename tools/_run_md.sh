cd /root/repo
timeout 1200 python -m pytest tests -q -m gpu 2>&1 | tail -15 > gpurun_out/t_all.log
B200_RADIAL_TABLE=0 timeout 600 python -m pytest tests/test_gpu_model.py -q -m gpu 2>&1 | tail -5 > gpurun_out/t_notable.log
