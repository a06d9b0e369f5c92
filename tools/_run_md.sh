cd /root/repo
timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -5 > gpurun_out/t_all.log
timeout 600 python bench.py --no-finetune > gpurun_out/b_md.json 2> gpurun_out/b_md.err
timeout 600 python bench.py --workload finetune > gpurun_out/b_ft.json 2> gpurun_out/b_ft.err
