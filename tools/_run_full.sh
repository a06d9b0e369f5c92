cd /root/repo
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/s_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-finetune > gpurun_out/s_ncu.log 2>&1
