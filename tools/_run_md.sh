cd /root/repo
timeout 900 python bench.py --no-cpu-baseline --steps 4 > gpurun_out/md_bench.json 2> gpurun_out/md_bench.err
echo "rc=$?" >> gpurun_out/md_bench.err
