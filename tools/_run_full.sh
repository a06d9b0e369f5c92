cd /root/repo
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/r_bench.json 2> gpurun_out/r_bench.err
