cd /root/repo
mkdir -p gpurun_out
N=$1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/md_bench_n$N.json 2> gpurun_out/md_bench_n$N.err
echo "md rc=$?" >> gpurun_out/md_bench_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --workload finetune --steps 5 --warmup 3 > gpurun_out/ft_bench_n$N.json 2> gpurun_out/ft_bench_n$N.err
echo "ft rc=$?" >> gpurun_out/ft_bench_n$N.err
