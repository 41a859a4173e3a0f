# round 2: 2-GPU run of the bench contract (weak + strong legs, e2e with NUMA binding) and the reference arm under torchrun
cd $GRAFT_REPO_ROOT
N=${N:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02t_bench_n$N.json 2> gpurun_out/r02t_bench_n$N.err
tail -3 gpurun_out/r02t_bench_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/r02t_ref_n$N.json 2>> gpurun_out/r02t_bench_n$N.err
cut -c1-600 gpurun_out/r02t_ref_n$N.json
