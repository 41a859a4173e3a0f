# round 2: the bench contract at N = 4 and N = 8 (weak + strong legs, e2e) on one box
cd $GRAFT_REPO_ROOT
for N in 4 8; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29530+N)) bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02y_bench_n$N.json 2> gpurun_out/r02y_bench_n$N.err
  tail -2 gpurun_out/r02y_bench_n$N.err
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > gpurun_out/r02y_ref_n8.json 2> gpurun_out/r02y_ref_n8.err
tail -c 600 gpurun_out/r02y_ref_n8.json
