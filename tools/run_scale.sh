set -x
cd $GRAFT_REPO_ROOT
TAG=${TAG:-r01g}
: > gpurun_out/${TAG}_scale.jsonl
for N in 1 2 4 8; do
  if [ $N = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-extras --no-cpu >> gpurun_out/${TAG}_scale.jsonl 2>> gpurun_out/${TAG}_scale.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) bench.py --gpus $N --steps 10 --warmup 3 >> gpurun_out/${TAG}_scale.jsonl 2>> gpurun_out/${TAG}_scale.err
  fi
done
cat gpurun_out/${TAG}_scale.jsonl | cut -c1-200
tail -5 gpurun_out/${TAG}_scale.err
