set -x
cd $GRAFT_REPO_ROOT
TAG=${TAG:-r01e}
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/${TAG}_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
cat gpurun_out/${TAG}_pytest_gpu.log
tail -3 gpurun_out/${TAG}_bench.err
