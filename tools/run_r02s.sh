cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_gpu_misc.py -x -q 2>&1 | tail -12 > gpurun_out/r02s_pytest.log; cat gpurun_out/r02s_pytest.log
timeout 120 python tools/time_fused.py
