cd $GRAFT_REPO_ROOT
timeout 1200 python -m pytest tests/test_gpu_tridiag.py -x -q 2>&1 | tail -8 > gpurun_out/r02p_pytest.log; cat gpurun_out/r02p_pytest.log
python tools/time_clobber.py
python tools/exp_small.py bfb 128 345600,3145728 0,0,0 2>&1 | grep -v floor
