cd $GRAFT_REPO_ROOT
timeout 1200 python -m pytest tests/test_gpu_tridiag.py tests/test_gpu_misc.py -x -q 2>&1 | tail -5 > gpurun_out/r02n_pytest.log; cat gpurun_out/r02n_pytest.log
python tools/trace_shared.py 345600 128 10 2>&1 | grep bfb | head -1
EXP_NRHS=10 python tools/exp_small.py bfb 128 345600,1572864 0,0,0 5,4,0 5,6,0 2>&1 | grep -v floor
EXP_NRHS=4 python tools/exp_small.py bfb 128 345600 0,0,0 2>&1 | grep -v floor
EXP_NRHS=10 EXP_DT=f32 python tools/exp_small.py bfb 128 345600 0,0,0 2>&1 | grep -v floor
