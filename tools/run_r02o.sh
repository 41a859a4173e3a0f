cd $GRAFT_REPO_ROOT
timeout 1200 python -m pytest tests/test_gpu_tridiag.py -x -q 2>&1 | tail -8 > gpurun_out/r02o_pytest.log; cat gpurun_out/r02o_pytest.log
EXP_CM=1 python tools/exp_small.py cr 128 345600,2048 0,0,0 0,0,2 2>&1 | grep -v floor
EXP_CM=1 EXP_NRHS=43 python tools/exp_small.py cr 128 2048 0,0,0 0,0,2 2>&1 | grep -v floor
EXP_NRHS=43 python tools/exp_small.py cr 128 2048 0,0,0 0,0,2 2>&1 | grep -v floor
python tools/exp_small.py cr 128 2048,4736 0,0,0 0,0,2 2>&1 | grep -v floor
python tools/exp_small.py cr 256 2048 0,0,0 0,0,2 2>&1 | grep -v floor
EXP_CM=1 EXP_DT=f32 python tools/exp_small.py cr 128 345600 0,0,0 0,0,2 2>&1 | grep -v floor
