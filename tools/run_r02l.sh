cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_tridiag.py -x -q -k "reference_layout" 2>&1 | tail -3
EXP_CM=1 EXP_NRHS=10 python tools/exp_small.py thomas 128 786432 0,1,0 0,2,0 0,4,0 2>&1 | grep -v floor
EXP_CM=1 EXP_NRHS=10 python tools/exp_small.py bfb 128 786432 0,1,0 0,2,0 0,4,0 2>&1 | grep -v floor
