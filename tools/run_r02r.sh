cd $GRAFT_REPO_ROOT
EXP_NRHS=10 timeout 300 python tools/exp_small.py bfb 128 1572864 0,0,0 0,0,31 0,0,32 0,4,0 0,5,0 0,4,31 2>&1 | grep -v floor
EXP_NRHS=10 timeout 300 python tools/exp_small.py bfb 72 345600 0,0,0 2>&1 | grep -v floor
