cd $GRAFT_REPO_ROOT
EXP_CM=1 EXP_NRHS=10 python tools/exp_small.py thomas 128 786432 0,0,0 3,0,0 4,0,0 5,0,0 6,0,0 2>&1 | grep -v floor
EXP_CM=1 EXP_NRHS=10 python tools/exp_small.py bfb 128 786432 0,0,0 4,0,0 5,0,0 2>&1 | grep -v floor
EXP_CM=1 EXP_NRHS=4 python tools/exp_small.py thomas 128 345600 0,0,0 7,0,0 10,0,0 2>&1 | grep -v floor
