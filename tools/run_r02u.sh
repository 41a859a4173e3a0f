cd $GRAFT_REPO_ROOT
EXP_CM=1 python tools/exp_small.py thomas 72 21600 0,0,0 0,0,13 16,0,13 8,0,13 2>&1 | grep -v floor
EXP_CM=1 python tools/exp_small.py thomas 128 345600 0,0,0 0,0,13 16,0,13 2>&1 | grep -v floor
EXP_CM=1 python tools/exp_small.py bfb 72 21600 0,0,0 0,0,13 16,0,13 2>&1 | grep -v floor
