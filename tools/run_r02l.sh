cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_tridiag.py tests/test_gpu_misc.py -x -q -k "reference_layout or clobber or parity_reference or fused" 2>&1 | tail -5
EXP_CM=1 EXP_NRHS=10 python tools/exp_small.py thomas 128 786432 0,0,0 4,0,0 2>&1 | grep -v floor
EXP_CM=1 EXP_NRHS=43 python tools/exp_small.py thomas 128 2048 0,0,0 1,0,0 2>&1 | grep -v floor
EXP_CM=1 EXP_NRHS=4 python tools/exp_small.py thomas 128 345600 0,0,0 2>&1 | grep -v floor
EXP_CM=1 EXP_NRHS=10 python tools/exp_small.py thomas 72 345600 0,0,0 2>&1 | grep -v floor
EXP_CM=1 EXP_NRHS=10 EXP_DT=f32 python tools/exp_small.py thomas 128 786432 0,0,0 2>&1 | grep -v floor
