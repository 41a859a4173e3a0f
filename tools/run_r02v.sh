cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_tridiag.py -x -q -k "cr or long or parity" 2>&1 | tail -5
python tools/exp_small.py cr 2048 2048,256 0,0,0 0,0,2 2>&1 | grep -v floor
python tools/exp_small.py cr 512 2048,20000 0,0,0 0,0,2 2>&1 | grep -v floor
python tools/exp_small.py cr 1000 2048 0,0,0 0,0,2 2>&1 | grep -v floor
EXP_CM=1 python tools/exp_small.py cr 2048 2048 0,0,0 0,0,2 2>&1 | grep -v floor
EXP_DT=f32 python tools/exp_small.py cr 2048 2048 0,0,0 0,0,2 2>&1 | grep -v floor
