cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_tridiag.py -x -q 2>&1 | tail -5 > gpurun_out/r02e_pytest.log; cat gpurun_out/r02e_pytest.log
python tools/exp_small.py thomas 72 32,4736,21600,345600 0,0,13 0,8,0 0,0,10 > gpurun_out/r02e.log 2>&1
python tools/exp_small.py thomas 128 21600,345600 0,0,13 0,8,0 >> gpurun_out/r02e.log 2>&1
EXP_DT=f32 python tools/exp_small.py thomas 72 21600 0,0,13 0,0,10 >> gpurun_out/r02e.log 2>&1
EXP_DT=f32 python tools/exp_small.py thomas 128 21600,345600 0,0,13 0,0,10 >> gpurun_out/r02e.log 2>&1
python tools/exp_small.py bfb 72 21600 0,0,0 >> gpurun_out/r02e.log 2>&1
cat gpurun_out/r02e.log
