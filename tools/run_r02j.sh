cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_tridiag.py tests/test_gpu_diffusion.py -x -q -k "cr or diffusion" 2>&1 | tail -15 > gpurun_out/r02j_pytest.log; cat gpurun_out/r02j_pytest.log
python tools/exp_small.py cr 128 345600 0,0,0 0,0,14 > gpurun_out/r02j.log 2>&1
python tools/exp_small.py cr 72 21600,345600 0,0,0 0,0,14 >> gpurun_out/r02j.log 2>&1
EXP_DT=f32 python tools/exp_small.py cr 128 345600 0,0,0 0,0,14 >> gpurun_out/r02j.log 2>&1
cat gpurun_out/r02j.log
