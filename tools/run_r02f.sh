cd $GRAFT_REPO_ROOT
python tools/exp_small.py thomas 72 32,4736,21600 0,0,13 0,8,0 0,0,10 > gpurun_out/r02f.log 2>&1
EXP_DT=f32 python tools/exp_small.py thomas 72 21600 0,0,13 0,0,10 >> gpurun_out/r02f.log 2>&1
python tools/exp_small.py bfb 72 32,21600 0,0,0 0,0,4 >> gpurun_out/r02f.log 2>&1
cat gpurun_out/r02f.log
