cd $GRAFT_REPO_ROOT
python tools/exp_small.py thomas 72 32,4736,21600 0,0,10 4,9,10 2,9,10 4,5,10 2,5,10 4,9,12 2,9,12 > gpurun_out/r02c_exp_small.log 2>&1
python tools/exp_small.py bfb 72 32,4736,21600 0,0,4 4,9,4 2,9,4 >> gpurun_out/r02c_exp_small.log 2>&1
cat gpurun_out/r02c_exp_small.log
