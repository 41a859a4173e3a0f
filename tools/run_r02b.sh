set -x
cd $GRAFT_REPO_ROOT
python tools/exp_small.py thomas 72 32,4736,21600 0,0,0 0,8,0 0,4,0 0,2,0 0,8,8 0,4,8 0,3,8 0,2,8 > gpurun_out/r02b_exp_small.log 2>&1
python tools/exp_small.py bfb 72 32,4736,21600 0,0,0 0,0,5 0,0,11 0,0,6 0,0,4 2,0,0 4,0,0 0,0,9 >> gpurun_out/r02b_exp_small.log 2>&1
python tools/exp_small.py cr 72 21600 >> gpurun_out/r02b_exp_small.log 2>&1
EXP_COLD=1 python tools/exp_small.py thomas 72 21600 0,0,0 0,8,0 >> gpurun_out/r02b_exp_small.log 2>&1
cat gpurun_out/r02b_exp_small.log
python tools/sanitize_cases.py > gpurun_out/r02b_cases.log 2>&1; tail -3 gpurun_out/r02b_cases.log
