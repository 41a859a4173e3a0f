cd $GRAFT_REPO_ROOT
python tools/exp_small.py thomas 72 32,21600 0,0,0 0,0,10 > gpurun_out/r02d.log 2>&1
PROF_NCOL=32 PROF_NLEV=72 ncu --metrics gpu__time_duration.sum --clock-control none --csv python tools/prof_case.py thomas bfb 2>&1 | grep -i "k_tridiag\|fill\|copy" | cut -c1-200 >> gpurun_out/r02d.log
cat gpurun_out/r02d.log
