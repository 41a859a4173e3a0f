cd $GRAFT_REPO_ROOT
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_tridiag_cr' -s 2 -c 1 -f -o gpurun_out/r02k_full_cr python tools/prof_case.py cr > gpurun_out/ncu_cr.log 2>&1
tail -3 gpurun_out/ncu_cr.log
ls -la gpurun_out/*.ncu-rep
