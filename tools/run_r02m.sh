cd $GRAFT_REPO_ROOT
for c in cm10_thomas; do
PROF_NCOL=196608 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_tridiag_cm' -s 2 -c 1 -f -o gpurun_out/r02m_full_$c python tools/prof_case.py $c > gpurun_out/ncu_$c.log 2>&1
tail -2 gpurun_out/ncu_$c.log
done
