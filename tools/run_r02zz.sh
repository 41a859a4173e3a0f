# ncu source-level profile of k_tridiag_bfb_shared (after the plain run of the same program exited 0)
cd $GRAFT_REPO_ROOT
PROF_NCOL=345600 timeout 120 python tools/prof_case.py shared_bfb > gpurun_out/prof_plain.log 2>&1 || exit 1
PROF_NCOL=345600 timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tridiag_bfb_shared -s 2 -c 1 -f -o gpurun_out/r02zz_full_shared_bfb python tools/prof_case.py shared_bfb > gpurun_out/ncu_zz.log 2>&1
tail -1 gpurun_out/ncu_zz.log
ncu -i gpurun_out/r02zz_full_shared_bfb.ncu-rep --page source --print-source cuda,sass --csv > gpurun_out/r02zz_src.csv 2> gpurun_out/ncu_zz2.log
python tools/ncu_lines.py gpurun_out/r02zz_src.csv 0.7 > gpurun_out/r02zz_bfb_shared_lines.txt 2>&1
python tools/ncu_summary.py gpurun_out/r02zz_full_shared_bfb.ncu-rep > gpurun_out/r02zz_bfb_shared_summary.txt 2>&1
head -60 gpurun_out/r02zz_bfb_shared_lines.txt
