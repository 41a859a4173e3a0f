# round 2: ncu --set full captures of the kernels added this round (one launch each), after the plain runs above exited 0
cd $GRAFT_REPO_ROOT
TAG=r02x
cap() {  # name, kernel regex, env..., case
  local name=$1 re=$2; shift 2
  env "$@" timeout 600 ncu --set full --clock-control none --import-source on -k regex:$re -s 2 -c 1 -f -o gpurun_out/${TAG}_full_$name python tools/prof_case.py ${CASE} > gpurun_out/ncu_$name.log 2>&1
  tail -1 gpurun_out/ncu_$name.log
}
CASE=shared_bfb cap shared_bfb k_tridiag_bfb_shared PROF_NCOL=345600
CASE=cm10_thomas cap cm10_thomas k_tridiag_cm PROF_NCOL=196608
CASE=cm10_bfb cap cm10_bfb k_tridiag_cm PROF_NCOL=196608
CASE=col cap col_cr k_tridiag_cr_warp PROF_NCOL=345600
CASE=cr cap cr_long k_tridiag_cr_cta PROF_NCOL=2048 PROF_NLEV=2048
CASE=fused cap fused k_solve_remap PROF_NCOL=393216
ls -la gpurun_out/*.ncu-rep | tail -8
