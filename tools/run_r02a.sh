# round 2, GPU call A: full GPU suite, bench (ours + reference arm), sanitizer families, config-1/2 ncu captures
set -x
cd $GRAFT_REPO_ROOT
TAG=r02a
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${TAG}_pytest_gpu.log
cat gpurun_out/${TAG}_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
tail -3 gpurun_out/${TAG}_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>> gpurun_out/${TAG}_bench.err
for tool in memcheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_cases.py > gpurun_out/${TAG}_sanitizer_${tool}.log 2>&1
  tail -5 gpurun_out/${TAG}_sanitizer_${tool}.log
done
for c in thomas bfb; do
  PROF_NCOL=21600 PROF_NLEV=72 timeout 300 ncu --set full --clock-control none --import-source on -k regex:'k_tridiag' -s 2 -c 1 -f -o gpurun_out/${TAG}_full_c1_$c python tools/prof_case.py $c > gpurun_out/ncu_c1_$c.log 2>&1
done
PROF_NCOL=21600 timeout 300 ncu --set full --clock-control none --import-source on -k regex:'k_lin_interp' -s 2 -c 1 -f -o gpurun_out/${TAG}_full_c2_li72 python tools/prof_case.py li72 > gpurun_out/ncu_c2_li72.log 2>&1
ls -la gpurun_out | tail -12
