set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r01j_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/r01j_bench.json 2> gpurun_out/r01j_bench.err
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01j_bench_reference.json 2>> gpurun_out/r01j_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01j_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-extras > gpurun_out/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_tridiag|k_lin_interp' -s 8 -c 2 -f -o gpurun_out/r01j_full_bench python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-extras > gpurun_out/ncu2.log 2>&1
for c in shared_thomas shared_bfb col_thomas col_bfb bfb cr fused li; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_tridiag|k_solve_remap|k_lin_interp' -s 2 -c 1 -f -o gpurun_out/r01j_full_$c python tools/prof_case.py $c > gpurun_out/ncu_$c.log 2>&1
done
ls -la gpurun_out | tail -20
