# round 2, final GPU call: GPU suite, smoke, ncu --set full of the two timed kernels -> profiles/traffic.json, bench (ours +
# reference arm), launch list. Every ncu command runs only after the same program has exited 0 without ncu.
set -x
cd $GRAFT_REPO_ROOT
TAG=r02zf
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print(\"smoke ok\")" > gpurun_out/${TAG}_smoke.txt 2>&1; tail -2 gpurun_out/${TAG}_smoke.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; tail -3 gpurun_out/${TAG}_pytest_gpu.log
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-extras --no-strong"
timeout 600 $CMD > gpurun_out/${TAG}_bench_short.json 2> gpurun_out/${TAG}_bench_short.err && \
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:k_tridiag_tma_shared|k_lin_interp' -s 6 -c 2 -f -o gpurun_out/${TAG}_full_bench $CMD > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_full.log
python tools/make_traffic.py gpurun_out/${TAG}_full_bench.ncu-rep "profiles/${TAG}_ncu_full_bench_summary.txt (ncu --set full --clock-control none, $CMD, launches 7 and 8 of the two kernels)" > gpurun_out/${TAG}_traffic.json 2> gpurun_out/${TAG}_traffic.err
cp profiles/traffic.json gpurun_out/${TAG}_traffic_file.json
python tools/ncu_summary.py gpurun_out/${TAG}_full_bench.ncu-rep > gpurun_out/${TAG}_ncu_full_bench_summary.txt 2>&1
timeout 1200 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
tail -3 gpurun_out/${TAG}_bench.err
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2>> gpurun_out/${TAG}_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench.csv $CMD > gpurun_out/ncu1.log 2>&1
ls -la gpurun_out | tail -14
