set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/s24_pytest_final.log
python bench.py > gpurun_out/s24_bench.log 2>&1
python bench.py --no-cpu-baseline --workload cfg1 > gpurun_out/s24_bench_cfg1.log 2>&1
python bench.py --no-cpu-baseline --workload cfg1 --batch 64 > gpurun_out/s24_bench_cfg1_b64.log 2>&1
python bench.py --no-cpu-baseline --workload cfg3 > gpurun_out/s24_bench_cfg3.log 2>&1
python bench.py --no-cpu-baseline --workload cfg4 > gpurun_out/s24_bench_cfg4.log 2>&1
python bench.py --no-cpu-baseline --workload cfg5 > gpurun_out/s24_bench_cfg5.log 2>&1
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/s24_smoke.log 2>&1
for f in gpurun_out/s24_*.log; do echo "== $f"; tail -1 $f | cut -c1-200; done
