set -x
python bench.py --no-cpu-baseline --workload cfg1 > gpurun_out/s24_bench_cfg1.log 2>&1
python bench.py --no-cpu-baseline --workload cfg3 > gpurun_out/s24_bench_cfg3.log 2>&1
python bench.py --no-cpu-baseline --workload cfg5 > gpurun_out/s24_bench_cfg5.log 2>&1
python bench.py --no-cpu-baseline --workload cfg4 > gpurun_out/s24_bench_cfg4.log 2>&1
python tools/scratch/cfg1_time.py 1 once && ncu --set full --clock-control none --import-source on -k regex:"conv_(in|out)_fwd" -s 2 -c 2 -o gpurun_out/conv_thin python tools/scratch/cfg1_time.py 1 once > gpurun_out/conv_thin_ncu.log 2>&1
for f in gpurun_out/s24_*.log; do echo "== $f"; tail -1 $f | cut -c1-200; done
