set -x
python bench.py --workload cfg3 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/s24_cfg3_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r1_s24_cfg3_launches.csv python bench.py --workload cfg3 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/s24_cfg3_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"conv_out_fwd_mma|pixel_shuffle2|conv_in_fwd" -c 4 -o gpurun_out/s24_cfg3_thin python bench.py --workload cfg3 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/s24_cfg3_ncu_full.log 2>&1
tail -2 gpurun_out/s24_cfg3_plain.log | cut -c1-200
ls -la gpurun_out/ | tail -5
