rm -f gpurun_out/r06_configs.jsonl gpurun_out/*.ncu-rep
python bench.py > gpurun_out/r06_bench_default.jsonl 2> gpurun_out/r06_bench_default.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r06_bench_reference.jsonl 2>> gpurun_out/r06_bench_default.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r06_launches_default.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
PT_SLICES=1 PT_NO_GRAPH=1 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shade|k_shadow' -c 6 -f -o gpurun_out/prof_r06_atrium python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
PT_FRAME_BLOCKS_PER_SM=9 PT_SLICES=1 PT_NO_GRAPH=1 ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shadow' -c 2 -f -o gpurun_out/prof_r06_atrium_fullgrid python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_trace_batch' -c 2 -f -o gpurun_out/prof_r06_micro python bench.py --workload micro --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu4.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_ploc|k_sort|k_morton|k_scene|k_emit|k_depth' -c 400 --csv --log-file gpurun_out/r06_launches_ploc.csv python bench.py --builder ploc --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu5.log 2>&1
python tools/run_configs.py --out gpurun_out/r06_configs.jsonl
ls -la gpurun_out | tail -20
