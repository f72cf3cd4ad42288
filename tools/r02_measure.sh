mkdir -p gpurun_out
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_reference.json 2> gpurun_out/r02_bench_reference.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02_ncu_launch.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:hist_smem -s 12 -c 6 -o gpurun_out/r02_prof_hist python bench.py --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/r02_ncu_full.log 2>&1
timeout 200 python tools/hist_bench.py --reps 3 > gpurun_out/r02_hist_bench.txt 2>&1
timeout 200 python tools/col_cost.py > gpurun_out/r02_col_cost.txt 2>&1
timeout 300 python bench.py --config c3 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c3.json 2> gpurun_out/r02_bench_c3.err
timeout 200 python bench.py --config c5a --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c5a.json 2> gpurun_out/r02_bench_c5a.err
timeout 200 python bench.py --config c1 --steps 10 --warmup 3 > gpurun_out/r02_bench_c1.json 2> gpurun_out/r02_bench_c1.err
timeout 500 python bench.py --config c5b --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c5b.json 2> gpurun_out/r02_bench_c5b.err
timeout 500 python bench.py --config c4 --rows 25000000 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c4_25M.json 2> gpurun_out/r02_bench_c4_25M.err
for f in gpurun_out/r02_bench_*.json; do echo $f; tail -c 300 $f; echo; done
tail -3 gpurun_out/r02_ncu_full.log
