#!/bin/bash
# Round-2 measurement batch on ONE B200 (every command under its own timeout; outputs under gpurun_out/, copied to profiles/ by hand).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout=900 > gpurun_out/r02_pytest_gpu.txt 2>&1; tail -3 gpurun_out/r02_pytest_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -2
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_reference.json 2> gpurun_out/r02_bench_reference.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02_ncu_launch.log 2>&1
timeout 300 python bench.py --config c3 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c3.json 2> gpurun_out/r02_bench_c3.err
timeout 200 python bench.py --config c5a --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c5a.json 2> gpurun_out/r02_bench_c5a.err
timeout 200 python bench.py --config c1 --steps 10 --warmup 3 > gpurun_out/r02_bench_c1.json 2> gpurun_out/r02_bench_c1.err
timeout 500 python bench.py --config c5b --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c5b.json 2> gpurun_out/r02_bench_c5b.err
timeout 500 python bench.py --config c4 --rows 25000000 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_c4_25M.json 2> gpurun_out/r02_bench_c4_25M.err
for f in gpurun_out/r02_bench_*.json; do echo $f; python - "$f" <<'PY'
import json, sys
l = [x for x in open(sys.argv[1]) if x.startswith("{")]
if l:
    d = json.loads(l[-1])
    print(d.get("ms_per_step"), d.get("value"), (d.get("e2e") or {}).get("ms_total"), (d.get("roofline") or {}).get("frac"), d.get("parity"))
PY
done
