set -x
timeout 700 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
mkdir -p gpurun_out/r1f
python bench.py > gpurun_out/r1f/bench_default.json 2> gpurun_out/r1f/bench_default.err
python bench.py --impl reference > gpurun_out/r1f/bench_reference.json 2>/dev/null
for integ in rk101 rk323 rk435; do python bench.py --integrator $integ --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/r1f/bench_c4_$integ.json 2>/dev/null; done
python bench.py --size 4096 --nz 1 --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/r1f/bench_c2_box.json 2>/dev/null
python bench.py --workload c2disc --size 4096 --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/r1f/bench_c2disc.json 2>/dev/null
python bench.py --workload c3 --size 8192 --integrator rk323 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r1f/bench_c3_fixed.json 2>/dev/null
python bench.py --workload c3 --size 8192 --integrator rk323 --adaptive --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r1f/bench_c3_adaptive.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1f/launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r1f/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:smeb_stage_tile -s 6 -c 4 -o gpurun_out/r1f/tile python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r1f/ncu_full.log 2>&1
ls -la gpurun_out/r1f
for f in gpurun_out/r1f/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d.get("roofline") or {}
    print(sys.argv[1].split('/')[-1], round(d["ms_per_step"],4), "%.4g"%d["value"], r.get("frac"), (d.get("e2e") or {}).get("value"), d["clocks"])
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
