mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02_pytest_c10.log 2>&1; echo "exit $?" >> gpurun_out/r02_pytest_c10.log)
tail -4 gpurun_out/r02_pytest_c10.log
timeout 300 python bench.py --no-cpu --steps 10 > gpurun_out/r02_bench_c10_stage.json 2> gpurun_out/r02_bench_c10_stage.err; tail -2 gpurun_out/r02_bench_c10_stage.err
timeout 300 python bench.py --no-cpu --steps 10 --no-stage > gpurun_out/r02_bench_c10_nostage.json 2> /dev/null
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02_bench_c10_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["ms_per_step"],3), round(d["value"]/1e6,1), "e2e", round(d["e2e"]["ms_per_step"],3), "rays", round(d["e2e_host_rays"]["ms_per_step"],3), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()})
    except Exception as e: print(f, "ERR", e)
PY
