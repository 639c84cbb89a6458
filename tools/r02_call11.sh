mkdir -p gpurun_out
(timeout 1200 python -m pytest tests/test_gpu_golden.py tests/test_gpu_parity.py tests/test_gpu_surfaces.py tests/test_gpu_io.py -q -x > gpurun_out/r02_pytest_c11.log 2>&1; echo "exit $?" >> gpurun_out/r02_pytest_c11.log)
tail -3 gpurun_out/r02_pytest_c11.log
timeout 300 python bench.py --no-cpu --no-host-rays --steps 10 > gpurun_out/r02_bench_c11_tiled.json 2>/dev/null
BRE_COMPACT_TILED=0 timeout 300 python bench.py --no-cpu --no-host-rays --steps 10 > gpurun_out/r02_bench_c11_plain.json 2>/dev/null
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02_bench_c11_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["ms_per_step"],3), round(d["value"]/1e6,1), "e2e", round(d["e2e"]["ms_per_step"],3), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()})
    except Exception as e: print(f, "ERR", e)
PY
