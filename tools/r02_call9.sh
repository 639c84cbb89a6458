mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_golden.py -q -x > gpurun_out/r02_pytest_c9.log 2>&1; echo "exit $?" >> gpurun_out/r02_pytest_c9.log)
tail -4 gpurun_out/r02_pytest_c9.log
timeout 300 python bench.py --no-cpu --no-host-rays --steps 10 > gpurun_out/r02_bench_c9_os.json 2> gpurun_out/r02_bench_c9_os.err; tail -2 gpurun_out/r02_bench_c9_os.err
BRE_SORT_ONESWEEP=0 timeout 300 python bench.py --no-cpu --no-host-rays --steps 10 > gpurun_out/r02_bench_c9_rs.json 2> gpurun_out/r02_bench_c9_rs.err
timeout 300 python bench.py --no-cpu --no-host-rays --steps 5 --workload smoke50m_4k > gpurun_out/r02_bench_c9_50m.json 2> gpurun_out/r02_bench_c9_50m.err
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02_bench_c9_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["ms_per_step"],3), round(d["value"]/1e6,1), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()})
    except Exception as e: print(f, "ERR", e)
PY
