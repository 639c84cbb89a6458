mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu_c8.log 2>&1; echo "exit $?" >> gpurun_out/r02_pytest_gpu_c8.log)
tail -6 gpurun_out/r02_pytest_gpu_c8.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 400 python bench.py > gpurun_out/r02_bench_c8.json 2> gpurun_out/r02_bench_c8.err; tail -3 gpurun_out/r02_bench_c8.err
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_ref_c8.json 2> gpurun_out/r02_bench_ref_c8.err; tail -3 gpurun_out/r02_bench_ref_c8.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r02_bench_c8.json").read().strip().splitlines()[-1])
print(round(d["ms_per_step"],3), round(d["value"]/1e6,1), "e2e", round(d["e2e"]["value"]/1e6,1), round(d["e2e"]["ms_per_step"],3), "hostrays", d.get("e2e_host_rays",{}).get("ms_per_step"), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()}, "launches", d["gpu_launches_per_step"], d["gpu_launches_per_step_all"], "roofline", d["roofline"]["frac"], d["roofline"]["traffic"])
print(d.get("cpu_baseline"))
r=json.loads(open("gpurun_out/r02_bench_ref_c8.json").read().strip().splitlines()[-1])
print("REF", r["value"], r["ms_per_step"], r["cpu_baseline"]["sample"][:400]); print(r["cpu_baseline"].get("port"))
PY
