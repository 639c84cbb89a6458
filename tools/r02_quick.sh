# usage: bash tools/r02_quick.sh TAG  (one GPU): tile + golden tests, then the bench line of the default workload
TAG=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tile.py tests/test_gpu_golden.py -x -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -3 gpurun_out/${TAG}_pytest.log
timeout 300 python bench.py --no-cpu > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; tail -2 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/${TAG}_bench.json").read().strip().splitlines() if l.startswith("{")][-1])
print(round(d["ms_per_step"],3), round(d["value"]/1e6,1), "e2e", round(d["e2e"]["value"]/1e6,1), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()}, d["roofline"]["frac"])
PY
