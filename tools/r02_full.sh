# usage: bash tools/r02_full.sh TAG  (one GPU): whole GPU suite, then bench lines (default workload; packet path)
TAG=$1
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -3 gpurun_out/${TAG}_pytest.log
timeout 300 python bench.py --no-cpu > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; tail -2 gpurun_out/${TAG}_bench.err
timeout 300 python bench.py --no-cpu --no-tile --steps 5 > gpurun_out/${TAG}_bench_notile.json 2> gpurun_out/${TAG}_bench_notile.err
python - <<PY
import json
for f in ("gpurun_out/${TAG}_bench.json","gpurun_out/${TAG}_bench_notile.json"):
    d=json.loads([l for l in open(f).read().strip().splitlines() if l.startswith("{")][-1])
    print(round(d["ms_per_step"],3), round(d["value"]/1e6,1), "e2e", round(d["e2e"]["value"]/1e6,1), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()}, d["roofline"]["frac"])
PY
