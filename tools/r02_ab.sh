# usage: bash tools/r02_ab.sh TAG "ENV1" "ENV2" ...  (one GPU): tile tests once, then one bench line per environment setting
TAG=$1; shift
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tile.py tests/test_gpu_golden.py -x -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -2 gpurun_out/${TAG}_pytest.log
i=0
for E in "$@"; do
  i=$((i+1))
  env $E timeout 300 python bench.py --no-cpu --steps 8 > gpurun_out/${TAG}_bench_$i.json 2> gpurun_out/${TAG}_bench_$i.err
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/${TAG}_bench_$i.json").read().strip().splitlines() if l.startswith("{")][-1])
print("$E", round(d["ms_per_step"],3), round(d["value"]/1e6,1), "e2e", round(d["e2e"]["value"]/1e6,1), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()})
PY
done
