# One gpurun call (1 GPU): GPU tests, smoke, bench (both arms), ncu launch list and full capture, other workloads.
# Outputs land in gpurun_out/; the summaries kept for the record are copied to profiles/ (r02_*).
TAG=${1:-r02_v20}
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log)
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1
timeout 400 python bench.py > gpurun_out/${TAG}_bench_smoke10m.json 2> gpurun_out/${TAG}_bench.err
timeout 400 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_launches_smoke10m.csv python bench.py --no-cpu --no-host-rays --steps 1 --warmup 1 > gpurun_out/${TAG}_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_tile_walk|k_leaf_tile|k_tau_tile|k_trace_flat|k_compact_push|k_os_scatter|k_gather_aos|k_interleave|k_karras|k_refit_direct|k_refit_climb|k_emit_fast" -s 17 -c 15 -o gpurun_out/${TAG}_step -f python bench.py --no-cpu --no-host-rays --no-stage --steps 1 --warmup 1 > gpurun_out/${TAG}_ncu_full.log 2>&1
timeout 300 python bench.py --workload fog1m_1080p --no-cpu > gpurun_out/${TAG}_bench_fog1m_1080p.json 2>/dev/null
timeout 600 python bench.py --workload smoke50m_4k --no-cpu --steps 5 > gpurun_out/${TAG}_bench_smoke50m_4k.json 2>/dev/null
timeout 300 python bench.py --no-cpu --no-tile > gpurun_out/${TAG}_bench_smoke10m_packets.json 2>/dev/null
tail -2 gpurun_out/${TAG}_pytest_gpu.log; tail -1 gpurun_out/${TAG}_smoke.log; tail -2 gpurun_out/${TAG}_ncu_full.log
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${TAG}_bench_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        if d.get("impl")=="reference": print(f, d["value"], d["ms_per_step"]); continue
        print(f, round(d["ms_per_step"],3), round(d["value"]/1e6,1), "e2e", round(d["e2e"]["value"]/1e6,1), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()}, "frac", round(d["roofline"]["frac"],3))
    except Exception as e: print(f, "ERR", e)
PY
