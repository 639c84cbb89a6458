# One gpurun call (1 GPU): smoke, bench (both arms), ncu launch list, ncu captures, other workloads.
# Outputs land in gpurun_out/ (<= 64 MiB in total or nothing comes back: the raw pages are exported to CSV on the box
# and only the K5 report, which carries the source page, is kept); summaries for the record are copied to profiles/.
TAG=${1:-r02_v25}
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest_gpu.log); tail -2 gpurun_out/${TAG}_pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1
timeout 400 python bench.py > gpurun_out/${TAG}_bench_smoke10m.json 2> gpurun_out/${TAG}_bench.err
timeout 400 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${TAG}_launches_smoke10m.csv python bench.py --no-cpu --no-host-rays --steps 1 --warmup 1 > gpurun_out/${TAG}_ncu_launches.log 2>&1
# (a) K5 kernels: full set + source (the timed step's instances are the last two)
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_tile_walk|k_leaf_tile" -c 4 -o gpurun_out/${TAG}_k5 -f python bench.py --no-cpu --no-host-rays --no-stage --steps 1 --warmup 1 > gpurun_out/${TAG}_ncu_k5.log 2>&1
ncu -i gpurun_out/${TAG}_k5.ncu-rep --page raw --csv > gpurun_out/${TAG}_k5_raw.csv 2>/dev/null
# (b) every kernel of the step: full set, no source; only the CSV travels (skipped with QUICK=1)
[ "$QUICK" = "1" ] || timeout 900 ncu --set full --clock-control none -k "regex:k_tile_walk|k_leaf_tile|k_tau_tile|k_trace_flat|k_compact_push|k_os_scatter|k_os_hist|k_gather_pos|k_karras|k_refit_direct|k_refit_climb|k_emit_fast|k_top_cut|k_morton" -c 44 -o /tmp/${TAG}_step -f python bench.py --no-cpu --no-host-rays --no-stage --steps 1 --warmup 1 > gpurun_out/${TAG}_ncu_full.log 2>&1
[ "$QUICK" = "1" ] || ncu -i /tmp/${TAG}_step.ncu-rep --page raw --csv > gpurun_out/${TAG}_step_raw.csv 2>/dev/null
timeout 300 python bench.py --workload fog1m_1080p --no-cpu > gpurun_out/${TAG}_bench_fog1m_1080p.json 2>/dev/null
timeout 600 python bench.py --workload smoke50m_4k --no-cpu --steps 5 > gpurun_out/${TAG}_bench_smoke50m_4k.json 2>/dev/null
[ "$QUICK" = "1" ] || timeout 300 python bench.py --no-cpu --no-tile > gpurun_out/${TAG}_bench_smoke10m_packets.json 2>/dev/null
tail -1 gpurun_out/${TAG}_smoke.log; tail -2 gpurun_out/${TAG}_ncu_full.log; du -sh gpurun_out
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${TAG}_bench_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        if d.get("impl")=="reference": print(f, d["value"], d["ms_per_step"]); continue
        print(f, round(d["ms_per_step"],3), round(d["value"]/1e6,1), "e2e", round(d["e2e"]["value"]/1e6,1), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()}, "frac", round(d["roofline"]["frac"],3))
    except Exception as e: print(f, "ERR", e)
PY
