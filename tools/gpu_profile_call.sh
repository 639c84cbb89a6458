# One gpurun call: GPU tests, smoke, bench (both arms), K5a block sweep, ncu launch list and full capture, other workloads.
# Outputs land in gpurun_out/; the summaries kept for the record are copied to profiles/ (r01_*_v15*).
mkdir -p gpurun_out
run() { timeout 200 python bench.py --no-cpu > gpurun_out/ab_$1.json 2>/dev/null; }
(timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_v15.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu_v15.log)
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_v15.log 2>&1
timeout 300 python bench.py > gpurun_out/bench_v15.json 2> gpurun_out/bench_v15.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_v15.json 2> gpurun_out/bench_ref_v15.err
run w_base1; for k in 5 6 7 8; do BRE_K5A_BLOCKS=$k run w_k$k; done; run w_base2
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v15.csv python bench.py --no-cpu --steps 1 --warmup 1 > gpurun_out/ncu_v15.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k "regex:k_traverse|k_leaf_fat|k_leaf_pairs|k_interleave|k_gather_aos|k_karras|k_refit_direct|k_refit_climb|k_emit_fast|k_rs_scatter|k_trace_flat|k_tau_table|k_compact" -s 18 -c 16 -o gpurun_out/step_v15 -f python bench.py --no-cpu --steps 1 --warmup 1 > gpurun_out/ncu_full_v15.log 2>&1
timeout 300 python bench.py --workload fog1m_1080p --no-cpu > gpurun_out/bench_v15_fog1m.json 2>/dev/null
timeout 600 python bench.py --workload smoke50m_4k --no-cpu --steps 5 > gpurun_out/bench_v15_smoke50m_4k.json 2>/dev/null
tail -2 gpurun_out/pytest_gpu_v15.log; tail -1 gpurun_out/smoke_v15.log
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/ab_w_*.json")+glob.glob("gpurun_out/bench_v15*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["ms_per_step"],3), round(d["value"]/1e6,1), round(d["e2e"]["value"]/1e6,1), {k:round(v,3) for k,v in d["config"]["phase_ms"].items()})
    except Exception as e: print(f, "ERR", e)
PY
