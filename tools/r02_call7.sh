mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_tile.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py -q > gpurun_out/r02_tile_tests.log 2>&1; echo "exit $?" >> gpurun_out/r02_tile_tests.log)
tail -5 gpurun_out/r02_tile_tests.log
timeout 300 python bench.py --no-cpu --steps 10 > gpurun_out/r02_bench_c7_tile.json 2> gpurun_out/r02_bench_c7_tile.err
tail -2 gpurun_out/r02_bench_c7_tile.err
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02_bench_c7_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["ms_per_step"],3), round(d["value"]/1e6,1), {k:round(v,3) for k,v in d["config"]["phase_ms"].items()}, d["config"]["leaf_pairs_per_ray"], d["config"]["fast_nodes_per_ray"], d["config"]["fast_disk_tests_per_ray"])
    except Exception as e: print(f, "ERR", e)
PY
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_tau_tile|k_leaf_tile" -s 4 -c 2 -o gpurun_out/r02_c7_tile -f python bench.py --no-cpu --steps 1 --warmup 1 > gpurun_out/r02_ncu_c7.log 2>&1
tail -2 gpurun_out/r02_ncu_c7.log
