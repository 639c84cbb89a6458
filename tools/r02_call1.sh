# r02 call 1: tile path parity + first timing
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_tile.py -x -q > gpurun_out/r02_tile_tests.log 2>&1; echo "exit $?" >> gpurun_out/r02_tile_tests.log)
tail -25 gpurun_out/r02_tile_tests.log
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu_c1.log 2>&1; echo "exit $?" >> gpurun_out/r02_pytest_gpu_c1.log)
tail -5 gpurun_out/r02_pytest_gpu_c1.log
timeout 300 python bench.py --no-cpu > gpurun_out/r02_bench_c1_tile.json 2> gpurun_out/r02_bench_c1_tile.err
timeout 300 python bench.py --no-cpu --no-tile > gpurun_out/r02_bench_c1_notile.json 2> gpurun_out/r02_bench_c1_notile.err
tail -3 gpurun_out/r02_bench_c1_tile.err
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02_bench_c1_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["ms_per_step"],3), round(d["value"]/1e6,1), {k:round(v,3) for k,v in d["config"]["phase_ms"].items()}, d["config"]["leaf_pairs_per_ray"], d["config"]["fast_nodes_per_ray"], d["config"]["fast_disk_tests_per_ray"])
    except Exception as e: print(f, "ERR", e)
PY
