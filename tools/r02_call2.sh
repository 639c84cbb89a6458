mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_tile.py -q > gpurun_out/r02_tile_tests.log 2>&1; echo "exit $?" >> gpurun_out/r02_tile_tests.log)
tail -5 gpurun_out/r02_tile_tests.log
timeout 300 python bench.py --no-cpu --steps 5 > gpurun_out/r02_bench_c2_tile.json 2> gpurun_out/r02_bench_c2_tile.err
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_traverse_tile|k_leaf_tile|k_tau_table" -s 6 -c 3 -o gpurun_out/r02_c2_tile -f python bench.py --no-cpu --steps 1 --warmup 1 > gpurun_out/r02_ncu_c2.log 2>&1
tail -3 gpurun_out/r02_ncu_c2.log
ls -la gpurun_out/*.ncu-rep
