# repeat the 16x16 tile test under env variants to localise a race
mkdir -p gpurun_out
for E in "BRE_LAZY_PAYLOAD=1" "BRE_LAZY_PAYLOAD=0"; do
  f=0
  for i in 1 2 3 4 5 6 7 8; do
    env $E timeout 120 python -m pytest tests/test_gpu_tile.py -x -q -k "image_sizes" > gpurun_out/flaky.log 2>&1 || { f=$((f+1)); grep -E "^FAILED|AssertionError: assert" gpurun_out/flaky.log | head -2; }
  done
  echo "$E failures $f / 8"
done
