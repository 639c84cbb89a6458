# usage: bash tools/r02_multi.sh N   (inside a gpurun --gpus N call)
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = "2" ]; then
  (timeout 300 python -m pytest tests/test_gpu_multi.py -q > gpurun_out/r02_pytest_multi_n2.log 2>&1; echo "exit $?" >> gpurun_out/r02_pytest_multi_n2.log)
  tail -15 gpurun_out/r02_pytest_multi_n2.log
fi
timeout 150 $TR --master-port 29541 bench.py --gpus $N --no-cpu > gpurun_out/r02_bench_n${N}_strong.json 2> gpurun_out/r02_bench_n${N}_strong.err
tail -3 gpurun_out/r02_bench_n${N}_strong.err
timeout 150 $TR --master-port 29542 bench.py --gpus $N --no-cpu --scaling weak > gpurun_out/r02_bench_n${N}_weak.json 2> gpurun_out/r02_bench_n${N}_weak.err
timeout 200 $TR --master-port 29543 bench.py --gpus $N --no-cpu --steps 5 --workload smoke50m_4k > gpurun_out/r02_bench_n${N}_smoke50m_4k.json 2> gpurun_out/r02_bench_n${N}_smoke50m_4k.err
tail -3 gpurun_out/r02_bench_n${N}_smoke50m_4k.err
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02_bench_n${N}_*.json")):
    try:
        d=json.loads([l for l in open(f).read().strip().splitlines() if l.startswith("{")][-1]); print(f, d["n_gpus"], d["scaling"], round(d["ms_per_step"],3), round(d["value"]/1e6,1), "e2e", round(d["e2e"]["value"]/1e6,1), round(d["e2e"]["ms_per_step"],2), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()})
    except Exception as e: print(f,"ERR",e)
PY
