# usage: bash tools/r02_pshard.sh N   (inside a gpurun --gpus N call): photon-sharded render check + bench lines
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29540 tools/check_photon_shard.py > gpurun_out/r02_pshard_check_n${N}.log 2>&1; echo "check exit $?"
grep -E "^rank|Error|error" gpurun_out/r02_pshard_check_n${N}.log | head -20
timeout 150 $TR --master-port 29541 bench.py --gpus $N --no-cpu > gpurun_out/r02_bench_n${N}_pshard.json 2> gpurun_out/r02_bench_n${N}_pshard.err
tail -3 gpurun_out/r02_bench_n${N}_pshard.err
timeout 150 $TR --master-port 29542 bench.py --gpus $N --no-cpu --shard tiles > gpurun_out/r02_bench_n${N}_tshard.json 2> gpurun_out/r02_bench_n${N}_tshard.err
timeout 200 $TR --master-port 29543 bench.py --gpus $N --no-cpu --steps 5 --workload smoke50m_4k > gpurun_out/r02_bench_n${N}_pshard_smoke50m_4k.json 2> gpurun_out/r02_bench_n${N}_pshard_smoke50m_4k.err
tail -3 gpurun_out/r02_bench_n${N}_pshard_smoke50m_4k.err
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02_bench_n${N}_*shard*.json")):
    try:
        d=json.loads([l for l in open(f).read().strip().splitlines() if l.startswith("{")][-1]); print(f, d["n_gpus"], d["scaling"], round(d["ms_per_step"],3), round(d["value"]/1e6,1), "e2e", round(d["e2e"]["value"]/1e6,1), round(d["e2e"]["ms_per_step"],2), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()})
    except Exception as e: print(f,"ERR",e)
PY
