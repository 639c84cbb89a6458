# usage: bash tools/gpu_scale_call.sh N   (inside a gpurun --gpus N call)
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29531 bench.py --gpus $N --no-cpu > gpurun_out/bench_v15_n${N}_p2p.json 2> gpurun_out/bench_v15_n${N}_p2p.err
timeout 200 $TR --master-port 29532 bench.py --gpus $N --no-cpu --exchange nccl > gpurun_out/bench_v15_n${N}_nccl.json 2> gpurun_out/bench_v15_n${N}_nccl.err
timeout 300 $TR --master-port 29533 bench.py --gpus $N --no-cpu --steps 5 --workload smoke50m_4k > gpurun_out/bench_v15_n${N}_smoke50m_4k.json 2> gpurun_out/bench_v15_n${N}_smoke50m_4k.err
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/bench_v15_n${N}_*.json")):
    try:
        d=json.loads([l for l in open(f).read().strip().splitlines() if l.startswith("{")][-1]); print(f, d["n_gpus"], round(d["ms_per_step"],3), round(d["value"]/1e6,1), round(d["e2e"]["value"]/1e6,1), round(d["e2e"]["ms_per_step"],2), {k:round(v,3) for k,v in d["config"]["phase_ms"].items()})
    except Exception as e: print(f,"ERR",e)
PY
