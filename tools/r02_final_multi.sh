# usage: bash tools/r02_final_multi.sh N  (inside gpurun --gpus N): final-build multi-GPU lines (photon sharding),
# K6 start policy A/B (BRE_TAU_DEFER), and on 2 GPUs the multi-GPU tests
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = "2" ]; then
  (timeout 500 python -m pytest tests/test_gpu_multi.py -q > gpurun_out/r02_final_pytest_multi_n2.log 2>&1; echo "exit $?" >> gpurun_out/r02_final_pytest_multi_n2.log)
  tail -4 gpurun_out/r02_final_pytest_multi_n2.log
fi
for D in ${DEFERS:-1 0}; do
  BRE_TAU_DEFER=$D timeout 150 $TR --master-port 2955$D bench.py --gpus $N --no-cpu > gpurun_out/r02_final_n${N}_pshard_defer$D.json 2> gpurun_out/r02_final_n${N}_pshard_defer$D.err
  BRE_TAU_DEFER=$D timeout 200 $TR --master-port 2956$D bench.py --gpus $N --no-cpu --steps 5 --workload smoke50m_4k > gpurun_out/r02_final_n${N}_pshard_smoke50m_4k_defer$D.json 2> gpurun_out/r02_final_n${N}_pshard_smoke50m_4k_defer$D.err
done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r02_final_n${N}_*.json")):
    try:
        d=json.loads([l for l in open(f).read().strip().splitlines() if l.startswith("{")][-1]); print(f, d["n_gpus"], round(d["ms_per_step"],3), round(d["value"]/1e6,1), "e2e", round(d["e2e"]["value"]/1e6,1), round(d["e2e"]["ms_per_step"],2), {k:round(v,3) for k,v in d["detail"]["phase_ms"].items()})
    except Exception as e: print(f,"ERR",e)
PY
