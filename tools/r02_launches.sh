mkdir -p gpurun_out
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r02_launches.csv python bench.py --no-cpu --no-host-rays --steps 1 --warmup 1 > gpurun_out/r02_launches.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/r02_launches.csv")) if len(r)>10]
hdr=rows[0]; ik=hdr.index("Kernel Name"); iv=hdr.index("Metric Value")
seq=[(r[ik][:60], float(r[iv].replace(",",""))) for r in rows[1:]]
# last step only: find last k_trace_flat
last=max(i for i,(k,v) in enumerate(seq) if "k_trace_flat" in k)
# the one before (full step between the two k_trace_flat launches)
prev=max(i for i,(k,v) in enumerate(seq[:last]) if "k_trace_flat" in k) if any("k_trace_flat" in k for k,v in seq[:last]) else 0
for k,v in seq[prev:last]: print("%9.1f us  %s"%(v/1000 if v>1e4 else v, k))
PY
