#!/bin/bash
# One 8-GPU box: the copy ceiling at N = 2, 4, 8 and the bench at N = 8, 4, 2 (run under gpurun --gpus 8).
mkdir -p gpurun_out
for N in ${PCIE_N:-2 4 8}; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
    tools/pcie_ceiling.py 2>/dev/null | tail -1 > gpurun_out/pcie_$N.json
  cat gpurun_out/pcie_$N.json
done
for N in 8 4 2; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + N)) \
    bench.py --gpus $N --steps 100 --warmup 5 --no-cpu-baseline ${BENCH_FLAGS:---only euroc_sequence,hilti_multiframe} 2>gpurun_out/scale_$N.err | tail -1 > gpurun_out/scale_$N.json
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/scale_$N.json"))
    print($N, d["value"], d["e2e"]["value"], {k: (v.get("value"), v.get("e2e", {}).get("value")) for k, v in d["extra"]["configs"].items() if isinstance(v, dict) and "value" in v})
    sw = d["extra"].get("matcher_sweep_inputs")
    if sw: print({k: v.get("value") for k, v in sw.items()})
except Exception as e:
    print($N, "failed", e); print(open("gpurun_out/scale_$N.err").read()[-1500:])
PY
done
