#!/bin/bash
# Fill / drain cost of one b200_reads_to_sam call when a device gets only 1/8 of the 100 M reads (the 8-GPU case), and the
# sub-batch size that pays for it: one GPU, 12.5 M reads per call.
mkdir -p gpurun_out
for sr in 0 524288 262144; do
  python bench.py --reads 12500000 --steps 4 --warmup 2 --no-sub --no-cpu-baseline --sub-reads $sr > gpurun_out/sub_$sr.json 2>/dev/null
done
python - <<PY
import json
for f in ("sub_0", "sub_524288", "sub_262144"):
    try:
        r = json.loads(open("gpurun_out/%s.json" % f).read().strip().split("\n")[-1])
        print(json.dumps({"run": f, "value": r["value"], "e2e": r["e2e"]["value"], "sub_batches": r["e2e"].get("sub_batches"), "thread_seconds": r["e2e"]["thread_seconds"]}))
    except Exception as e:
        print(f, "ERR", e)
PY
