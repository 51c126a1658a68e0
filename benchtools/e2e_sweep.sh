#!/bin/bash
# e2e host-buffer variants of the cfg2 step, one bench.py --quick run each (tuning tool)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_guards.py -x -q -k "host_pipe or host_pipes" 2>&1 | tail -5 > gpurun_out/e2e_pytest.log
i=0
while read -r mode; do
  [ -z "$mode" ] && continue
  i=$((i+1))
  python bench.py --quick --steps 20 --warmup 5 --regions 3 --e2e-mode $mode > gpurun_out/e2e_$i.json 2> gpurun_out/e2e_$i.err
  python - <<PY
import json
try:
    r=json.load(open("gpurun_out/e2e_$i.json")); e=r["e2e"]
    print("$mode", "| ms/step %.4f"%e["ms_per_step"], "parity", e["parity"]["ok"], "value %.3g"%e["value"], "h2d %.1f MB"%(e["h2d_bytes_per_step"]/1e6))
except Exception as ex:
    print("$mode", "FAILED", ex)
PY
done < "${1:-benchtools/e2e_modes.txt}"
python bench.py --quick --e2e-variants --steps 20 --warmup 5 --regions 3 > gpurun_out/e2e_full.json 2> gpurun_out/e2e_full.err; python -c "import json; r=json.load(open(\"gpurun_out/e2e_full.json\")); print(json.dumps(r[\"e2e\"], indent=1))"
