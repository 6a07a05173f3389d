#!/bin/bash
# where the e2e step's 0.07 ms over the device-timed step goes: drop one leg of the pipeline at a time
set -x
mkdir -p gpurun_out
true
true
rm -f gpurun_out/e2ediag.txt
for skip in none h2d h2d,d2d h2d,d2h h2d,d2d,d2h none; do
  s=$skip; [ $s = none ] && s=""
  EAGS_BENCH_E2E_SKIP=$s timeout 300 python bench.py --no-mapping --no-cpu --steps 200 --warmup 5 2>/dev/null | tail -1 > /tmp/e.json
  python - "$skip" <<'PY' >> gpurun_out/e2ediag.txt
import json, sys
d = json.load(open("/tmp/e.json"))
print("skip", sys.argv[1], "| device ms/step %.4f" % d["ms_per_step"], "| e2e ms/step %.4f" % d["e2e"]["ms_per_step"],
      "| h2d alone %.3f / %.3f" % (d["e2e"]["h2d_copy_alone"]["before"]["median_ms"], d["e2e"]["h2d_copy_alone"]["after"]["median_ms"]),
      "| eager %.4f nosync %.4f" % (d["eager"]["ms_per_step"], d["eager"]["nosync"]["ms_per_step"]))
PY
done
cat gpurun_out/e2ediag.txt
