#!/bin/bash
# run bench.py once per library variant (B200AMG_LIB) and print the level-0 op times: kernel tuning A/B
for lib in "$@"; do
  if [ "$lib" = base ]; then unset B200AMG_LIB; else export B200AMG_LIB=$PWD/$lib; fi
  timeout 200 python bench.py --no-cpu-baseline --steps 20 > gpurun_out/v.json 2> gpurun_out/v.err || { echo "$lib FAILED"; tail -2 gpurun_out/v.err; continue; }
  python - "$lib" <<PY
import json,sys
d=json.loads(open("gpurun_out/v.json").read().strip().splitlines()[-1])
o=d["roofline"]["op_ms"]
print(sys.argv[1], round(d["value"],1), {k:o[k] for k in ("presmooth@L0","residual@L0","restrict@L0","prolong@L0","postsmooth@L0")})
PY
done
