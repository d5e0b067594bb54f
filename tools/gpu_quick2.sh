#!/bin/bash
# Quick GPU iteration (round 2): raster parity tests + a short bench (selected legs) with per-kernel times printed.
# usage (under gpurun): bash tools/gpu_quick2.sh <tag> "<pytest args>" "<bench legs>"
tag=${1:-q}; tests=${2:-"tests/test_raster3_gpu.py tests/test_rasters_gpu.py"}; legs=${3:-"surfel,pings_like,large_map"}
mkdir -p gpurun_out/quick
timeout 900 python -m pytest $tests -m gpu -x -q 2>&1 | tail -15
timeout 400 python bench.py --no-cpu-baseline --legs "$legs" > gpurun_out/quick/$tag.json 2> gpurun_out/quick/$tag.err || tail -20 gpurun_out/quick/$tag.err
python - "$tag" <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/quick/{sys.argv[1]}.json").read().strip().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], "serial", d["e2e"].get("serial_value"))
print(" ".join(f"{k}={v['ms']}" for k, v in d["roofline"]["kernels"].items()))
x = d.get("extra", {})
for k, v in x.items():
    if isinstance(v, dict):
        print(k, v.get("value"), v.get("ms_per_step", v.get("ms_per_view")), v.get("phases_ms") or v.get("kernels_ms") or
              {kk: vv["ms"] for kk, vv in v.get("roofline", {}).get("kernels", {}).items()} or "")
    else:
        print(k, v)
PY
