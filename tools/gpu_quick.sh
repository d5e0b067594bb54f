#!/bin/bash
# Quick GPU iteration: raster parity tests + a short bench with the per-phase times printed.
# usage (under gpurun): bash tools/gpu_quick.sh [tag] [pytest files...]
tag=${1:-q}; shift
mkdir -p gpurun_out/quick
tests=${@:-tests/test_raster3_gpu.py tests/test_rasters_gpu.py tests/test_renderer_gpu.py}
timeout 400 python -m pytest $tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/quick/$tag.json 2> gpurun_out/quick/$tag.err || tail -20 gpurun_out/quick/$tag.err
python - "$tag" <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/quick/{sys.argv[1]}.json").read().strip().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], "serial", d["e2e"].get("serial_value"))
print(" ".join(f"{k}={v['ms']}" for k, v in d["roofline"]["phases"].items()))
x = d["extra"]
print("c3 phases", x["surfel_ssim"].get("phases_ms"))
print("mapping", x.get("mapping_step"))
print("sdf", x["sdf"]["value"], x["sdf"]["ms_per_step"], "train", x["sdf"]["train_fwd_bwd_bs10000"]["ms_per_step"],
      "surfel", x["surfel_ssim"]["value"], x["surfel_ssim"]["ms_per_step"], "mlp", x["mlp_batch"]["ms_per_step"])
PY
