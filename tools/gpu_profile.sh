#!/bin/bash
# One GPU call that produces everything kept under profiles/ for a round: tests, smoke, both bench arms,
# the ncu launch list of the bench command and `--set full` captures of the library's kernels.
# usage (under gpurun): bash tools/gpu_profile.sh <tag>
tag=${1:-vX}; out=gpurun_out/$tag; mkdir -p $out
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) > $out/pytest.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep smoke) > $out/smoke.log
timeout 400 python bench.py --impl reference > $out/bench_ref.json 2> $out/bench_ref.err
timeout 400 python bench.py > $out/bench_ours.json 2> $out/bench_ours.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/ncu_launches.log 2>&1
# C2 step kernels: skip the warm-up steps' launches, capture one full step of library kernels
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^(r3_|band_|depth_sort|tile_|bin_|photo_)" \
  --launch-skip 45 -c 15 -o $out/c2_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/ncu_c2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^(rs_render|ssim_|sdf_query|mlp_tc|np_)" \
  --launch-skip 6 -c 8 -o $out/c3_full -f python tools/dev_c3_kernels.py > $out/ncu_c3.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:"^(sdf_query|mlp_tc2?_forward|np_query)" \
  --launch-skip 3 -c 4 -o $out/neural_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/ncu_neural.log 2>&1
cat $out/pytest.log $out/smoke.log; tail -c 400 $out/bench_ours.json; ls -la $out
