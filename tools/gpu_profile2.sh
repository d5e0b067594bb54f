#!/bin/bash
# One GPU call that produces what is kept under profiles/ for round 2: tests, smoke, both bench arms, the ncu launch
# lists of BOTH arms, and `--set full` captures of the library's kernels (raw CSV pages written next to the reports so
# that the summaries can be made without shipping the .ncu-rep files back).
# usage (under gpurun): bash tools/gpu_profile2.sh <tag>
tag=${1:-r02}; out=gpurun_out/$tag; mkdir -p $out
(timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4) > $out/pytest.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep smoke) > $out/smoke.log
timeout 500 python bench.py --impl reference > $out/bench_ref.json 2> $out/bench_ref.err
timeout 500 python bench.py > $out/bench_ours.json 2> $out/bench_ours.err
NCU="ncu --clock-control none"
timeout 500 $NCU --metrics gpu__time_duration.sum -c 2500 --csv --log-file $out/launches_ours.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $out/ncu_launches.log 2>&1
timeout 500 $NCU --metrics gpu__time_duration.sum -c 400 --csv --log-file $out/launches_ref.csv \
  python bench.py --impl reference --steps 2 --warmup 1 --legs none > $out/ncu_launches_ref.log 2>&1
cap() {  # name, kernel regex, skip, count, command...
  local name=$1 re=$2 skip=$3 cnt=$4; shift 4
  timeout 700 $NCU --set full --import-source on -k regex:"$re" --launch-skip $skip -c $cnt -o $out/$name -f "$@" > $out/ncu_$name.log 2>&1
  ncu -i $out/$name.ncu-rep --page raw --csv > $out/${name}_raw.csv 2>/dev/null
}
# C2 step: skip the warm-up steps' launches, capture one full step of library kernels
cap c2 "^(r3_|band_|bucket_)" 42 14 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --legs none
cap c3 "^(rs_render|rs_pre|ssim_)" 10 5 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --legs surfel
cap c4 "^(spawn_|mlp_|np_|depth2normal|exposure|photo_)" 60 24 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --legs mapping
cap sdf "^(sdf_query)" 12 2 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --legs sdf
rm -f $out/*.ncu-rep
cat $out/pytest.log $out/smoke.log; tail -c 300 $out/bench_ours.json; ls -la $out
