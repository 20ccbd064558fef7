#!/bin/bash
# One parameterised GPU-box runner (replaces round 1's one-off scripts).  Usage: scripts/gpu_round2.sh TAG [steps...]
#   steps: test  = pytest -m gpu              bench = default bench.py line (c2 + extras)
#          wl:X  = bench.py --workload X      ncu:X = launch list + full capture of workload X's kernels
TAG=${1:-r02}; shift
mkdir -p gpurun_out
for s in "$@"; do
  case $s in
    test) timeout 1500 python -m pytest tests -m gpu -x -q -s 2>&1 | tail -40 > gpurun_out/${TAG}_pytest.txt; tail -5 gpurun_out/${TAG}_pytest.txt ;;
    test:*) timeout 900 python -m pytest tests -m gpu -x -q -s -k "${s#test:}" 2>&1 | tail -40 > gpurun_out/${TAG}_pytest_k.txt; tail -15 gpurun_out/${TAG}_pytest_k.txt ;;
    bench) timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; tail -3 gpurun_out/${TAG}_bench.err; head -c 600 gpurun_out/${TAG}_bench.json; echo ;;
    wl:*) w=${s#wl:}; timeout 600 python bench.py --workload $w --steps 5 > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err; tail -3 gpurun_out/${TAG}_bench_$w.err; head -c 400 gpurun_out/${TAG}_bench_$w.json; echo ;;
    ncu:*) w=${s#ncu:}
      CMD="python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline"
      $CMD > gpurun_out/${TAG}_plain_$w.log 2>&1 &&
      ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_$w.csv $CMD > gpurun_out/${TAG}_ncu1_$w.log 2>&1
      ncu --set full --clock-control none --import-source on -k regex:"grid4d_|regroup|mlp_head|proj_" -s 12 -c 6 -o gpurun_out/${TAG}_prof_$w $CMD > gpurun_out/${TAG}_ncu2_$w.log 2>&1
      ls -la gpurun_out/${TAG}_prof_$w* ;;
    *) echo "unknown step $s" ;;
  esac
done
