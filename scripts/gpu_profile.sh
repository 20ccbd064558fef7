#!/bin/bash
# round-2 profiling pass (one gpurun call): plain run of each command first, then the launch list / full capture of it.
#   usage: TAG=r02x scripts/gpu_profile.sh [proj] [field] [c1r] [c3] [c5h768]
mkdir -p gpurun_out
prof() {   # tag, kernel regex, skip, count, command...
  local tag=$1 rx=$2 skip=$3 cnt=$4; shift 4
  "$@" > gpurun_out/${TAG:-r02}_plain_$tag.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG:-r02}_launches_$tag.csv "$@" > gpurun_out/${TAG:-r02}_ncu1_$tag.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:"$rx" -s $skip -c $cnt -o gpurun_out/${TAG:-r02}_prof_$tag "$@" > gpurun_out/${TAG:-r02}_ncu2_$tag.log 2>&1
  ls -la gpurun_out/${TAG:-r02}_prof_$tag* 2>/dev/null
}
for what in "$@"; do
  case $what in
    proj)   prof proj "proj_|gemm_bf16|cast_colsum" 9 16 python bench.py --extra-only projection_h768 --steps 2 ;;
    field)  prof field "field" 2 2 python bench.py --extra-only field --steps 2 ;;
    c1r)    prof c1r "grid4d_|small1d" 8 4 python bench.py --workload c1r --steps 2 --warmup 3 --no-cpu-baseline ;;
    c3)     prof c3 "grid4d_|regroup" 12 6 python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline ;;
    c5h768) prof c5h768 "grid4d_|proj_|gemm_bf16|cast_colsum|geo2xyzt" 27 12 python bench.py --workload c5h768 --steps 2 --warmup 3 --no-cpu-baseline ;;
  esac
done
