#!/bin/bash
# Run ON THE GPU BOX (via gpurun) from the repo root: collects the per-round ncu evidence into gpurun_out/.
#   tools/profile_round.sh <tag> [extra bench.py flags, e.g. --quick-per-env]
#   1. steady-state launch list (gpu__time_duration, --clock-control none) of two steps
#   2. one full-set capture (--set full --import-source on) of one step
# Each ncu run is preceded by the same command without ncu (must exit 0), as the profiling recipe requires.
set -u
R=${1:-r02}
shift || true
export GSW_PROFILE_RANGE=1
python bench.py --quick "$@" --steps 2 --warmup 3 > gpurun_out/${R}_plain_a.log 2>&1 &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/${R}_launches.csv python bench.py --quick "$@" --steps 2 --warmup 3 > gpurun_out/${R}_ncu_a.log 2>&1
echo "launch list rc=$?"
python bench.py --quick "$@" --steps 1 --warmup 3 > gpurun_out/${R}_plain_b.log 2>&1 &&
ncu --profile-from-start off --set full --clock-control none --import-source on -o gpurun_out/${R}_step \
    python bench.py --quick "$@" --steps 1 --warmup 3 > gpurun_out/${R}_ncu_b.log 2>&1
echo "full capture rc=$?"
