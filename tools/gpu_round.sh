#!/bin/bash
# One GPU-box pass: gpu tests, default bench line, ncu launch list, one full capture of findNeighborsKernel and one
# full capture of every other kernel of one steady-state step.
# usage (under gpurun): bash tools/gpu_round.sh <tag> [skip-tests]
tag=${1:-rXX}; out=gpurun_out
mkdir -p $out
if [ -z "$2" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
  tail -3 $out/${tag}_pytest.log
fi
timeout 900 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?"
tail -c 3000 $out/${tag}_bench.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $out/${tag}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > $out/${tag}_ncu_launch.log 2>&1; echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:findNeighborsKernel -s 3 -c 1 -f -o $out/${tag}_nb \
  python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > $out/${tag}_ncu_full.log 2>&1; echo "ncu full exit $?"
# the large streaming kernels of one steady-state step (full set; the small tree kernels: see gpu_kernels_small below)
if [ -n "$3" ]; then
timeout 900 ncu --set full --clock-control none -k 'regex:onesweep|sfcKeys|gather8x4|gatherMulti|minMaxPartial|histogram|nodeCounts|haloRadii|geoCenters|findHalos' -s 90 -c 36 -f -o $out/${tag}_kernels \
  python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > $out/${tag}_ncu_kernels.log 2>&1; echo "ncu kernels exit $?"
fi
