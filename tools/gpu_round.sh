#!/bin/bash
# One GPU visit: parity tests, smoke, bench (both arms), ncu launch list, one ncu --set full capture.
# Everything lands under gpurun_out/<tag>/ ; run as: gpurun --timeout 1500 -- 'bash tools/gpu_round.sh r01b'
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
nproc > $OUT/nproc.txt
python -c "import torchdiffeq; print(torchdiffeq.__version__)" > $OUT/torchdiffeq.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
if [ -n "$SWEEP" ]; then timeout 600 python tools/sweep.py --out $OUT/sweep_dopri5.md > $OUT/sweep.log 2>&1; echo "sweep rc=$?" | tee -a $OUT/rc.txt; fi
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" | tee -a $OUT/rc.txt
if [ -z "$SKIP_NCU" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $OUT/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-certify --no-fi --no-segway --no-sweep --min-seconds 0 > $OUT/ncu_launch.log 2>&1; echo "ncu-list rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -c 40 -f -o $OUT/full \
  python tools/prof_kernels.py all > $OUT/ncu_full.log 2>&1; echo "ncu-full rc=$?" | tee -a $OUT/rc.txt
fi
tail -3 $OUT/pytest_gpu.log; tail -2 $OUT/smoke.log; cat $OUT/bench.json | cut -c1-1500
