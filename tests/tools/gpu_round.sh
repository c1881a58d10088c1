#!/bin/bash
# One GPU call: tests, smoke, bench, then ncu launch list + full capture of the traversal kernel.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.log 2>&1; nproc >> gpurun_out/smi.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/bench_ref.json 2>&1; cat gpurun_out/bench_ref.json
BENCH_SMALL="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$BENCH_SMALL > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv $BENCH_SMALL > gpurun_out/ncu1.log 2>&1
echo "ncu launches rc=$?"
$BENCH_SMALL > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:walk_kernel -s 3 -c 1 -f -o gpurun_out/prof $BENCH_SMALL > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out
