#!/bin/bash
# bench lines of the other BASELINE configs (one GPU): flu-shaped HKY+G4 fp64/fp32, the north-star
# gate shape GTR+G4 5000 x 10k, the codon path.
mkdir -p gpurun_out
python bench.py --workload hky_g4_flu --steps 100 --cpu-seconds 5 > gpurun_out/bench_hky_f64.json 2> gpurun_out/bench_others.err; tail -c 900 gpurun_out/bench_hky_f64.json
python bench.py --workload hky_g4_flu --precision float32 --steps 100 --no-cpu-baseline > gpurun_out/bench_hky_f32.json 2>> gpurun_out/bench_others.err; tail -c 600 gpurun_out/bench_hky_f32.json
python bench.py --workload gtr_g4_5000x10k --steps 30 --warmup 5 --cpu-seconds 10 > gpurun_out/bench_gtr_5000x10k.json 2>> gpurun_out/bench_others.err; tail -c 900 gpurun_out/bench_gtr_5000x10k.json
python bench.py --workload codon_500x2000 --steps 20 --warmup 3 --cpu-seconds 10 > gpurun_out/bench_codon.json 2>> gpurun_out/bench_others.err; tail -c 900 gpurun_out/bench_codon.json
tail -5 gpurun_out/bench_others.err
