#!/bin/bash
# Quick GPU check: parity tests (stop at first failure) + probe timings.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/pytest_gpu.log
timeout 300 python tests/tools/gpu_probe.py jc69_20x1000 jc69_1000x10k hky_g4_flu 2>&1 | python tests/tools/fmt_probe.py | tee gpurun_out/probe.log
