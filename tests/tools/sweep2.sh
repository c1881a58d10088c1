python tests/tools/gpu_probe.py jc69_20x1000 hky_g4_flu:64:500 jc69_1000x10k hky_g4_flu 2>&1 | python tests/tools/fmt_probe.py
for nt in 768 512; do echo "== NT=$nt"; B2P_NT=$nt python tests/tools/gpu_probe.py jc69_1000x10k hky_g4_flu --nocheck 2>&1 | python tests/tools/fmt_probe.py; done
echo "== 2 CTAs/SM"; B2P_CTAS_PER_SM=2 python tests/tools/gpu_probe.py jc69_1000x10k hky_g4_flu 2>&1 | python tests/tools/fmt_probe.py
echo "== f32"; python tests/tools/gpu_probe.py jc69_1000x10k hky_g4_flu --f32 2>&1 | python tests/tools/fmt_probe.py
