for cfg in "8 1 32" "8 2 64" "16 1 32" "16 2 64" "16 1 23" "16 2 34" "8 1 23"; do
set -- $cfg
echo "== W=$1 NC=$2 CPT=$3"
B2P_WARPS=$1 B2P_NC=$2 B2P_COLS_PER_TILE=$3 python tests/tools/gpu_probe.py jc69_1000x10k hky_g4_flu --nocheck 2>&1 | python tests/tools/fmt_probe.py
done
