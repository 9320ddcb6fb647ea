#!/bin/bash
# halo3x3 in-kernel probes (B2SEG_PROBE bits: 1 no stores, 2 no stats, 4 no MMAs, 8 MMAs x3, 16 no A traffic, 32 no TMEM read)
for pr in "$@"; do echo "== PROBE $pr"; B2SEG_PROBE=$pr NOFLUSH=1 python tools/conv_bench.py 256,64,64,64,64 2>&1 | grep -E "fprop|dgrad "; done
