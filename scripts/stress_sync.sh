#!/bin/bash
# Synchronisation stress: with the arithmetic switched off (PVA_TILED_DEBUG=4/5) the consumers run as
# far ahead as the barriers allow, which is how the parity-aliasing race was found.  Any failure here
# is a protocol bug even if the normal mode never shows it.
for d in 4 5 6; do for b in 1 3 5 6 7 9 12 16 20; do
  out=$(PVA_TILED_DEBUG=$d timeout 60 python scripts/dbg_once.py $b 6 0 2>&1 | grep -E "all ok|rror" | head -1)
  echo "debug=$d batch=$b: $out"
done; done
