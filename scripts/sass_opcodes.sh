#!/bin/bash
# Per-kernel counts of the SASS mnemonics that prove which hardware path a kernel uses (B200_PROFILING.md):
# UTC*MMA (tcgen05.mma), LDTM/STTM (tcgen05.ld/st), UTMALDG (TMA tensor load), UBLKCP (bulk copy),
# UTCBAR (tcgen05.commit), HMMA (legacy mma.sync), SYNCS (mbarrier), MUFU (special function unit).
# Usage: scripts/sass_opcodes.sh > profiles/sass_opcodes.txt   (needs the built liblwc_b200.so)
set -euo pipefail
LIB="$(dirname "$0")/../llm-weighted-consensus_b200/liblwc_b200.so"
echo "# cuobjdump -sass $(basename "$LIB"), $(nvcc --version | tail -2 | head -1)"
echo "# kernel | UTCxMMA LDTM STTM UTMALDG UBLKCP UTCBAR HMMA SYNCS MUFU FFMA total"
cuobjdump -sass "$LIB" | awk '
  /Function :/ { if (name != "") flush(); name=$3; delete c; total=0; next }
  /^[ \t]+\/\*[0-9a-f]+\*\// {
    total++
    if ($0 ~ /UTC[A-Z]*MMA/) c["mma"]++
    if ($0 ~ /LDTM/) c["ldtm"]++
    if ($0 ~ /STTM/) c["sttm"]++
    if ($0 ~ /UTMALDG/) c["tma"]++
    if ($0 ~ /UBLKCP/) c["blk"]++
    if ($0 ~ /UTCBAR/) c["bar"]++
    if ($0 ~ /[ \t]HMMA/) c["hmma"]++
    if ($0 ~ /SYNCS/) c["syncs"]++
    if ($0 ~ /MUFU/) c["mufu"]++
    if ($0 ~ /FFMA/) c["ffma"]++
  }
  function flush() { printf "%s | %d %d %d %d %d %d %d %d %d %d %d\n", name, c["mma"], c["ldtm"], c["sttm"], c["tma"], c["blk"], c["bar"], c["hmma"], c["syncs"], c["mufu"], c["ffma"], total }
  END { if (name != "") flush() }' | while IFS='|' read -r name rest; do echo "$(echo "$name" | c++filt | sed 's/(.*//') |$rest"; done | sort
