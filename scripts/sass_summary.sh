#!/bin/bash
# Per-kernel SASS mnemonic counts of simubox_b200/libsimubox_b200.so (runs on the build machine: cuobjdump only).
# DMMA = fp64 mma.sync (the pruned DFT's reduction engine), UBLKCP = cp.async.bulk (TMA engine, 1-D),
# UTMALDG / UTMASTG = tensor-map TMA, SYNCS = mbarrier ops, UTC*MMA / LDTM / STTM = tcgen05 (none: no fp64 kind),
# ACQBULK / PREEXIT = programmatic dependent launch, REDUX / VIMNMX = warp reductions and integer min/max.
LIB=${1:-simubox_b200/libsimubox_b200.so}
cuobjdump -sass "$LIB" | awk '
  /Function :/ { fn=$3; next }
  fn != "" {
    for (i = 1; i <= NF; ++i) {
      m=$i; sub(/\..*/, "", m)
      if (m ~ /^(DMMA|HMMA|UBLKCP|UTMALDG|UTMASTG|SYNCS|LDTM|STTM|ACQBULK|PREEXIT|REDUX|LDGSTS|DSETP|DADD|DMUL|DFMA|ATOMS|ATOMG|RED)$/ || m ~ /^UTC.*MMA$/) c[fn, m]++
    }
    if ($0 ~ /\/\*[0-9a-f]{4}\*\//) n[fn]++
  }
  END {
    for (f in n) {
      line = sprintf("%-110s instr %6d", f, n[f])
      split("DMMA UBLKCP UTMALDG SYNCS ACQBULK PREEXIT REDUX LDGSTS DFMA DADD DMUL DSETP ATOMS RED", ms, " ")
      for (k = 1; k <= 14; ++k) if (c[f, ms[k]] > 0) line = line sprintf("  %s %d", ms[k], c[f, ms[k]])
      print line
    }
  }' | sort
echo "--- totals"
cuobjdump -sass "$LIB" | grep -oE "\b(DMMA|HMMA|UBLKCP|UTMALDG|UTMASTG|SYNCS|LDTM|STTM|ACQBULK|PREEXIT|REDUX|UTC[A-Z]*MMA)\b" | sort | uniq -c
