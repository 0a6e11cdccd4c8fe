#!/bin/bash
# sweep of the raw-frame kernel's tuning overrides (column segments, threads per CTA)
for cfg in "0 512" "0 448" "0 256" "7 512" "2 512" "2 448"; do
  set -- $cfg
  echo "== NSEG=$1 THREADS=$2"
  LVC_PIL_NSEG=$1 LVC_PIL_THREADS=$2 timeout 100 python scripts/perf_hfraw.py 2>&1 | grep -E "raw +(1080|720)"
done
