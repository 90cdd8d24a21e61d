#!/bin/bash
# experiment sweep for the update kernel tunables (run under gpurun)
for cfg in "8 1 0" "8 0 0" "4 1 0" "16 1 0" "8 1 2" "4 1 4" "4 1 3"; do
  set -- $cfg
  echo "== COLS=$1 SERP=$2 BPS=$3"
  TCI_RRLU_COLS=$1 TCI_RRLU_SERPENTINE=$2 TCI_RRLU_BLOCKS_PER_SM=$3 python scripts/quick_rrlu.py 2>&1 | awk 'NR%3==0'
done
