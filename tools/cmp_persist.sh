#!/bin/bash
# default plan vs one-tile kernel (ME_PERSISTENT=0) vs resident kernel forced (ME_PERSISTENT=2) on a list of shapes
cd "$(dirname "$0")/.."
for s in "1080 1920 16 32 64" "1080 1920 16 64 8" "1080 1920 16 24 32" "1080 1920 16 16 32" "1080 1920 16 48 16" "720 1280 16 32 64" \
         "1080 1920 8 32 23" "2160 3840 8 64 4" "1080 1920 32 32 23" "1080 1920 32 64 6" "1080 1920 16 100 2" "64 640 16 32 64"; do
  for m in 1 0 2; do
    echo -n "ME_PERSISTENT=$m "; ME_PERSISTENT=$m python tools/run_one.py $s | sed -E "s/\{.*'resident_kernel': ([0-9]).*/resident=\1/"
  done
done
