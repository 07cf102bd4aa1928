for s in "1080 1920 16 32 64" "1080 1920 16 48 16" "720 1280 16 32 64" "1080 1920 8 32 23" "1080 1920 32 32 23" "2160 3840 16 32 16"; do
  for m in 2; do echo -n "ME_PERSISTENT=$m "; ME_PERSISTENT=$m python tools/run_one.py $s | sed -E "s/\{.*'resident_kernel': ([0-9]).*/resident=\1/"; done
done
