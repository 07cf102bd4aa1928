for s in "1080 1920 8 32 23" "720 1280 8 32 51" "2160 3840 8 32 6" "1080 1920 8 40 16" "1080 1920 8 48 12" "1080 1920 16 40 24" "1080 1920 16 28 32" "720 1280 16 48 32" "1080 1920 32 24 32" "1080 1920 32 40 16" "720 1280 32 32 53"; do
  for m in 0 2; do echo -n "ME_PERSISTENT=$m "; ME_PERSISTENT=$m python tools/run_one.py $s | sed -E "s/\{.*'tile_blocks_x': ([0-9]+).*'slabs_x': ([0-9]+), 'slabs_y': ([0-9]+).*'resident_kernel': ([0-9]).*'tile_blocks_y': ([0-9]+).*/tbx=\1 slabs=\2x\3 resident=\4 rows=\5/"; done
done
