for s in "1080 1920 16 128 4" "1080 1920 8 128 4" "1080 1920 32 128 4" "720 1280 16 128 8" "1080 1920 32 64 12" "1080 1920 16 120 4"; do
  for m in 0 2; do echo -n "ME_PERSISTENT=$m "; ME_PERSISTENT=$m python tools/run_one.py $s | sed -E "s/\{.*'tile_blocks_x': ([0-9]+).*'slabs_x': ([0-9]+), 'slabs_y': ([0-9]+).*'resident_kernel': ([0-9]).*'tile_blocks_y': ([0-9]+).*/tbx=\1 slabs=\2x\3 resident=\4 rows=\5/"; done
done
