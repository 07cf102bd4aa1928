for s in "1080 1920 16 32 64" "1080 1920 8 32 23" "720 1280 8 32 51" "1080 1920 8 48 12" "720 1280 16 48 32" "1080 1920 32 32 23" "720 1280 32 32 53" "1080 1920 16 16 32" "1080 1920 32 16 32" "1080 1920 32 48 12" "360 640 16 32 64" "2160 3840 8 32 6"; do
  for m in 1 0; do echo -n "ME_PERSISTENT=$m "; ME_PERSISTENT=$m python tools/run_one.py $s | sed -E "s/\{.*'tile_blocks_x': ([0-9]+).*'slabs_x': ([0-9]+), 'slabs_y': ([0-9]+).*'resident_kernel': ([0-9]).*'tile_blocks_y': ([0-9]+).*/tbx=\1 slabs=\2x\3 resident=\4 rows=\5/"; done
done
