for s in "720 1280 8 32 51" "2160 3840 8 32 16" "1080 1920 8 32 23" "1080 1920 32 48 12" "720 1280 32 48 24" "1080 1920 16 24 32" "1080 1920 16 32 64"; do
  for m in 1; do echo -n "ME_PERSISTENT=$m "; ME_PERSISTENT=$m python tools/run_one.py $s | sed -E "s/\{.*'tile_blocks_x': ([0-9]+).*'slabs_x': ([0-9]+), 'slabs_y': ([0-9]+).*'resident_kernel': ([0-9]).*'tile_blocks_y': ([0-9]+).*/tbx=\1 slabs=\2x\3 resident=\4 rows=\5/"; done
done
ME_TMR=2 python tools/run_one.py 2160 3840 8 32 16 | cut -c1-70
