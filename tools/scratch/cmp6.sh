for s in "1080 1920 16 8 64" "1080 1920 16 16 32" "1080 1920 16 24 32" "720 1280 16 16 64" "2160 3840 16 8 16"; do
  python tools/run_one.py $s | sed -E "s/\{.*'tile_blocks_x': ([0-9]+).*'slabs_x': ([0-9]+), 'slabs_y': ([0-9]+).*'resident_kernel': ([0-9]).*'tile_blocks_y': ([0-9]+).*/tbx=\1 slabs=\2x\3 resident=\4 rows=\5/"
done
for s in "1080 1920 8 8 64" "1080 1920 8 16 32" "1080 1920 32 16 32" "1080 1920 32 8 64"; do
  python tools/run_one.py $s | sed -E "s/\{.*'tile_blocks_x': ([0-9]+).*'slabs_x': ([0-9]+), 'slabs_y': ([0-9]+).*'resident_kernel': ([0-9]).*'tile_blocks_y': ([0-9]+).*/tbx=\1 slabs=\2x\3 resident=\4 rows=\5/"
done
