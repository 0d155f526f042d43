for shp in "16 4096 4096" "64 4096 4096"; do
for cfg in "" "cta_group=1 block_n=32 split_k=1" "cta_group=1 block_n=64 split_k=1" "cta_group=1 block_n=32 split_k=2" "cta_group=1 block_n=64 split_k=2" "cta_group=1 block_n=128 split_k=1" "cta_group=1 block_n=32 split_k=1 stages=12"; do
  for mode in cold ""; do
  echo "== $shp $mode $cfg"
  TLB200_LIB=tools/_ab/libtlb200_trace.so timeout 120 python tools/trace_tiles.py $shp small $mode $cfg 2>&1 | grep -E "main loop|rror|epilogue warp done|'stages'" | sed "s/.*'stages': \([0-9]*\).*'split_k': \([0-9]*\).*/stages \1 split \2/" | tr "\n" " "; echo
  done
done; done
