for cfg in "" "cta_group=1 block_n=64 split_k=1 kgroup=1" "cta_group=1 block_n=64 split_k=1 kgroup=2" "cta_group=1 block_n=64 split_k=1 kgroup=4" "cta_group=1 block_n=128 split_k=2 kgroup=1" "cta_group=1 block_n=128 split_k=2 kgroup=2" "cta_group=1 block_n=128 split_k=2 kgroup=4" "cta_group=1 block_n=32 split_k=1 kgroup=4 stages=12"; do
  echo "== 64 8192 8192 cold $cfg"
  TLB200_LIB=tools/_ab/libtlb200_trace.so timeout 120 python tools/trace_tiles.py 64 8192 8192 small cold $cfg 2>&1 | grep -E "main loop|rror|epilogue warp done|'stages'" | cut -c1-200
done
