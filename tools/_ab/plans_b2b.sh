# back-to-back (L2-warm) timing of a few explicit plans per small shape
run() { python tools/run_gemm.py $1 $2 $3 bf16 bf16 200 ${@:4} 2>&1 | tail -1 | sed 's/run_gemm //'; }
for shp in "512 2048 4096" "1024 1024 1024" "2048 512 8192" "256 4096 4096"; do
  echo "== $shp"
  run $shp
  run $shp cta_group=2 block_n=64 split_k=1
  run $shp cta_group=2 block_n=128 split_k=1
  run $shp cta_group=2 block_n=128 split_k=2
  run $shp cta_group=2 block_n=256 split_k=2
  run $shp cta_group=2 block_n=256 split_k=4
  run $shp cta_group=1 block_n=64 split_k=1
  run $shp cta_group=1 block_n=128 split_k=1
  run $shp cta_group=1 block_n=128 split_k=2
  run $shp cta_group=1 block_n=256 split_k=4
done
