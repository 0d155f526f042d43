for shp in "16 4096 4096" "64 8192 8192" "512 2048 4096" "1024 1024 1024" "2048 512 8192" "2048 2048 2048"; do
  echo "== $shp"
  ncu --metrics gpu__time_duration.sum --clock-control none --csv python tools/run_gemm.py $shp bf16 bf16 3 2>/dev/null | grep -E "gemm_tcgen05|reduce|repitch" | tail -4 | awk -F'","' '{gsub(/"/,"",$NF); printf("   ours   %-70s %s %s\n", substr($5,1,70), $NF, $(NF-1))}'
  ncu --metrics gpu__time_duration.sum --clock-control none --csv python tools/run_cublas.py $shp 3 2>/dev/null | grep -E "nvjet|cutlass|gemm|splitK|reduce" | tail -3 | awk -F'","' '{gsub(/"/,"",$NF); printf("   cublas %-70s %s %s\n", substr($5,1,70), $NF, $(NF-1))}'
done
