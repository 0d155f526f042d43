run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 tools/bench_nshard.py $2 10 $3 2>&1 | grep NSHARD | cut -c1-330; }
for cfg in "" "epilogue=3" "epilogue=1"; do export NSHARD_CFG=$cfg; run 29601 65536 512; run 29602 65536 8192; run 29603 16384 8192; done
