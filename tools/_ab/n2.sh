run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 tools/bench_nshard.py $2 10 $3 2>&1 | grep NSHARD; }
echo "== new (lane-parallel, rotated)"; run 29601 65536 512; run 29602 65536 8192; run 29603 16384 8192
export TLB200_LIB=$PWD/tools/_ab/libtlb200_serial_peer.so
echo "== old (one lane, serial)"; run 29604 65536 512; run 29605 65536 8192; run 29606 16384 8192
