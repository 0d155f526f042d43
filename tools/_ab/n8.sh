run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 tools/bench_nshard.py $2 10 $3 2>&1 | grep NSHARD; }
echo "== new (lane-parallel, rotated)"; run 29601 65536 8192; run 29602 65536 512
TLB200_LIB=$PWD/tools/_ab/libtlb200_serial_peer.so bash -c 'echo "== old (one lane, serial)"'
export TLB200_LIB=$PWD/tools/_ab/libtlb200_serial_peer.so
run 29604 65536 8192; run 29605 65536 512
unset TLB200_LIB
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 8 --steps 20 --warmup 3 2>/dev/null | grep '^{' > gpurun_out/bench_n8_r2a.json
cut -c1-160 gpurun_out/bench_n8_r2a.json
