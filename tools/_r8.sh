run() { env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 tools/ranked_bench.py c3 2>/dev/null | grep '^{' ; }
run A=1
run GRITPACK_COMM_CTAS=16 GRITPACK_COMM_RESERVE=16
run GRITPACK_COMM_CTAS=32 GRITPACK_COMM_RESERVE=8
run GRITPACK_COMM_SLICES=16
run GRITPACK_COMM_CTAS=16 GRITPACK_COMM_RESERVE=16 GRITPACK_COMM_SLICES=16
