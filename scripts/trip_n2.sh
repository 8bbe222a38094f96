N=2
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 4 --warmup 3 > gpurun_out/bench_n$N.log 2> gpurun_out/bench_n$N.err; echo "bench n=$N rc=$?"
tail -n 1 gpurun_out/bench_n$N.log | cut -c1-400; tail -n 3 gpurun_out/bench_n$N.err
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --gpus $N --workload decode --steps 2 --warmup 3 > gpurun_out/bench_decode_n$N.log 2> gpurun_out/bench_decode_n$N.err; echo "decode n=$N rc=$?"
tail -n 1 gpurun_out/bench_decode_n$N.log | cut -c1-400; tail -n 3 gpurun_out/bench_decode_n$N.err
