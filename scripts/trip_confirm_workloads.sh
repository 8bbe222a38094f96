# HEAD confirmation of the two other workloads (cfg4 decode, cfg5 long-context prefill)
mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-3} gpurun_out/$name.log; }
run bench_decode 900 python bench.py --workload decode --steps 3 --warmup 3
run bench_longctx 900 python bench.py --workload longctx --steps 2 --warmup 3
