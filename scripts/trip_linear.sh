mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TAILN=3 run bench_prefill 900 python bench.py --steps 8 --warmup 3
TAILN=3 run bench_decode 900 python bench.py --workload decode --steps 3 --warmup 3
