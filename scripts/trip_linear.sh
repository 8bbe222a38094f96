mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log | cut -c1-160; }
TAILN=1 run bench_a 900 python bench.py --workload decode --steps 3 --warmup 3 --no-cpu-baseline
CRAM_B200_LINEAR_RING=4 TAILN=1 run bench_ring4 900 python bench.py --workload decode --steps 3 --warmup 3 --no-cpu-baseline
TAILN=1 run bench_b 900 python bench.py --workload decode --steps 3 --warmup 3 --no-cpu-baseline
CRAM_B200_LINEAR_RING=4 TAILN=1 run bench_ring4b 900 python bench.py --workload decode --steps 3 --warmup 3 --no-cpu-baseline
CRAM_B200_STREAMING_PROJ=1 CRAM_B200_LINEAR_RING=4 TAILN=1 run bench_proj_ring4 900 python bench.py --workload decode --steps 3 --warmup 3 --no-cpu-baseline
