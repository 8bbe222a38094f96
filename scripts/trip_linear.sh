mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log | cut -c1-160; }
TAILN=1 run bench_a 900 python bench.py --workload decode --steps 3 --warmup 3 --no-cpu-baseline
CRAM_B200_PDL=0 TAILN=1 run bench_nopdl 900 python bench.py --workload decode --steps 3 --warmup 3 --no-cpu-baseline
CRAM_B200_STREAMING_PROJ=1 TAILN=1 run bench_proj 900 python bench.py --workload decode --steps 3 --warmup 3 --no-cpu-baseline
CRAM_B200_STREAMING_PROJ=1 CRAM_B200_PDL=0 TAILN=1 run bench_proj_nopdl 900 python bench.py --workload decode --steps 3 --warmup 3 --no-cpu-baseline
