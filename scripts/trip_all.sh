# full GPU trip: every -m gpu test, smoke, default bench, reference arm, the two other workloads, launch list, ncu --set full of attention
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TAILN=25 run t_gpu 1500 python -m pytest tests -m gpu -q -x
run smoke 600 python __graft_entry__.py --smoke
TAILN=3 run bench 1500 python bench.py
TAILN=3 run bench_ref 900 python bench.py --impl reference --steps 2 --warmup 1
TAILN=3 run bench_decode 1500 python bench.py --workload decode --steps 3 --warmup 3
TAILN=3 run bench_longctx 1800 python bench.py --workload longctx --steps 2 --warmup 3
run prof_attn 300 python scripts/profile_attn.py 10
