mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
python scripts/debug_linear_caps.py
TAILN=6 run t_linear 300 python -m pytest tests/test_linear_decode_gpu.py tests/test_engine_gpu.py -q -x
TAILN=12 run prof_chain 200 python scripts/profile_linear.py 10 chain
TAILN=3 run bench_decode 900 python bench.py --workload decode --steps 3 --warmup 3
