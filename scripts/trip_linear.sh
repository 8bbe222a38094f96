mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TAILN=30 run t_linear 300 python -m pytest tests/test_linear_decode_gpu.py -q -x
TAILN=12 run prof_linear 300 python scripts/profile_linear.py 10
