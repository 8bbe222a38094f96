mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TAILN=30 run t_attn 900 python -m pytest tests/test_attention_gpu.py -q -x
run prof_attn 300 python scripts/profile_attn.py 10
