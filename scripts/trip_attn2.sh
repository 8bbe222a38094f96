mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TAILN=4 run t_attn 900 python -m pytest tests/test_attention_gpu.py tests/test_engine_gpu.py tests/test_sweep_gpu.py -q -x
TAILN=3 run prof_attn 300 python scripts/profile_attn.py 10
TAILN=3 run prof_attn2 300 python scripts/profile_attn.py 10
