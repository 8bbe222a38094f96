mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TAILN=15 run t_glue 600 python -m pytest tests/test_glue_gpu.py tests/test_linear_decode_gpu.py tests/test_engine_gpu.py -q -x
python scripts/time_glue.py 2>&1 | head -5
