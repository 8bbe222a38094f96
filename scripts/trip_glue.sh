mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TAILN=4 run t_glue 600 python -m pytest tests/test_glue_gpu.py -q -x
echo "--- previous build"; CRAM_B200_LIB=cram_b200/_build/libcram_b200_prev.so python scripts/time_glue.py
echo "--- this build"; python scripts/time_glue.py
