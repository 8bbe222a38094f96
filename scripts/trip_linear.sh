mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TAILN=4 run tr_qkv4p 100 python scripts/trace_linear.py run qkv 4 1
TAILN=4 run tr_o4 100 python scripts/trace_linear.py run o 4 1
TAILN=4 run tr_o8 100 python scripts/trace_linear.py run o 8 1
TAILN=4 run tr_down 100 python scripts/trace_linear.py run down 8 1
TAILN=4 run tr_gu 100 python scripts/trace_linear.py run gu 1 1
