run() { echo "== $*"; env "$@" ONLY=${ONLY:-gate_up} timeout 100 python scripts/bench_gemm_w4.py 2>&1 | grep -o "'shape': '[a-z_ ]*'\|'fused_ms': [0-9.]*\|'dequant+gemm_ms': [0-9.]*" | paste - - - | tr '\n' ' '; echo; }
run B200_W4_DEBUG=296
run B200_W4_DEBUG=40
run B200_W4_DEBUG=0
