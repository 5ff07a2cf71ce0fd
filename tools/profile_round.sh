#!/bin/bash
# One gpurun call that refreshes the round's ncu evidence (B200_PROFILING.md recipe).  Every ncu command is preceded
# by the identical plain command exiting 0.  Outputs under gpurun_out/prof_$TAG/.
#   gpurun --timeout 1500 -- 'bash tools/profile_round.sh r1b'
TAG=${1:-r1}
MODE=${2:-all}   # all | list (launch list only)
OUT=gpurun_out/prof_$TAG
mkdir -p $OUT
# warm-up 1: the ncu pass replays every launch of every bench pass (r1g: 22 k metric rows, 580 s of box time at --warmup 3)
BENCH="python bench.py --steps 1 --warmup ${PROFILE_WARMUP:-1} --no-cpu ${PROFILE_BENCH_FLAGS:-}"
# (1) launch list of one bench command: device time + DRAM bytes of every launch
$BENCH > $OUT/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file $OUT/launches.csv $BENCH > $OUT/ncu_bench.log 2>&1
[ "$MODE" = list ] && { ls -la $OUT; exit 0; }
# (2) --set full on the dominant shapes, each timed alone by the microbench tools (2 launches each)
full() {  # name, kernel regex, skip, count, command...
  local name=$1 rx=$2 skip=$3 cnt=$4; shift 4
  "$@" > $OUT/plain_$name.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -f -o $OUT/$name "$@" \
      > $OUT/ncu_$name.log 2>&1
}
full gemm_geglu_k320  gemm_kernel 1 1 python tools/bench_gemm.py "geglu M217728"
full gemm_lin_k320    gemm_kernel 1 1 python tools/bench_gemm.py "linear M217728 N320 K320 res1"
full gemm_conv_ds1    gemm_kernel 1 1 python tools/bench_gemm.py "conv3x3 n42 H72 Cin320"
full gemm_conv_ds4    gemm_kernel 1 1 python tools/bench_gemm.py "conv3x3 n42 H18 Cin1280"
full attn_3d_ds2      attn_kernel 1 1 python tools/bench_attn.py 3d_ds2
full attn_2d_ds1      attn_kernel 1 1 python tools/bench_attn.py 2d_ds1
full gn               "gn_(stats|apply)" 5 2 python tools/bench_kernels.py gn
full gn_finalize      gn_finalize 20 3 python bench.py --steps 1 --warmup 1 --no-cpu --no-gpu-baseline
full upconv_ds1       gemm_kernel 1 1 python tools/bench_gemm.py "upconv"
ls -la $OUT
