V=stable-virtual-camera_b200/lib/variants
for p in 0 1; do for h in 1 2; do for emu in 0 4 8; do echo "== ping=$p halves=$h emu=$emu"; SEVA_B200_LIB=$V/libseva_xbase.so SEVA_ATTN_PING=$p SEVA_ATTN_HALVES=$h SEVA_ATTN_EMU=$emu python tools/bench_attn.py 3d_ds2 2>&1 | tail -n 1; done; done; done
for h in 1 2; do SEVA_ATTN_PING=1 SEVA_ATTN_HALVES=$h SEVA_B200_LIB=$V/libseva_xtrace.so SEVA_ATTN_TRACE_FILE=gpurun_out/attn_trace5_h$h.txt python tools/bench_attn.py 3d_ds2 --once; done
