"""Developer tool: summarise the clock64 stamps a -DSEVA_DEBUG -DSEVA_ATTN_TRACE build of attn.cu writes for block (0,0,0)
(SEVA_ATTN_TRACE_FILE=<path>).  Rows: role (0 / 1 = softmax warp 0 of Q tile 0 / 1, 2 = MMA warp), key block, 8 stamps."""
import sys
import numpy as np
d = np.loadtxt(sys.argv[1], dtype=np.int64)
t = d[:, 2:].reshape(3, 64, 8)
t0 = t[t > 0].min()
rel = np.where(t > 0, t - t0, -1)
lo, hi = 10, 60
for r in range(3):
    print("role", r, "mean period", np.diff(rel[r, lo:hi, 0]).mean())
names_s = ["s_ready", "ld_done", "max_done", "pre_exp", "exp_done", "p1_arrived", "half0_stored", "-"]
for r in range(2):
    x = rel[r, lo:hi]
    order = [0, 1, 2, 3, 6, 4, 5]
    seq = x[:, order]
    print("softmax", r, "stamps", [names_s[k] for k in order])
    print("   mean deltas", np.round(np.diff(seq, axis=1).mean(0), 1), " end->next s_ready", (rel[r, lo + 1:hi, 0] - rel[r, lo:hi - 1, 5]).mean())
x = rel[2, lo:hi]
cols = [k for k in range(8) if (x[:, k] >= 0).all()]
print("mma stamps", cols, "mean deltas", np.round(np.diff(x[:, cols], axis=1).mean(0), 1), "end->next", (rel[2, lo + 1:hi, 0] - rel[2, lo:hi - 1, cols[-1]]).mean())
j = 30
print("sample j=30: sm0", rel[0, j], "sm1", rel[1, j], "mma", rel[2, j])
