"""Developer tool: mean / median cycles per phase from an attention trace (tools/attn_lab.py --trace out.npy, library built
with -DATT_TRACE): softmax warp of lane quarter 0 (slots 0 start, 2 S in registers, 5 exponent reference fixed, 6 exponentials
done, 7 P stored) and the MMA issuer (role 4: 0 QK issued, 1 p_full seen, 2 PV issued).
    python tools/attn_trace_summary.py gpurun_out/trace.npy
"""
import numpy as np, sys
t = np.load(sys.argv[1]).astype(np.int64)   # [cta, role, blk, slot]
ncta, nrole, nblk, nslot = t.shape
# softmax role 0: slots 0 (start), 2 (S in regs), 5 (after max/exchange/o_full/rescale), 6 (exp done), 7 (P stored)
r = t[:, 0]
ok = (r[:, :, 0] > 0) & (r[:, :, 7] > 0)
def d(a, b):
    x = (r[:, :, b] - r[:, :, a])[ok]
    return x.mean(), np.median(x)
for name, a, b in (("wait s_full + ld", 0, 2), ("max + xchg + o_full + rescale", 2, 5), ("exp", 5, 6), ("P store", 6, 7)):
    print(name, "mean %.0f median %.0f" % d(a, b))
nxt = (r[:, 1:, 0] - r[:, :-1, 7])[ok[:, 1:] & ok[:, :-1]]
print("loop back / item gap: mean %.0f median %.0f p90 %.0f" % (nxt.mean(), np.median(nxt), np.percentile(nxt, 90)))
per = (r[:, 1:, 0] - r[:, :-1, 0])[ok[:, 1:] & ok[:, :-1]]
print("block period: mean %.0f median %.0f" % (per.mean(), np.median(per)))
m = t[:, 4]
okm = (m[:, :, 0] > 0) & (m[:, :, 2] > 0)
print("mma: qk issue -> pv issued mean %.0f ; p_full seen -> pv issued %.0f" % ((m[:, :, 2] - m[:, :, 0])[okm].mean(), (m[:, :, 2] - m[:, :, 1])[okm].mean()))
