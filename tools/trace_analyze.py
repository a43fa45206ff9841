"""Schedule analysis from the per-task trace (gpb_set_trace): phase durations per task type."""
import sys, os, math
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G

def gs(batch, rt, j): return batch * (1 + j * rt - j * (j + 1) // 2 + j)
def decode(batch, nt, rt, t):
    G = int(os.environ.get("GPB_TICKET_GROUP", "0")) or batch
    G = min(G, batch)
    per_group = G * (nt * rt - nt * (nt - 1) // 2)
    gi, t = divmod(t, per_group)
    gb = min(G, batch - gi * G)
    b, i, j = decode1(gb, nt, rt, t)
    return (b + gi * G, i, j)

def decode1(batch, nt, rt, t):
    # ticket order of gpb_chol.cu: group j = [partial diag(j+1)] [fused (j+1, j) + diag(j+1)] [plain (i, j)]
    if t < batch: return (t, 0, 0)
    j = 0
    while j + 1 < nt and t >= gs(batch, rt, j + 1): j += 1
    local = t - gs(batch, rt, j)
    nA = batch if j + 1 < rt else 0; nB = batch if j + 1 < nt else 0
    if local < nB: return (local, j + 1, j + 1)
    local -= nB
    if local < nA: return (local, j + 1, j)
    local -= nA
    per = rt - j - 2
    return (local // per, j + 2 + local % per, j)

def main():
    n, B = int(sys.argv[1]), int(sys.argv[2])
    ghz = 1.965
    ctx = G.Context(0)
    rng = np.random.default_rng(3)
    x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(B)]
    for _ in range(2): ds.loglik_batch(props)
    ctx.set_trace(1)
    ds.loglik_batch(props)
    print("timing", ctx.last_timing())
    tr = ctx.get_trace().astype(np.int64)
    ctx.set_trace(0)
    npad = (n + 63) // 64 * 64; nt = npad // 64; rt = nt + 1
    T = len(tr)
    dec = np.array([decode(B, nt, rt, t) for t in range(T)])
    isdiag = dec[:, 1] == dec[:, 2]
    g0 = tr[:, 0].min()
    us = lambda a: a / (ghz * 1e3)
    fused = (dec[:, 1] == dec[:, 2] + 1) & (dec[:, 1] < nt)
    for name, m in (("diag block 0 / partial diag", isdiag), ("fused (j+1,j)+diag(j+1)", fused), ("plain offdiag", ~isdiag & ~fused)):
        t = tr[m]; j = dec[m, 2]
        has = j > 0
        print(f"--- {name}: {m.sum()} tasks")
        print("   recv->first data   us: mean %.2f  p50 %.2f  p90 %.2f" % tuple(np.percentile(us(t[has, 2] - t[has, 1]), [50, 50, 90]) if has.any() else (0, 0, 0)))
        ml = us(t[has, 3] - t[has, 2]); ch = j[has]
        print("   mainloop us per chunk: mean %.2f  p50 %.2f p90 %.2f (ideal 2.08)" % (np.mean(ml / ch), np.median(ml / ch), np.percentile(ml / ch, 90)))
        nst = (j[has] * 4).sum()
        print("   mainloop blocked on data: %.1f%% of mainloop time; %.1f%% of stages not ready at first probe, mean wait %.2f us when not ready" % (100.0 * t[has, 9].sum() / max(1, (t[has, 3] - t[has, 2]).sum()), 100.0 * t[has, 8].sum() / max(1, nst), us(t[has, 9].sum() / max(1, t[has, 8].sum()))))
        print("   mainloop end->dep ok  us: p50 %.2f  p90 %.2f  mean %.2f" % (*np.percentile(us(t[:, 4] - t[:, 3]), [50, 90]), us(t[:, 4] - t[:, 3]).mean()))
        print("   dep ok->compute done  us: p50 %.2f  p90 %.2f" % tuple(np.percentile(us(t[:, 5] - t[:, 4]), [50, 90])))
        print("   compute done->release us: p50 %.2f  p90 %.2f" % tuple(np.percentile(us(t[:, 6] - t[:, 5]), [50, 90])))
        print("   total task us: mean %.2f" % us(t[:, 6] - t[:, 1]).mean())
        print("   producer: posted->recv us p50 %.2f | tile issued - recv p50 %.2f p90 %.2f | last C stage issued -> mainloop end p50 %.2f | LT issued - mainloop end p50 %.2f p90 %.2f" % (
            np.median(us(t[:, 1] - t[:, 10])), *np.percentile(us(t[:, 11] - t[:, 1]), [50, 90]),
            np.median(us(t[has, 3] - t[has, 12])), *np.percentile(us(t[:, 13] - t[:, 3]), [50, 90])))
    # busy fraction per SM
    span = (tr[:, 7].max() - g0) / 1e3
    busy = us(tr[:, 6] - tr[:, 1]).sum()
    print(f"kernel span {span:.1f} us; sum task time {busy:.1f} us -> avg CTA busy {busy/span:.1f}")
    if B > 1:
        # per block column: when the chain reaches it (median over the proposals of the release of
        # diag(j)), and how many CTA-microseconds of tasks ran between consecutive columns
        rel = np.zeros((B, nt))
        for t in range(T):
            b, i, j = dec[t]
            if i == j == 0: rel[b, 0] = (tr[t, 7] - g0) / 1e3
            elif fused[t]: rel[b, i] = (tr[t, 7] - g0) / 1e3
        med = np.median(rel, axis=0)
        start = (tr[:, 0] - g0) / 1e3; end = (tr[:, 7] - g0) / 1e3
        print("col: diag released at (us, median over proposals) | step | busy CTAs on average in that window | of which waiting on a dependency")
        wait_us = us(tr[:, 4] - tr[:, 3])
        prev = 0.0
        for j in range(nt):
            lo, hi = prev, med[j]
            ov = np.clip(np.minimum(end, hi) - np.maximum(start, lo), 0, None).sum()
            # waiting time that falls in the window (approximate: waits are attributed to the task's end window)
            w = wait_us[(end > lo) & (end <= hi)].sum()
            print(f"  {j:3d}: {med[j]:8.1f} | {hi - lo:6.1f} | {ov / max(hi - lo, 1e-9):6.1f} | {w / max(hi - lo, 1e-9):6.1f}")
            prev = hi
        print(f"  end: {span:8.1f} | {span - prev:6.1f}")
    if B == 1:
        # critical chain: release time of diag(j) relative to start
        d = {(dec[t, 1], dec[t, 2]): t for t in range(T)}
        rel = [(tr[d[(j, j)], 7] - g0) / 1e3 for j in range(nt)]
        print("diag release times (us), every 8th:", [round(v, 1) for v in rel[::8]], "per column avg", (rel[-1] - rel[0]) / (nt - 1))
        for j in (8, 32, 56):
            if j + 1 >= nt: continue
            a = tr[d[(j + 1, j)]]; b = tr[d[(j + 1, j + 1)]]; dj = tr[d[(j, j)]]
            print(f"  col {j}: diag({j}) rel {(dj[7]-g0)/1e3:.1f} | A recv {(a[0]-g0)/1e3:.1f} rel {(a[7]-g0)/1e3:.1f} (dep->done {us(a[5]-a[4]):.1f}, done->rel {us(a[6]-a[5]):.1f}) | diag({j+1}) recv {(b[0]-g0)/1e3:.1f} factor {us(b[5]-b[4]):.1f} rel {(b[7]-g0)/1e3:.1f}")
if __name__ == "__main__":
    main()
