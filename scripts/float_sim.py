"""Sizing tool (not a solver): a float64 dense revised simplex with the reference's pivot rules
(FirstProfitableWithMemory in phase 1, SteepestDescent in phase 2, Bland ratio ties) on a golden
fixture, to ESTIMATE how many pivots an exact solve needs and how large det(B) of the row-scaled
integer basis gets (log2 |det| = sum of log2 |pivot| + log2 of the row scales).  Floating point
decides ties differently, so only the order of magnitude means anything.

    python scripts/float_sim.py tests/golden/mps_prefix/netlib_25FV47.npz [max_pivots]
"""
import math
import sys
import time
from math import lcm

import numpy as np


def main():
    z = np.load(sys.argv[1])
    limit = int(sys.argv[2]) if len(sys.argv) > 2 else 10 ** 6
    m, n = int(z["m"]), int(z["n"])
    cp, ri = z["col_ptr"], z["row_idx"]
    vn, vd = z["val_num"], z["val_den"]
    A = np.zeros((m, n))
    rowscale = [1] * m
    for j in range(n):
        for t in range(cp[j], cp[j + 1]):
            A[ri[t], j] = vn[t] / vd[t]
            rowscale[ri[t]] = lcm(rowscale[ri[t]], int(vd[t]))
    b = z["rhs_num"] / z["rhs_den"]
    for i in range(m):
        rowscale[i] = lcm(rowscale[i], int(z["rhs_den"][i]))
    c = z["cost_num"] / z["cost_den"]
    real = dict(zip(z["piv_rows"].tolist(), z["piv_cols"].tolist()))
    art_rows = [i for i in range(m) if i not in real]
    na = len(art_rows)
    # tableau columns: artificial 0..na-1, then provider columns
    basis = [art_rows.index(i) if i not in real else na + real[i] for i in range(m)]
    Binv = np.eye(m)
    xb = b.copy()
    log2scale = sum(math.log2(s) for s in rowscale)
    print(f"m={m} n={n} artificial={na} sum log2(row scale)={log2scale:.0f} bits, max row scale={max(rowscale)}")
    log2det = 0.0
    maxdet = 0.0
    tol = 1e-9
    t0 = time.time()
    piv = 0
    for phase in (1, 2):
        if phase == 1:
            cost = np.concatenate([np.ones(na), np.zeros(n)])
        else:
            cost = np.concatenate([np.full(na, np.inf), c])
        last = None
        while piv < limit:
            cb = np.array([cost[j] for j in basis])
            pi = cb @ Binv
            rc_struct = (cost[na:] - pi @ A)
            rc = np.concatenate([cost[:na] - pi[art_rows] if phase == 1 else np.full(na, np.inf), rc_struct])
            rc[basis] = 0.0
            neg = np.nonzero(rc < -tol)[0]
            if len(neg) == 0:
                break
            if phase == 1:
                if last is None:
                    j = neg[0]
                else:
                    after = neg[neg >= last]
                    j = after[0] if len(after) else neg[0]
                last = j
            else:
                j = neg[np.argmin(rc[neg])]
            col = Binv[:, art_rows[j]] if j < na else Binv @ A[:, j - na]
            pos = np.nonzero(col > tol)[0]
            if len(pos) == 0:
                print("unbounded"); return
            ratios = xb[pos] / col[pos]
            rmin = ratios.min()
            ties = pos[ratios <= rmin + 1e-12 * max(1.0, abs(rmin))]
            r = min(ties, key=lambda i: basis[i])
            p = col[r]
            log2det += math.log2(abs(p))
            maxdet = max(maxdet, log2det)
            Binv[r] /= p
            xb[r] /= p
            f = col.copy(); f[r] = 0
            Binv -= np.outer(f, Binv[r])
            xb -= f * xb[r]
            xb[np.abs(xb) < 1e-11] = 0
            basis[r] = j
            piv += 1
            if piv % 2000 == 0:
                nnz = np.count_nonzero(np.abs(Binv) > 1e-12)
                print(f"  phase {phase} pivot {piv}: log2|det B| = {log2det:.0f} (+{log2scale:.0f} scale), nnz(B^-1) = {nnz} "
                      f"({100.0 * nnz / m / m:.1f} %), {time.time() - t0:.0f} s", flush=True)
        obj = sum(cost[j] * x for j, x in zip(basis, xb) if np.isfinite(cost[j]))
        print(f"phase {phase} done after {piv} pivots, objective {obj:.6f}, log2|det B| = {log2det:.0f}, max {maxdet:.0f}")
        if phase == 1 and abs(obj) > 1e-6:
            print("infeasible (float)"); return


if __name__ == "__main__":
    main()
