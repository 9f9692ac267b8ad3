"""TEST INFRASTRUCTURE ONLY -- the float64 pre-pass of logistic regression in numpy.

Restates lib/encoding/logistic_regression.go of the reference (which IS under /root/reference, so this row is pinned
to the reference's own source and known answers, logistic_regression_test.go:20-72):
  StandardiseWith (:964-978)          x' = (x - mean) / sd per feature
  Augment (:1014-1023)                X = [1.0, x'_1 .. x'_d]
  ComputeAllApproxCoefficients (:367-403)
      level 1:  a1[s]   = X[s] * (2*y - 1)
      level j:  aj[ri]  = ypart * (((1.0 * X[i1]) * X[i2]) ...), tuples (i1..ij) in lexicographic order,
                ypart = y - y*(-1)^j - 1
  AggregateApproxCoefficients (:433-455)   record-by-record sums in index order
  Float64ToInt642DArrayWithPrecision (:1558-1574)   int64(math.Round(x * precision))
  packing (:306-313, :46-118)          level j at offset j * (d+1)^j  (consistent for k <= 2, the only case used)
Every operation is an IEEE-754 double operation rounded to nearest with no fused multiply-add; np.add.accumulate adds
strictly in order, so the summation order is the reference's. Only tests/ may import this."""
import numpy as np


def coefficients(x, y, means, sds, k, precision):
    """x: float64[n][d], y: int64[n] -> packed int64 coefficients of one data provider."""
    y = np.asarray(y, dtype=np.int64)
    n, d = len(y), len(means)
    x = np.asarray(x, dtype=np.float64).reshape(n, d)
    assert 1 <= k <= 2
    ncoef = sum((d + 1) ** (j + 1) for j in range(k))
    out = np.zeros(ncoef, dtype=np.int64)
    if n == 0:
        return out
    X = np.concatenate([np.ones((n, 1)), (x - np.asarray(means, np.float64)) / np.asarray(sds, np.float64)], axis=1)
    yf = y.astype(np.float64)
    levels = [X * (2 * yf - 1)[:, None]]
    if k >= 2:
        ypart = (y - y * 1 - 1).astype(np.float64)  # (-1)^2 = 1
        xp = (1.0 * X[:, :, None]) * X[:, None, :]
        levels.append((ypart[:, None, None] * xp).reshape(n, -1))
    for j, terms in enumerate(levels):
        acc = np.add.accumulate(terms, axis=0)[-1] if n > 1 else terms[0]
        acc = 0.0 + acc  # the reference starts every sum from 0.0
        vals = acc * precision
        t = np.trunc(vals)  # math.Round: to nearest, halves away from zero (vals - t is exact)
        exact = np.where(np.abs(vals - t) >= 0.5, t + np.sign(vals), t)
        off = j * (d + 1) ** j
        out[off:off + exact.size] = exact.astype(np.int64)
    return out


def coefficients_loop(x, y, means, sds, k, precision):
    """The same with plain Python floats, one operation at a time (small inputs; cross-check of the numpy form)."""
    import math
    n = len(y)
    d = len(means)
    ncoef = sum((d + 1) ** (j + 1) for j in range(k))
    acc = [[0.0] * ((d + 1) ** (j + 1)) for j in range(k)]
    for r in range(n):
        X = [1.0] + [(float(x[r][i]) - means[i]) / sds[i] for i in range(d)]
        yy = int(y[r])
        for s in range(d + 1):
            acc[0][s] += X[s] * (2 * float(yy) - 1)
        if k >= 2:
            ypart = float(yy - yy * 1 - 1)
            ri = 0
            for i1 in range(d + 1):
                for i2 in range(d + 1):
                    xp = 1.0
                    xp *= X[i1]
                    xp *= X[i2]
                    acc[1][ri] += ypart * xp
                    ri += 1
    out = [0] * ncoef
    for j in range(k):
        off = j * (d + 1) ** j
        for i, v in enumerate(acc[j]):
            w = v * precision
            t = float(math.trunc(w))
            out[off + i] = int(t + math.copysign(1.0, w)) if abs(w - t) >= 0.5 else int(t)
    return np.array(out, dtype=np.int64)
