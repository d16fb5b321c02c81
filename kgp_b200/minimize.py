"""Host-side hyper-parameter optimiser of the engine (scalar logic only).

Same algorithm as GPML/util/minimize.m:70-170 (Polack-Ribiere CG, Wolfe-Powell
line search, constants minimize.m:42-46), which is what `GPML.train` runs through
"hyp = minimize(hyp, @gp, -n_iter, ...)" (kgp/backend/gpml.py:13-15).  Each
function evaluation is one kgp_eval(want=1) on the device.  ``f(X) -> (value,
grad)``; a raised exception / NaN / Inf during extrapolation bisects the step
(minimize.m:94-103).  Returns (X, fX list, number of function evaluations).

This file is a Python restatement of minimize.m and keeps its notice, as the
GPML licence (FreeBSD) asks:

  minimize.m -- Copyright (C) 2001 - 2010 by Carl Edward Rasmussen, 2010-01-03.
  GPML Toolbox 4.0 -- Copyright (c) 2005-2016 Carl Edward Rasmussen & Hannes
  Nickisch.  All rights reserved.  Redistribution and use in source and binary
  forms, with or without modification, are permitted provided that the
  following conditions are met: 1. Redistributions of source code must retain
  the above copyright notice, this list of conditions and the following
  disclaimer.  2. Redistributions in binary form must reproduce the above
  copyright notice, this list of conditions and the following disclaimer in the
  documentation and/or other materials provided with the distribution.
  THIS SOFTWARE IS PROVIDED BY CARL EDWARD RASMUSSEN & HANNES NICKISCH "AS IS"
  AND ANY EXPRESS OR IMPLIED WARRANTIES, INCLUDING, BUT NOT LIMITED TO, THE
  IMPLIED WARRANTIES OF MERCHANTABILITY AND FITNESS FOR A PARTICULAR PURPOSE ARE
  DISCLAIMED.  IN NO EVENT SHALL CARL EDWARD RASMUSSEN & HANNES NICKISCH OR
  CONTRIBUTORS BE LIABLE FOR ANY DIRECT, INDIRECT, INCIDENTAL, SPECIAL,
  EXEMPLARY, OR CONSEQUENTIAL DAMAGES HOWEVER CAUSED AND ON ANY THEORY OF
  LIABILITY ARISING IN ANY WAY OUT OF THE USE OF THIS SOFTWARE, EVEN IF ADVISED
  OF THE POSSIBILITY OF SUCH DAMAGE.

There is ONE copy of this logic in the repository: the tests (KAT-2 of the GPML
manual, 14.13 / 0.15) run on this file; the oracle has no optimiser of its own.
"""
import numpy as np

INT, EXT, MAX, RATIO, SIG = 0.1, 3.0, 20, 10.0, 0.1
RHO = SIG / 2
REALMIN = np.finfo(np.float64).tiny


def minimize(X, f, length, red=1.0):
    X = np.asarray(X, dtype=np.float64).copy()
    nevals = [0]

    def feval(v):
        nevals[0] += 1
        val, g = f(v)
        return float(val), np.asarray(g, dtype=np.float64)

    i = 0
    ls_failed = False
    f0, df0 = feval(X)
    fX = [f0]
    i += (length < 0)
    s = -df0; d0 = -s @ s
    x3 = red / (1.0 - d0)
    f3, df3, d3 = f0, df0, d0
    while i < abs(length):
        i += (length > 0)
        X0, F0, dF0 = X.copy(), f0, df0.copy()
        M = MAX if length > 0 else min(MAX, -length - i)
        while True:                                   # extrapolation
            x2, f2, d2, f3, df3 = 0.0, f0, d0, f0, df0
            success = False
            while not success and M > 0:
                try:
                    M -= 1; i += (length < 0)
                    f3, df3 = feval(X + x3 * s)
                    if not np.isfinite(f3) or not np.all(np.isfinite(df3)):
                        raise FloatingPointError
                    success = True
                except Exception:
                    x3 = (x2 + x3) / 2
            if f3 < F0:
                X0, F0, dF0 = X + x3 * s, f3, df3
            d3 = df3 @ s
            if d3 > SIG * d0 or f3 > f0 + x3 * RHO * d0 or M == 0:
                break
            x1, f1, d1 = x2, f2, d2
            x2, f2, d2 = x3, f3, d3
            A = 6 * (f1 - f2) + 3 * (d2 + d1) * (x2 - x1)
            B = 3 * (f2 - f1) - (2 * d1 + d2) * (x2 - x1)
            disc = B * B - A * d1 * (x2 - x1)
            with np.errstate(all='ignore'):
                x3 = x1 - d1 * (x2 - x1) ** 2 / (B + np.sqrt(disc)) if disc >= 0 else np.nan
            if not np.isfinite(x3) or x3 < 0:
                x3 = x2 * EXT
            elif x3 > x2 * EXT:
                x3 = x2 * EXT
            elif x3 < x2 + INT * (x2 - x1):
                x3 = x2 + INT * (x2 - x1)
        x4 = f4 = d4 = None
        while (abs(d3) > -SIG * d0 or f3 > f0 + x3 * RHO * d0) and M > 0:   # interpolation
            if d3 > 0 or f3 > f0 + x3 * RHO * d0:
                x4, f4, d4 = x3, f3, d3
            else:
                x2, f2, d2 = x3, f3, d3
            with np.errstate(all='ignore'):
                if f4 > f0:
                    x3 = x2 - (0.5 * d2 * (x4 - x2) ** 2) / (f4 - f2 - d2 * (x4 - x2))
                else:
                    A = 6 * (f2 - f4) / (x4 - x2) + 3 * (d4 + d2)
                    B = 3 * (f4 - f2) - (2 * d2 + d4) * (x4 - x2)
                    disc = B * B - A * d2 * (x4 - x2) ** 2
                    x3 = x2 + (np.sqrt(disc) - B) / A if disc >= 0 else np.nan
            if not np.isfinite(x3):
                x3 = (x2 + x4) / 2
            x3 = max(min(x3, x4 - INT * (x4 - x2)), x2 + INT * (x4 - x2))
            f3, df3 = feval(X + x3 * s)
            if f3 < F0:
                X0, F0, dF0 = X + x3 * s, f3, df3
            M -= 1; i += (length < 0)
            d3 = df3 @ s
        if abs(d3) < -SIG * d0 and f3 < f0 + x3 * RHO * d0:       # line search succeeded
            X = X + x3 * s; f0 = f3; fX.append(f0)
            s = (df3 @ df3 - df0 @ df3) / (df0 @ df0) * s - df3   # Polack-Ribiere
            df0 = df3
            d3 = d0; d0 = df0 @ s
            if d0 > 0:
                s = -df0; d0 = -s @ s
            x3 = x3 * min(RATIO, d3 / (d0 - REALMIN))
            ls_failed = False
        else:
            X, f0, df0 = X0, F0, dF0
            if ls_failed or i > abs(length):
                break
            s = -df0; d0 = -s @ s
            x3 = 1.0 / (1.0 - d0)
            ls_failed = True
    return X, fX, nevals[0]
