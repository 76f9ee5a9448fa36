"""numpy restatement of ilqr_planner/ilqr_planner/src/utils/primitives.cpp:13-96 (test infrastructure).

Used to generate the goldens and to check the product's `ilqrb_build_psi`.
"""
import math

import numpy as np


def _binom(n, k):  # primitives.cpp:13-17
    if k == 0 or k == n:
        return 1
    return _binom(n - 1, k - 1) + _binom(n - 1, k)


def build_psi(kind, dim, K):
    dim, K = int(dim), int(K)
    Ts = np.arange(dim, dtype=np.float64)
    psi = np.zeros((dim, K))
    if kind == "rbf":  # :19-33
        bw = dim / K
        avg = bw / 2
        sig = bw
        for i in range(K):
            psi[:, i] = 1 / (2 * math.pi * sig) * np.exp(-1 * (Ts - avg) * (Ts - avg) / (2 * sig * sig))
            avg += bw
    elif kind == "bernstein":  # :35-50
        order = K - 1
        T = Ts / Ts.max()
        for i in range(K):
            b = _binom(order, i)
            for j in range(dim):
                psi[j, i] = b * math.pow(T[j], i) * math.pow(1 - T[j], order - i)
    elif kind == "unitstep":  # :52-69   (C round(): half away from zero)
        bw = int(math.floor(dim / K + 0.5))
        low, high = 0, bw
        for i in range(K):
            for j in range(dim):
                psi[j, i] = 1.0 / bw if (j >= low and j < high) else 0
            low += bw
            high += bw
    elif kind == "sawtooth":  # :71-88
        bw = int(math.ceil(dim / K))
        low, high = 0.0, float(bw)
        for i in range(K):
            for j in range(dim):
                psi[j, i] = ((j - low) / (bw - 1) - 0.5) if (j >= low and j < high) else 0
            low += bw
            high += bw
    elif kind == "linear":  # :90-96
        return np.hstack([build_psi("unitstep", dim, K), build_psi("sawtooth", dim, K)])
    else:
        raise ValueError(kind)
    return psi
