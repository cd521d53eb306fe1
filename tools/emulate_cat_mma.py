"""Lane-level CPU emulation of csrc/glm_categorical_mma.cu (no GPU): the tile layout of data.cu
(kLayoutMma), the fragment addressing of both DMMA phases, the softmax thread mapping and the CTA
reduction, with the same index expressions as the kernel, checked against numpy.  It validates the
ADDRESS ARITHMETIC of the kernel (which lane reads which word), not the CUDA code itself; the GPU
parity tests do that.  Used by tests/test_layout_cpu.py."""
import numpy as np


def tile_elem_offset(r, col, kpad):
    return ((r >> 3) * (kpad >> 2) + (col >> 2)) * 32 + (r & 7) * 4 + (col & 3)


def retile(x, R, kpad):
    n, k = x.shape
    tiles = (n + R - 1) // R
    dev = np.zeros(tiles * R * kpad)
    for i in range(n):
        t, r = divmod(i, R)
        for col in range(k):
            dev[t * R * kpad + tile_elem_offset(r, col, kpad)] = x[i, col]
    return dev, tiles


def dmma(c, a, b):
    """c: (32, 2) accumulator fragment, a, b: (32,) operand fragments of one warp -> new c."""
    lane = np.arange(32)
    A = np.zeros((8, 4))
    B = np.zeros((4, 8))
    A[lane >> 2, lane & 3] = a
    B[lane & 3, lane >> 2] = b
    D = A @ B
    out = c.copy()
    out[:, 0] += D[lane >> 2, 2 * (lane & 3)]
    out[:, 1] += D[lane >> 2, 2 * (lane & 3) + 1]
    return out


def emulate(x, y, alpha, beta):
    """x (n, k), y (n,) in 1..C, alpha (C,), beta (k, C) -> lp, dalpha (C,), dbeta (k, C)."""
    n, k = x.shape
    ncls = len(alpha)
    kpad = (k + 7) // 8 * 8
    R = 64 if kpad <= 64 else 32
    NTL = 1 if ncls <= 8 else 2
    ldl = 8 if NTL == 1 else 24
    kb, MT, RB = kpad >> 2, kpad >> 3, R >> 3
    KS, TPR = 8 // RB, 256 // R
    MPW = (MT + 7) // 8
    xdev, tiles = retile(x, R, kpad)
    ydev = np.zeros(tiles * R, dtype=np.int64)
    ydev[:n] = y
    params = np.zeros(ncls * kpad + ncls)
    for c in range(ncls):
        params[c * kpad:c * kpad + k] = beta[:, c]
    params[ncls * kpad:] = alpha
    bfrag = np.zeros(kb * NTL * 32 + 16)
    for q in range(ncls * kpad):
        c, kk = divmod(q, kpad)
        bfrag[((kk >> 2) * NTL + (c >> 3)) * 32 + (c & 7) * 4 + (kk & 3)] = params[q]
    alpha_s = params[ncls * kpad:]
    lane = np.arange(32)
    g = np.zeros((8, MPW, NTL, 32, 2))
    da = np.zeros((256, 4))
    lp_acc = np.zeros(256)
    dfrag = np.zeros((R // 4) * NTL * 32)
    for tile in range(tiles):
        xt = xdev[tile * R * kpad:(tile + 1) * R * kpad]
        ys = ydev[tile * R:(tile + 1) * R]
        linp = np.zeros(64 * ldl)
        for warp in range(8):
            rbk, ks = warp % RB, warp // RB
            kbeg = ks * (kb // KS)
            kend = kb if ks == KS - 1 else kbeg + kb // KS
            c = np.zeros((NTL, 32, 2))
            for kblk in range(kbeg, kend):
                a = xt[rbk * kb * 32 + lane + kblk * 32]
                for j in range(NTL):
                    c[j] = dmma(c[j], a, bfrag[(kblk * NTL + j) * 32 + lane])
            for j in range(NTL):
                base = (ks * R + rbk * 8 + (lane >> 2)) * ldl + 2 * (lane & 3) + j * 8
                linp[base] = c[j][:, 0]
                linp[base + 1] = c[j][:, 1]
        for tid in range(256):
            srow, sq = tid // TPR, tid % TPR
            row = tile * R + srow
            valid = row < n
            yc = ys[srow] - 1 if valid else -1
            # the TPR threads of a row exchange max / sum / ly by shuffle: emulate the group at once
            if sq != 0:
                continue
            v = np.full(ncls, 0.0)
            for c_ in range(ncls):
                t = linp[srow * ldl + c_]
                if KS == 2:
                    t += linp[(R + srow) * ldl + c_]
                v[c_] = t + alpha_s[c_]
            mx = v.max()
            e = np.exp(v - mx)
            inv = 1.0 / e.sum()
            if valid:
                lp_acc[tid] += np.log(inv) - mx + v[yc]
            for c_ in range(ncls):
                q, i = c_ % TPR, c_ // TPR
                dv = ((1.0 if c_ == yc else 0.0) - e[c_] * inv) if valid else 0.0
                da[srow * TPR + q, i] += dv
                dfrag[((srow >> 2) * NTL + (c_ >> 3)) * 32 + (c_ & 7) * 4 + (srow & 3)] = dv
        for warp in range(8):
            if MT >= 8:
                mt0, rs0, rs1, active = warp, 0, R // 4, True
            else:
                RS = 8 // MT
                per = (R // 4) // RS
                rpart = warp // MT
                mt0 = warp % MT
                active = rpart < RS
                rs0 = rpart * per
                rs1 = R // 4 if rpart == RS - 1 else rs0 + per
            if not active:
                continue
            # pointer walk of the kernel: two row steps (one row block of 8 rows) per iteration
            xr = (rs0 >> 1) * kb * 32 + mt0 * 64 + (lane >> 4) * 32 + (lane & 3) * 4 + ((lane >> 2) & 3)
            df = rs0 * NTL * 32 + lane
            for rs in range(rs0, rs1, 2):
                b0 = [dfrag[df + j * 32] for j in range(NTL)]
                b1 = [dfrag[df + (NTL + j) * 32] for j in range(NTL)]
                for m in range(MPW):
                    if m == 0 or mt0 + 8 * m < MT:
                        a0, a1 = xt[xr + m * 512], xt[xr + m * 512 + 16]
                        for j in range(NTL):
                            g[warp, m, j] = dmma(g[warp, m, j], a0, b0[j])
                            g[warp, m, j] = dmma(g[warp, m, j], a1, b1[j])
                xr = xr + kb * 32
                df = df + 2 * NTL * 32
    # CTA reduction
    RSn = 1 if MT >= 8 else 8 // MT
    rg = np.zeros(RSn * kpad * 8 * NTL)
    for warp in range(8):
        if MT >= 8:
            mt0, rpart, active = warp, 0, True
        else:
            mt0, rpart = warp % MT, warp // MT
            active = rpart < 8 // MT
        if not active:
            continue
        for m in range(MPW):
            mt = mt0 + 8 * m
            if mt < MT:
                for j in range(NTL):
                    dst = (rpart * kpad + mt * 8 + (lane >> 2)) * (8 * NTL) + j * 8 + 2 * (lane & 3)
                    rg[dst] = g[warp, m, j][:, 0]
                    rg[dst + 1] = g[warp, m, j][:, 1]
    dbeta = np.zeros((k, ncls))
    for idx in range(ncls * kpad):
        c, kk = divmod(idx, kpad)
        s = sum(rg[(q * kpad + kk) * (8 * NTL) + c] for q in range(RSn))
        if kk < k:
            dbeta[kk, c] = s
    dalpha = np.zeros(ncls)
    part = da.reshape(-1)
    for c in range(ncls):
        q, i = c % TPR, c // TPR
        dalpha[c] = sum(part[(r * TPR + q) * 4 + i] for r in range(R))
    return lp_acc.sum(), dalpha, dbeta


def reference(x, y, alpha, beta):
    lin = x @ beta + alpha
    mx = lin.max(axis=1, keepdims=True)
    e = np.exp(lin - mx)
    sm = e / e.sum(axis=1, keepdims=True)
    lp = float(np.sum(np.log(1.0 / e.sum(axis=1)) - mx[:, 0] + lin[np.arange(len(y)), y - 1]))
    d = -sm
    d[np.arange(len(y)), y - 1] += 1.0
    return lp, d.sum(axis=0), x.T @ d


if __name__ == "__main__":
    rng = np.random.default_rng(0)
    for n, k, c in [(70, 5, 3), (200, 32, 16), (130, 24, 9), (100, 72, 4), (90, 128, 16), (70, 250, 10), (65, 64, 8)]:
        x = rng.standard_normal((n, k))
        beta = rng.standard_normal((k, c)) * 0.3
        alpha = rng.standard_normal(c)
        y = rng.integers(1, c + 1, n)
        got, ref = emulate(x, y, alpha, beta), reference(x, y, alpha, beta)
        err = max(abs(got[0] - ref[0]) / abs(ref[0]), np.abs(got[1] - ref[1]).max(), np.abs(got[2] - ref[2]).max())
        print(n, k, c, "max err %.2e" % err)
        assert err < 1e-10
