"""Host-side emulation of the index arithmetic of `freq_split3d_mma_kernel` (hivae_b200/csrc/signal.cu).

The kernel's correctness rests on four mappings that are easy to get wrong and impossible to see in a diff: the m16n8k8
fragment layouts (which lane holds which element of A, B and C), the n-permutation of the T pass (tile nt, index g <-> column
4 g + nt), the k-permutation of the W pass (k index t <-> w = 2t, t + 4 <-> w = 2t + 1) and the XOR-swizzled slab addressing.
This script walks the same loops as the kernel, warp by warp and lane by lane, with every `mma.sync` replaced by an exact
matrix product placed through the PTX fragment layout, and compares the result with three plain einsums.  It needs no GPU:
`python tools/emulate_freq_mma.py`, and tests/test_cpu_oracle_and_host.py runs it for every template instance.

What it does not model: the TF32 rounding (products are exact here), timing, and the cp.async staging (a copy).
"""
import numpy as np


def mma_m16n8k8(acc, A, B):
    """acc[lane][0..3] += A . B with the operands given per lane in the PTX fragment layout of mma.m16n8k8 (tf32):
    A: a0 (row g, col t), a1 (g + 8, t), a2 (g, t + 4), a3 (g + 8, t + 4); B: b0 (k t, n g), b1 (t + 4, g);
    C: c0 (g, 2t), c1 (g, 2t + 1), c2 (g + 8, 2t), c3 (g + 8, 2t + 1); g = lane >> 2, t = lane & 3."""
    Am, Bm = np.zeros((16, 8)), np.zeros((8, 8))
    for lane in range(32):
        g, t = lane >> 2, lane & 3
        a, b = A[lane], B[lane]
        Am[g, t], Am[g + 8, t], Am[g, t + 4], Am[g + 8, t + 4] = a
        Bm[t, g], Bm[t + 4, g] = b
    D = Am @ Bm
    for lane in range(32):
        g, t = lane >> 2, lane & 3
        acc[lane][0] += D[g, 2 * t]
        acc[lane][1] += D[g, 2 * t + 1]
        acc[lane][2] += D[g + 8, 2 * t]
        acc[lane][3] += D[g + 8, 2 * t + 1]


def zero_acc():
    return {lane: [0.0] * 4 for lane in range(32)}


def emulate(T, N, vols=2, seed=0):
    """Runs the emulated kernel on random (non-symmetric, so transposes would show) matrices; returns the relative error of
    low against the einsum reference, the number of unwritten outputs, and max |low + high - x|."""
    rng = np.random.default_rng(seed)
    TC = 16 if T >= 16 else 8
    SLABS, HW = T // TC, N * N
    FP = (N * 32 if N == 32 else N * 24) + 4
    idx = (lambda h, w: h * 32 + (w ^ ((h & 3) << 3))) if N == 32 else (lambda h, w: h * 24 + w)   # slab_idx<N>
    KT, KN, MT = T // 8, N // 8, N // 16
    x = rng.standard_normal(vols * T * HW)
    ct, ch, cw = (rng.standard_normal(n * n) for n in (T, N, N))
    lo, hi = np.full_like(x, np.nan), np.full_like(x, np.nan)
    for blk in range(vols * SLABS):
        v, t0 = blk // SLABS, (blk % SLABS) * TC
        xv = x[v * T * HW:]
        S = np.full(TC * FP, np.nan)
        # ---- T pass: warp -> 32-column groups, four n-permuted tiles per group
        for warp in range(8):
            A = []
            for ks in range(KT):
                frag = {}
                for lane in range(32):
                    g, t = lane >> 2, lane & 3
                    r0 = (t0 + g) * T + 8 * ks + t
                    r1 = (t0 + g + 8) * T + 8 * ks + t
                    frag[lane] = [ct[r0], ct[r1] if TC == 16 else 0.0, ct[r0 + 4], ct[r1 + 4] if TC == 16 else 0.0]
                A.append(frag)
            for j in range(warp, HW // 32, 8):
                accs = []
                for nt in range(4):
                    acc = zero_acc()
                    for ks in range(KT):
                        B = {}
                        for lane in range(32):
                            g, t = lane >> 2, lane & 3
                            src = (8 * ks + t) * HW + 32 * j + 4 * g
                            B[lane] = [xv[src + nt], xv[src + 4 * HW + nt]]
                        mma_m16n8k8(acc, A[ks], B)
                    accs.append(acc)
                for lane in range(32):
                    g, t = lane >> 2, lane & 3
                    col = 32 * j + 8 * t
                    d = g * FP + idx(col // N, col % N)
                    for e in range(4):
                        S[d + e], S[d + 4 + e] = accs[e][lane][0], accs[e][lane][1]
                        if TC == 16:
                            S[d + 8 * FP + e], S[d + 8 * FP + 4 + e] = accs[e][lane][2], accs[e][lane][3]
        # ---- H pass, in place: warp -> (frame, 8-column block), all row tiles
        for warp in range(8):
            for it in range(warp, TC * KN, 8):
                fr, w8 = (it // KN) * FP, 8 * (it % KN)
                Bf = [{lane: [S[fr + idx(8 * ks + (lane & 3), w8 + (lane >> 2))],
                              S[fr + idx(8 * ks + (lane & 3) + 4, w8 + (lane >> 2))]] for lane in range(32)} for ks in range(KN)]
                outs = []
                for mt in range(MT):
                    acc = zero_acc()
                    for ks in range(KN):
                        A = {}
                        for lane in range(32):
                            g, t = lane >> 2, lane & 3
                            r0 = (16 * mt + g) * N + 8 * ks + t
                            A[lane] = [ch[r0], ch[r0 + 8 * N], ch[r0 + 4], ch[r0 + 8 * N + 4]]
                        mma_m16n8k8(acc, A, Bf[ks])
                    outs.append(acc)
                for mt in range(MT):
                    for lane in range(32):
                        g, t = lane >> 2, lane & 3
                        d, d2 = fr + idx(16 * mt + g, w8 + 2 * t), fr + idx(16 * mt + g + 8, w8 + 2 * t)
                        S[d], S[d + 1] = outs[mt][lane][0], outs[mt][lane][1]
                        S[d2], S[d2 + 1] = outs[mt][lane][2], outs[mt][lane][3]
        # ---- W pass, in place: warp -> (frame, 16-row block), all n-tiles; k-step index t <-> w = 2t, t + 4 <-> w = 2t + 1
        for warp in range(8):
            for it in range(warp, TC * MT, 8):
                fr = (it // MT) * FP
                Af = []
                for ks in range(KN):
                    frag = {}
                    for lane in range(32):
                        g, t = lane >> 2, lane & 3
                        r0 = 16 * (it % MT) + g
                        p, q = fr + idx(r0, 8 * ks + 2 * t), fr + idx(r0 + 8, 8 * ks + 2 * t)
                        frag[lane] = [S[p], S[q], S[p + 1], S[q + 1]]
                    Af.append(frag)
                outs = []
                for nt in range(KN):
                    acc = zero_acc()
                    for ks in range(KN):
                        B = {}
                        for lane in range(32):
                            g, t = lane >> 2, lane & 3
                            r = (8 * nt + g) * N + 8 * ks + 2 * t
                            B[lane] = [cw[r], cw[r + 1]]
                        mma_m16n8k8(acc, Af[ks], B)
                    outs.append(acc)
                for nt in range(KN):
                    for lane in range(32):
                        g, t = lane >> 2, lane & 3
                        r0, w = 16 * (it % MT) + g, 8 * nt + 2 * t
                        d, d2 = fr + idx(r0, w), fr + idx(r0 + 8, w)
                        S[d], S[d + 1] = outs[nt][lane][0], outs[nt][lane][1]
                        S[d2], S[d2 + 1] = outs[nt][lane][2], outs[nt][lane][3]
        # ---- epilogue: float4 i of the slab <-> frame 4i / HW, element 4i % HW
        off = v * T * HW + t0 * HW
        for i in range(TC * HW // 4):
            e = 4 * i
            a = (e // HW) * FP + idx((e % HW) // N, (e % HW) % N)
            for c in range(4):
                lo[off + e + c] = S[a + c]
                hi[off + e + c] = x[off + e + c] - S[a + c]
    X = x.reshape(vols, T, N, N)
    ref = np.einsum('at,vthw->vahw', ct.reshape(T, T), X)
    ref = np.einsum('bh,vahw->vabw', ch.reshape(N, N), ref)
    ref = np.einsum('cw,vabw->vabc', cw.reshape(N, N), ref)
    unwritten = int(np.isnan(lo).sum() + np.isnan(hi).sum())
    err = float(np.nanmax(np.abs(lo.reshape(ref.shape) - ref)) / np.abs(ref).max())
    return err, unwritten, float(np.nanmax(np.abs(lo + hi - x)))


INSTANCES = ((32, 32), (16, 32), (8, 32), (32, 16), (16, 16), (8, 16))   # the template instances hv_freq_split3d dispatches to

if __name__ == "__main__":
    for T, N in INSTANCES:
        err, unwritten, rec = emulate(T, N)
        print("T=%2d N=%2d  low vs einsum %.1e  unwritten %d  |low + high - x| %.1e" % (T, N, err, unwritten, rec))
