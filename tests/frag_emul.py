"""Lane-level numpy emulation of the decoder kernel's data movement (csrc/nsb_decoder.cu).

It executes mlp_forward / the backward chain exactly as a warp does -- A/B/C fragments of
mma.m16n8k8 per lane, the packed blob read at the same offsets with the same per-lane
indexing -- but with exact fp64 products instead of tensor-core arithmetic.  Run on the CPU,
it proves that the pack tables (nsb_api.cu), the slot/column maps (nsb_layout.h) and the
register chaining of the kernel compute the reference MLP, without a GPU.
Test infrastructure only.
"""
import ctypes as C

import numpy as np

from nice_slam_annotation_b200 import _lib

NAMES = ["B", "BIAS", "BC", "BO", "L0", "L1", "L2", "L3", "L4", "C0", "C1", "C2", "C3", "C4", "OUT", "TOUT",
         "TL0", "TL1", "TL2", "TL3", "TL4", "TL3X", "TC0", "TC1", "TC2", "TC3", "TC4", "FWD_END", "TOTAL",
         "CS", "NOUT", "EMBED"]
LANES = np.arange(32)
G, T = LANES >> 2, LANES & 3


def layout(kind):
    o = np.zeros(32, dtype=np.int32)
    _lib.check(_lib.lib().nsb_layout_offsets(kind, o.ctypes.data_as(C.c_void_p)))
    return dict(zip(NAMES, o.tolist()))


def pack(kind, params):
    """params: list of numpy arrays in canonical order -> packed blob (float64)."""
    L = _lib.lib()
    n = L.nsb_packed_numel(kind)
    tab = np.empty(n, dtype=np.int32)
    _lib.check(L.nsb_pack_table_host(kind, tab.ctypes.data_as(C.c_void_p)))
    flat = [np.asarray(p, dtype=np.float64).reshape(-1) for p in params]
    out = np.zeros(n)
    for i, e in enumerate(tab):
        if e >= 0:
            out[i] = flat[e >> 20][e & 0xfffff]
    return out


def mma(acc, a, b0, b1):
    """acc[32,4] += A*B with A frag a[32,4], B frag (b0,b1)[32]."""
    A = np.zeros((16, 8))
    B = np.zeros((8, 8))
    A[G, T], A[G + 8, T], A[G, T + 4], A[G + 8, T + 4] = a[:, 0], a[:, 1], a[:, 2], a[:, 3]
    B[T, G], B[T + 4, G] = b0, b1
    D = A @ B
    acc[:, 0] += D[G, 2 * T]
    acc[:, 1] += D[G, 2 * T + 1]
    acc[:, 2] += D[G + 8, 2 * T]
    acc[:, 3] += D[G + 8, 2 * T + 1]


def gemm(acc, a, blob, off, J, NT):
    """acc [NT][32,4], a [J][32,4]."""
    for j in range(J):
        if NT == 1:
            b = blob[off + (j * 32 + LANES) * 2][:, None] * 0
            b0 = blob[off + (j * 32 + LANES) * 2]
            b1 = blob[off + (j * 32 + LANES) * 2 + 1]
            mma(acc[0], a[j], b0, b1)
            continue
        for h in range(NT // 2):
            base = off + ((j * (NT // 2) + h) * 32 + LANES) * 4
            mma(acc[2 * h], a[j], blob[base], blob[base + 1])
            mma(acc[2 * h + 1], a[j], blob[base + 2], blob[base + 3])


def chain(c):
    return [np.stack([x[:, 0], x[:, 2], x[:, 1], x[:, 3]], 1) for x in c]


def bias_frag(blob, off):
    acc = []
    for n in range(4):
        v0, v1 = blob[off + 8 * n + 2 * T], blob[off + 8 * n + 2 * T + 1]
        acc.append(np.stack([v0, v1, v0, v1], 1).copy())
    return acc


def zeros(nt):
    return [np.zeros((32, 4)) for _ in range(nt)]


def a_from_rows(x, slot):
    """x [16, F] -> A fragments for slabs given slot(j, tau) -> feature index (or -1)."""
    def frag(j):
        out = np.zeros((32, 4))
        for q, (rows, tau) in enumerate(((G, T), (G + 8, T), (G, T + 4), (G + 8, T + 4))):
            f = np.array([slot(j, int(t)) for t in tau])
            ok = f >= 0
            out[ok, q] = x[rows[ok], f[ok]]
        return out
    return frag


def slot_natural(j, tau):
    return 8 * j + tau


def slot_gather(j, tau):
    return 32 * (j >> 2) + 8 * (tau & 3) + 2 * (j & 3) + (tau >> 2)


def c_to_rows(c, colmap, ncols):
    """C fragments [NT][32,4] -> dense [16, ncols] via colmap(n, gamma)."""
    out = np.zeros((16, ncols))
    for n, x in enumerate(c):
        for q, (rows, gam) in enumerate(((G, 2 * T), (G, 2 * T + 1), (G + 8, 2 * T), (G + 8, 2 * T + 1))):
            f = np.array([colmap(n, int(g)) for g in gam])
            ok = (f >= 0) & (f < ncols)
            out[rows[ok], f[ok]] = x[ok, q]
    return out


def col_natural(n, g):
    return 8 * n + g


def col_gather(n, g):
    return 8 * (g >> 1) + 2 * n + (g & 1)


def forward(kind, blob, e, c):
    """e [16,96] (or None for coarse), c [16, 32|64] -> out [16, 8] padded, relu masks."""
    lo = layout(kind)
    CS, embed = lo["CS"], bool(lo["EMBED"])
    cf = [a_from_rows(c, slot_gather)(j) for j in range(CS)]
    masks = []

    def relu(acc):
        masks.append([x > 0 for x in acc])
        return [np.where(x > 0, x, 0.0) for x in acc]

    def addb(acc, off):
        b = bias_frag(blob, off)
        return [x + y for x, y in zip(acc, b)]

    if embed:
        ef = [a_from_rows(e, lambda j, t: 8 * j + t)(j) for j in range(12)]
        h = None
        for i in range(5):
            acc = bias_frag(blob, lo["BIAS"] + 32 * i)
            if i == 0:
                gemm(acc, ef, blob, lo["L0"], 12, 4)
            elif i == 3:
                gemm(acc, ef, blob, lo["L3"], 12, 4)
                gemm(acc, h, blob, lo["L3"] + 12 * 256, 4, 4)
            else:
                gemm(acc, h, blob, lo["L%d" % i], 4, 4)
            acc = relu(acc)
            acc = addb(acc, lo["BC"] + 32 * i)
            gemm(acc, cf, blob, lo["C%d" % i], CS, 4)
            h = chain(acc)
    else:
        h = None
        for i in range(5):
            acc = bias_frag(blob, lo["BIAS"] + 32 * i)
            if i == 0:
                gemm(acc, cf, blob, lo["L0"], 4, 4)
            elif i == 3:
                gemm(acc, cf, blob, lo["L3"], 4, 4)
                gemm(acc, h, blob, lo["L3"] + 4 * 256, 4, 4)
            else:
                gemm(acc, h, blob, lo["L%d" % i], 4, 4)
            acc = relu(acc)
            h = chain(acc)
    bo0, bo1 = blob[lo["BO"] + 2 * T], blob[lo["BO"] + 2 * T + 1]
    out = [np.stack([bo0, bo1, bo0, bo1], 1).copy()]
    gemm(out, h, blob, lo["OUT"], 4, 1)
    return c_to_rows(out, col_natural, 8), masks


def backward(kind, blob, masks, dout):
    """dout [16, 8] (padded) -> d(e) [16,96] (two contributions summed), d(c) [16,32]."""
    lo = layout(kind)
    embed = bool(lo["EMBED"])
    dA = [a_from_rows(dout, slot_natural)(0)]
    dA[0][:, 2:] = 0.0
    dh = zeros(4)
    gemm(dh, dA, blob, lo["TOUT"], 1, 4)
    dc = zeros(4)
    dE = np.zeros((16, 96))
    for i in (4, 3, 2, 1, 0):
        a = chain(dh)
        if embed:
            gemm(dc, a, blob, lo["TC%d" % i], 4, 4)
        dh = [np.where(m, x, 0.0) for x, m in zip(dh, masks[i])]
        a = chain(dh)
        dh = zeros(4)
        if i in (3, 0):
            if embed:
                de = zeros(12)
                gemm(de, a, blob, lo["TL3X"] if i == 3 else lo["TL0"], 4, 12)
                dE += c_to_rows(de, col_natural, 96)
            else:
                gemm(dc, a, blob, lo["TL3X"] if i == 3 else lo["TL0"], 4, 4)
        if i != 0:
            gemm(dh, a, blob, lo["TL%d" % i], 4, 4)
    return dE, c_to_rows(dc, col_gather, 32)
