"""Pure-Python interpreter of the evaluation program the plan compiler emits
(minterpy_b200/csrc/mpx_plan.cu: compile_program; record format in mpx_common.h).

Test infrastructure: lets the CPU suite check the index-set compiler -- leaf records,
fold counts, irregular extensions, coefficient slots -- without a GPU, by evaluating the
program exactly the way the CUDA kernel walks it (mpx_eval.cu: eval_tree_kernel).
"""
import ctypes

import numpy as np

from minterpy_b200 import _lib

DESC_WORDS, DESC_HDR, COEF_CAP = 512, 4, 1024
D_IRREG, D_LAST = 1 << 12, 1 << 13


def compile_program(exponents):
    e = np.ascontiguousarray(exponents, dtype=np.int64)
    N, m = e.shape
    nch, fma = ctypes.c_int64(), ctypes.c_int64()
    lib = _lib.lib()
    _lib.check(lib.mpx_debug_program(e.ctypes.data, N, m, None, None, ctypes.byref(nch),
                                     ctypes.byref(fma)))
    desc = np.zeros(nch.value * DESC_WORDS, dtype=np.uint32)
    slots = np.zeros(nch.value * COEF_CAP, dtype=np.int32)
    _lib.check(lib.mpx_debug_program(e.ctypes.data, N, m, desc.ctypes.data, slots.ctypes.data,
                                     ctypes.byref(nch), ctypes.byref(fma)))
    return desc.reshape(-1, DESC_WORDS), slots.reshape(-1, COEF_CAP), fma.value


def l0p_for(max_exp0):
    L0 = max_exp0 + 1
    return 10 if L0 <= 10 else 18 if L0 <= 18 else 26 if L0 <= 26 else 34 if L0 <= 34 else 66


def run_program(desc, slots, x, coeffs, exponents, gp):
    """Value of sum_j coeffs[j] * N_j(x) for ONE point x, walking the program."""
    e = np.asarray(exponents)
    N, m = e.shape
    U = max(m - 2, 0)
    mx = e.max(axis=0)
    l0p = l0p_for(int(mx[0]))
    ncw = (l0p // 2 + 7) // 8
    order = np.lexsort(e.T)
    T0 = np.zeros(66)
    T0[0] = 1.0
    for j in range(mx[0]):
        T0[j + 1] = T0[j] * (x[0] - gp[j, 0])
    T1 = np.ones(int(mx[1]) + 1 if m >= 2 else 1)
    if m >= 2:
        for j in range(mx[1]):
            T1[j + 1] = T1[j] * (x[1] - gp[j, 1])
    xu = [x[2 + u] for u in range(U)]
    t = [1.0] * U
    for u in range(U):
        for j in range(int(e[order[0], 2 + u])):
            t[u] *= xu[u] - gp[j, 2 + u]
    acc = [0.0] * U
    acc1 = 0.0
    fma = 0
    for c in range(desc.shape[0]):
        dw = desc[c]
        packed = np.where(slots[c] >= 0, coeffs[np.maximum(slots[c], 0)], 0.0)
        cp = 0
        i = DESC_HDR
        end = DESC_HDR + int(dw[0])
        while i < end:
            w = int(dw[i])
            counts = [int(v) for v in dw[i + 1:i + 1 + ncw]]
            i += 1 + ncw
            e1, nf = w & 255, (w >> 8) & 15
            for K in range(l0p // 2, 0, -1):           # lengths walk downwards
                cnt = (counts[(K - 1) // 8] >> (4 * ((K - 1) % 8))) & 15
                for _ in range(cnt):
                    a = packed[cp]                     # T0[0] = 1: enters as an addend
                    for s in range(1, 2 * K):
                        a += packed[cp + s] * T0[s]
                        fma += 1
                    cp += 2 * K
                    acc1 += T1[e1] * a
                    e1 += 1
                    fma += 1
            if nf:
                adv, cur = 1, (w >> 14) & 255
                irregular, last = bool(w & D_IRREG), bool(w & D_LAST)
                if irregular:
                    cur, adv = int(dw[i]) & 255, int(dw[i]) >> 8
                low, acc1 = acc1, 0.0
                for u in range(nf):
                    av = acc[u] + t[u] * low
                    fma += 1
                    if u < nf - 1:
                        low, acc[u] = av, 0.0
                        tn = 1.0
                        if irregular:
                            for j in range(int(dw[i + 1 + u])):
                                tn *= xu[u] - gp[j, 2 + u]
                        t[u] = tn
                    else:
                        acc[u] = av
                        if not last:
                            for j in range(adv):
                                t[u] *= xu[u] - gp[cur + j, 2 + u]
                if irregular:
                    i += nf
    return (acc[U - 1] if U > 0 else acc1), fma


# ---- tile program of the tensor-core tree kernel (mpx_common.h: TileProgram) ---------------
T_DESC_WORDS, T_COEF_CAP = 128, 2048


def compile_tile_program(exponents):
    e = np.ascontiguousarray(exponents, dtype=np.int64)
    N, m = e.shape
    nch, fma = ctypes.c_int64(), ctypes.c_int64()
    lib = _lib.lib()
    _lib.check(lib.mpx_debug_tile_program(e.ctypes.data, N, m, None, None, ctypes.byref(nch),
                                          ctypes.byref(fma)))
    desc = np.zeros(nch.value * T_DESC_WORDS, dtype=np.uint32)
    slots = np.zeros(nch.value * T_COEF_CAP, dtype=np.int32)
    _lib.check(lib.mpx_debug_tile_program(e.ctypes.data, N, m, desc.ctypes.data, slots.ctypes.data,
                                          ctypes.byref(nch), ctypes.byref(fma)))
    return desc.reshape(-1, T_DESC_WORDS), slots.reshape(-1, T_COEF_CAP), fma.value


def run_tile_program(desc, slots, x, coeffs, exponents, gp):
    """Value of the polynomial at ONE point x, walking the tile program the way
    eval_tmma_kernel does (complete index sets, m >= 2)."""
    e = np.asarray(exponents)
    N, m = e.shape
    U = m - 2
    mx = e.max(axis=0)
    T0 = np.zeros(4 * ((mx[0] + 4) // 4) + 4)
    T0[0] = 1.0
    for j in range(mx[0]):
        T0[j + 1] = T0[j] * (x[0] - gp[j, 0])
    T1 = np.ones(int(mx[1]) + 1)
    for j in range(mx[1]):
        T1[j + 1] = T1[j] * (x[1] - gp[j, 1])
    t = [1.0] * max(U, 0)
    acc = [0.0] * max(U, 0)
    acc_cur = acc_next = 0.0
    fma = 0
    for c in range(desc.shape[0]):
        dw = desc[c]
        packed = np.where(slots[c] >= 0, coeffs[np.maximum(slots[c], 0)], 0.0)
        cp = 0
        i = DESC_HDR
        end = DESC_HDR + int(dw[0])
        while i < end:
            w0, w1 = int(dw[i]), int(dw[i + 1])
            i += 2
            ks, ncur, nnext = w0 & 31, (w0 >> 5) & 15, (w0 >> 9) & 15
            e1, ends, nf, last = (w0 >> 13) & 255, (w0 >> 21) & 1, (w0 >> 22) & 15, (w0 >> 26) & 1
            C = np.zeros(8)
            for s in range(ks):                       # one mma k-step: 4 exponents x 8 columns
                for col in range(8):
                    for q in range(4):
                        C[col] += packed[cp + (s >> 1) * 64 + (col * 4 + q) * 2 + (s & 1)] * T0[4 * s + q]
            fma += 32 * ks
            cp += 64 * ((ks + 1) // 2)
            for col in range(8):
                if col < ncur:
                    acc_cur += T1[e1 + col] * C[col]
                elif col < ncur + nnext:
                    acc_next += T1[col - ncur] * C[col]
                else:
                    assert C[col] == 0.0              # padding column
            fma += 8
            if ends:
                low = acc_cur
                for u in range(nf):
                    av = acc[u] + t[u] * low
                    fma += 1
                    if u < nf - 1:
                        low, acc[u], t[u] = av, 0.0, 1.0
                    else:
                        acc[u] = av
                        if not last:
                            t[u] *= x[2 + u] - gp[w1, 2 + u]
                if nf == 0:
                    assert last and U == 0
                else:
                    acc_cur = 0.0
                if not (nf == 0):
                    acc_cur, acc_next = acc_next, 0.0
    return (acc[U - 1] if U > 0 else acc_cur), fma
