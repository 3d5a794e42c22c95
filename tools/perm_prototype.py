"""NumPy prototype of the parallel formulation of Generator.permutation that csrc/perm.cu implements: (A) the sample
each Fisher-Yates draw uses, found per mask phase with bracketed prefix sums; (B) the composition of the swaps through
per-target step lists and pointer jumping.  `python tools/perm_prototype.py [n]` checks it against NumPy;
tests/test_host.py runs it on the CPU so the algorithm itself is covered without a GPU."""
import numpy as np, sys, time
from numpy.random import PCG64, Generator

def halves_of(seed, jump, count):
    raw = PCG64(seed).jumped(jump).random_raw((count + 1) // 2 + 1)
    h = np.empty(2 * len(raw), dtype=np.uint64); h[0::2] = raw & 0xffffffff; h[1::2] = raw >> 32
    return h

def resolve_draws(n, halves):
    """J[k] for k = 0..n-2 (i = n-1-k), phase by phase with interval iteration."""
    J = np.empty(n - 1, dtype=np.int64)
    S0 = 0; k0 = 0; i_hi = n - 1; iters_total = 0
    while i_hi >= 1:
        b = int(i_hi).bit_length() - 1            # 2^b <= i_hi < 2^(b+1)
        i_lo = max(1 << b, 1)
        mask = (1 << (b + 1)) - 1
        count = i_hi - i_lo + 1
        Wn = int(2 * count + 10 * np.sqrt(2 * count) + 64)
        v = (halves[S0:S0 + Wn] & np.uint64(mask)).astype(np.int64)
        A = v <= i_lo; C = (~A) & (v <= i_hi)
        a_before = np.cumsum(A) - A
        cidx = np.flatnonzero(C)
        theta = i_hi - v[cidx] - a_before[cidx]
        nc = len(cidx)
        lo = np.zeros(nc, dtype=np.int64); hi = np.arange(nc, dtype=np.int64)
        it = 0
        while True:
            acc = hi <= theta; rej = lo > theta; unc = ~(acc | rej)
            it += 1
            if not unc.any(): break
            lo = np.cumsum(acc) - acc
            hi = lo + (np.cumsum(unc) - unc)
        iters_total += it
        accepted = A.copy(); accepted[cidx] = acc
        t = np.cumsum(accepted)
        assert t[-1] >= count
        s_end = int(np.searchsorted(t, count))      # first index with t == count
        take = np.flatnonzero(accepted[:s_end + 1])
        assert len(take) == count
        J[k0:k0 + count] = v[take]
        k0 += count; S0 += s_end + 1; i_hi = i_lo - 1
    return J, S0, iters_total

def apply_swaps(n, J):
    """final permutation from j_i (i = n-1-k)."""
    i_of = np.arange(n - 1, 0, -1)                  # i for k
    j = np.empty(n, dtype=np.int64); j[0] = 0; j[i_of] = J
    steps = np.arange(1, n)
    real = steps[j[steps] != steps]                 # steps that move something
    # per-target sorted lists
    order = np.lexsort((real, j[real]))
    rs = real[order]; tg = j[rs]
    q = np.full(n, -1, dtype=np.int64)
    same = tg[:-1] == tg[1:]
    q[rs[:-1][same]] = rs[1:][same]
    parent = np.full(n, -1, dtype=np.int64)
    first = np.ones(len(rs), dtype=bool); first[1:] = ~same
    parent[tg[first]] = rs[first]
    # W by pointer jumping
    root = np.where(parent < 0, np.arange(n), parent)
    rounds = 0
    while True:
        nxt = root[root]
        rounds += 1
        if np.array_equal(nxt, root): break
        root = nxt
    W = root
    a = np.empty(n, dtype=np.int64)
    idx = np.arange(1, n)
    selfm = j[idx] == idx
    a[idx] = np.where(selfm, W[idx], np.where(q[idx] >= 0, W[np.maximum(q[idx], 0)], j[idx]))
    a[0] = W[parent[0]] if parent[0] >= 0 else 0
    return a, rounds

def permutation(n, seed, jump):
    """Generator(PCG64(seed).jumped(jump)).permutation(n) through the parallel formulation; also returns the number
    of samples consumed, bracket rounds (all phases) and pointer-jumping rounds."""
    if n < 2:
        return np.arange(n, dtype=np.int64), 0, 0, 0
    h = halves_of(seed, jump, 3 * n + 4096)
    J, used, its = resolve_draws(n, h)
    a, rounds = apply_swaps(n, J)
    return a, used, its, rounds


if __name__ == "__main__":
    for n in [2, 3, 5, 17, 1000, 4097, 100003, int(sys.argv[1]) if len(sys.argv) > 1 else 1000000]:
        t0 = time.time()
        a, used, its, rounds = permutation(n, 1337, 3)
        b = Generator(PCG64(1337).jumped(3)).permutation(n)
        print(n, np.array_equal(a, b), "samples", used, "iters", its, "jump rounds", rounds, "%.2fs" % (time.time() - t0))
