"""CPU model of the chunk-local bottom-up radix-tree construction used by k_fit_collapse (design check, not product code).

Sorted keys (duplicates allowed) -> cpl[i] = common prefix length of augmented keys i, i+1 (Karras' delta with the index tie-break).
A node is named by its split position i (between leaf i and i+1).  A subtree with leaf range [l, r] has parent r if
cpl[r] > cpl[l-1] (it is the left child) else l-1 (right child); cpl[-1] = cpl[n-1] = -1.  Node i is chunk-local iff its whole range
lies in one chunk: cpl[i] > min(cpl[chunk_lo-1 .. i-1]) and cpl[i] > min(cpl[i+1 .. chunk_hi]).
Checks against a direct top-down construction of the same radix tree."""
import sys
import numpy as np


def cpl_of(keys, bits=64):
    n = len(keys)
    out = np.full(n + 1, -1, np.int64)          # out[i+1] = cpl(i), so out[0] = cpl(-1) = -1, out[n] = cpl(n-1) = -1
    for i in range(n - 1):
        a, b = int(keys[i]), int(keys[i + 1])
        if a != b:
            out[i + 1] = bits - (a ^ b).bit_length()
        else:
            out[i + 1] = bits + 32 - (i ^ (i + 1)).bit_length()
    return out


def topdown(cpl, l, r, nodes):
    """reference: split of [l, r] is at the position of the minimum cpl inside (unique)"""
    if l == r:
        return
    seg = cpl[l + 1:r + 1]                       # cpl(l) .. cpl(r-1)
    m = int(seg.min())
    pos = [l + k for k, v in enumerate(seg) if v == m]
    assert len(pos) == 1
    s = pos[0]
    nodes[s] = (l, r)
    topdown(cpl, l, s, nodes)
    topdown(cpl, s + 1, r, nodes)


def bottomup(cpl, n, chunk):
    rng = {}
    pending = {}
    roots = []
    local = np.zeros(max(n - 1, 1), bool)
    for i in range(n - 1):
        lo = (i // chunk) * chunk
        hi = min(lo + chunk, n) - 1
        if i + 1 > hi:                            # the split itself is a chunk boundary
            continue
        left_min = cpl[lo:i + 1].min()            # cpl(lo-1) .. cpl(i-1)
        right_min = cpl[i + 2:hi + 2].min()       # cpl(i+1) .. cpl(hi)
        local[i] = cpl[i + 1] > left_min and cpl[i + 1] > right_min
    order = np.random.permutation(n)
    for j in order:
        l = r = int(j)
        while True:
            if l == 0 and r == n - 1:
                break
            right_parent = (l == 0) or (r != n - 1 and cpl[r + 1] > cpl[l])     # cpl(r) > cpl(l-1)
            p = r if right_parent else l - 1
            if not local[p]:
                roots.append((l, r))
                # top phase: continue globally (same rule)
            if p in pending:
                o = pending.pop(p)
                l, r = (l, o) if right_parent else (o, r)
                rng[p] = (l, r)
            else:
                pending[p] = l if right_parent else r
                # store OUR far end: as left child (p == r) we know l; as right child we know r
                break
    assert not pending
    return rng, local, roots


def main():
    rs = np.random.RandomState(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
    for trial in range(200):
        n = int(rs.randint(2, 700))
        chunk = int(rs.choice([4, 8, 16, 64]))
        kind = rs.randint(3)
        if kind == 0:
            keys = np.sort(rs.randint(0, 1 << 20, n).astype(np.uint64))
        elif kind == 1:
            keys = np.sort(rs.randint(0, 8, n).astype(np.uint64))           # heavy duplicates
        else:
            keys = np.sort((rs.randint(0, 1 << 10, n).astype(np.uint64) << np.uint64(rs.randint(0, 30))))
        cpl = cpl_of(keys)
        ref = {}
        topdown(cpl, 0, n - 1, ref)
        got, local, roots = bottomup(cpl, n, chunk)
        assert got == ref, (trial, n)
        for i, (l, r) in ref.items():
            inside = l // chunk == r // chunk
            assert bool(local[i]) == inside, (trial, i, l, r, chunk, local[i])
    print("ok")


main()
