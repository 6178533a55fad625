"""Model of libstdc++ std::sort (introsort) as a level-synchronous sequence of data-parallel Hoare partitions,
followed by one stable sort. This is the formulation the device kernel (order_candidates in detect.cu) uses; the
script checks it against the real std::sort compiled into oracle/_ref (ref_std_sort).  Development aid."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def heap_sort_segment(a, sc, f, l):
    """std::__partial_sort(first, last, last) = __heap_select (make_heap) + __sort_heap, comp(x,y) = sc[x] > sc[y]."""
    comp = lambda x, y: sc[x] > sc[y]
    n = l - f

    def push_heap(hole, top, val):
        parent = (hole - 1) // 2
        while hole > top and comp(a[f + parent], val):
            a[f + hole] = a[f + parent]
            hole = parent
            parent = (hole - 1) // 2
        a[f + hole] = val

    def adjust_heap(hole, length, val):
        top = hole
        child = hole
        while child < (length - 1) // 2:
            child = 2 * (child + 1)
            if comp(a[f + child], a[f + child - 1]):
                child -= 1
            a[f + hole] = a[f + child]
            hole = child
        if (length & 1) == 0 and child == (length - 2) // 2:
            child = 2 * (child + 1)
            a[f + hole] = a[f + child - 1]
            hole = child - 1
        push_heap(hole, top, val)

    if n < 2:
        return
    parent = (n - 2) // 2
    while True:
        adjust_heap(parent, n, a[f + parent])
        if parent == 0:
            break
        parent -= 1
    last = n
    while last > 1:
        last -= 1
        val = a[f + last]
        a[f + last] = a[f]
        adjust_heap(0, last, val)


def partition_phase(sc):
    n = len(sc)
    a = list(range(n))
    if n == 0:
        return a
    segs = [(0, n)]
    depth = 2 * (n.bit_length() - 1)
    while segs:
        new = []
        for f, l in segs:
            if l - f <= 16:
                continue
            if depth == 0:
                heap_sort_segment(a, sc, f, l)
                continue
            mid = f + (l - f) // 2
            A, B, Cc = f + 1, mid, l - 1
            comp = lambda x, y: sc[a[x]] > sc[a[y]]
            if comp(A, B):
                pick = B if comp(B, Cc) else (Cc if comp(A, Cc) else A)
            else:
                pick = A if comp(A, Cc) else (Cc if comp(B, Cc) else B)
            a[f], a[pick] = a[pick], a[f]
            pv = sc[a[f]]
            seg = np.asarray([sc[i] for i in a[f:l]])
            Ls = [f + i for i in range(1, l - f) if seg[i] <= pv]            # ascending
            Rs = [f + i for i in range(l - f - 1, -1, -1) if seg[i] >= pv]   # descending, includes f (the pivot)
            t = 0
            while t < len(Ls) and t < len(Rs) and Ls[t] < Rs[t]:
                t += 1
            for i in range(t):
                a[Ls[i]], a[Rs[i]] = a[Rs[i]], a[Ls[i]]
            if t == 0:
                cut = Ls[0]
            else:
                cut = min(Ls[t], Rs[t - 1]) if t < len(Ls) else Rs[t - 1]
            new.append((f, cut))
            new.append((cut, l))
        segs = new
        depth -= 1
    return a


def model_sort(sc):
    a = partition_phase(list(sc))
    order = sorted(range(len(a)), key=lambda i: -sc[a[i]])  # Python's sort is stable
    return [a[i] for i in order]


if __name__ == "__main__":
    import refbind
    ref = refbind.Ref()
    rng = np.random.default_rng(0)
    bad = 0
    cases = []
    for trial in range(300):
        n = int(rng.integers(0, 3000))
        hi = int(rng.choice([2, 5, 50, 1000, 10 ** 6]))
        cases.append(rng.integers(0, hi, n))
    cases.append(np.arange(5000))
    cases.append(np.arange(5000)[::-1])
    cases.append(np.zeros(4000, np.int64))
    # median-of-3 killer-ish patterns to reach the heapsort fallback
    for n in (2000, 4096):
        k = n // 2
        kill = np.zeros(n, np.int64)
        for i in range(k):
            if i % 2 == 0:
                kill[i] = i + 1
            else:
                kill[i] = k + i + (1 if i % 2 else 0)
            kill[k + i] = 2 * (i + 1)
        cases.append(kill)
        cases.append(-kill)
    for sc in cases:
        n = len(sc)
        pts = np.zeros(n, refbind.PWS_DTYPE)
        pts["score"] = sc
        pts["x"] = np.arange(n) % 65536
        pts["y"] = np.arange(n) // 65536
        want = ref.std_sort(pts)
        widx = want["y"].astype(np.int64) * 65536 + want["x"]
        got = np.asarray(model_sort([int(v) for v in sc]), np.int64)
        if not np.array_equal(widx, got):
            bad += 1
            print("MISMATCH n=%d" % n)
    print("cases", len(cases), "bad", bad)
