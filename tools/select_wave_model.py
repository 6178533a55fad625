"""CPU model of the wave-parallel greedy uniformity selection (design aid for select_keypoints_kernel).

Counts, per config image, how many rounds a window-based parallel resolution of EnforceKeyPointUniformity
(uniformity-enforcement-inl.h:57-207) needs, and checks that the accepted set equals the compiled reference's.
Rule of a round: every pending candidate re-tests the occupancy at its cell (monotone: a failure is final); a
pending candidate with no LOWER-index pending candidate in the 3x3 bins around it is final -> accepted, stamped.
"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import refbind
import __graft_entry__ as ge
import importlib
mod = ge.load()
syn = importlib.import_module("okvis2x_b200.synthetic")


def simulate(pts, lut, radius, H, W, max_kpt, window, binsz, exact_near=False):
    scaling = np.float32(15.0 / np.float32(radius))
    occ = np.zeros((H + 64, W + 64), np.int32)
    score = pts["score"].astype(np.float32)
    maxs = score[0]
    nsc1 = np.sqrt(np.sqrt(score / maxs)).astype(np.float32) * np.float32(255.0)
    cy = (pts["y"].astype(np.float32) * scaling + np.float32(16.0)).astype(np.int32)
    cx = (pts["x"].astype(np.float32) * scaling + np.float32(16.0)).astype(np.int32)
    n = len(pts)
    kept = []
    rounds = 0
    pos = 0
    fills = 0
    acc_per_round = []
    while pos < n and len(kept) < max_kpt:
        # fill: pre-test candidates until `window` survivors are found
        win = []
        while pos < n and len(win) < window:
            hi = min(n, pos + window)
            idx = np.arange(pos, hi)
            alive = ~(nsc1[idx] < occ[cy[idx], cx[idx]].clip(max=255).astype(np.float32))
            win.extend(idx[alive].tolist())
            pos = hi
        fills += 1
        win = np.array(win[:], np.int64)   # may exceed `window` slightly; fine for a model
        pending = np.ones(len(win), bool)
        accepted = np.zeros(len(win), bool)
        while pending.any():
            rounds += 1
            p = np.nonzero(pending)[0]
            fail = nsc1[win[p]] < occ[cy[win[p]], cx[win[p]]].clip(max=255).astype(np.float32)
            pending[p[fail]] = False
            p = p[~fail]
            if len(p) == 0:
                break
            bx, by = cx[win[p]] // binsz, cy[win[p]] // binsz
            nbx = (W + 64) // binsz + 2
            binmin = np.full((nbx + 2) * (nbx + 2), 1 << 30, np.int64)
            key = (by + 1) * (nbx + 2) + (bx + 1)
            np.minimum.at(binmin, key, p)
            m = np.full(len(p), 1 << 30, np.int64)
            for dy in (-1, 0, 1):
                for dx in (-1, 0, 1):
                    m = np.minimum(m, binmin[key + dy * (nbx + 2) + dx])
            acc = p[m == p]
            acc_per_round.append(len(acc))
            for a in acc:
                i = win[a]
                v = np.ceil(lut * (np.float32(0.99) * nsc1[i])).astype(np.int32) & 0xff
                sl = occ[cy[i] - 15:cy[i] + 16, cx[i] - 15:cx[i] + 16]
                np.minimum(sl + v, 255, out=sl)
            accepted[acc] = True
            pending[acc] = False
        kept.extend(win[accepted].tolist())
    kept = kept[:max_kpt]
    return kept, rounds, fills, acc_per_round


def main():
    ref = refbind.Ref()
    lut = ref.lut()
    cases = {
        "euroc-scene400": (syn.EUROC, syn.scene_image(752, 480, 1)),
        "euroc-scene1500": (syn.EUROC, syn.scene_image(752, 480, 1, rects=1500)),
        "euroc-noise": (syn.EUROC, syn.noise_image(752, 480, 2)),
        "d455-scene1500": (syn.D455, syn.scene_image(848, 480, 8, rects=1500)),
        "d455-noise": (syn.D455, syn.noise_image(848, 480, 9)),
        "hilti-scene": (syn.HILTI, syn.scene_image(720, 540, 3)),
    }
    for name, (cfg, img) in cases.items():
        H, W = img.shape
        raw = ref.maxima(img, cfg["abs_threshold"])
        pts = ref.std_sort(raw)
        want = ref.uniformity(raw, cfg["radius"], H, W, cfg["max_kpt"])   # sorts internally (tie order!)
        for window, binsz in ((256, 16), (1024, 16), (1024, 32), (2048, 16)):
            kept, rounds, fills, apr = simulate(pts, lut, cfg["radius"], H, W, cfg["max_kpt"], window, binsz)
            ok = len(kept) == len(want) and np.array_equal(pts[kept], want)
            print("%-16s n=%6d kept=%5d window=%5d bin=%2d rounds=%5d fills=%3d acc/round=%.2f ok=%s" % (
                name, len(pts), len(kept), window, binsz, rounds, fills, np.mean(apr) if apr else 0, ok))


if __name__ == "__main__":
    main()
