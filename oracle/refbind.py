"""TEST INFRASTRUCTURE ONLY. ctypes bindings of the oracle libraries.

  oracle/_ref/libbrisk_ref.so  - the reference's own external/brisk sources (see oracle/Makefile, ref_capi.cc)
  oracle/liboracle_port.so     - our CPU restatement of the matchers / geometry (oracle/port/*.cc)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_PATH = os.path.join(_HERE, "_ref", "libbrisk_ref.so")
PORT_PATH = os.path.join(_HERE, "liboracle_port.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
PWS_DTYPE = np.dtype([("score", "<i4"), ("x", "<u2"), ("y", "<u2")])
assert KP_DTYPE.itemsize == 28 and PWS_DTYPE.itemsize == 8

_vp, _i, _d, _u64, _f = C.c_void_p, C.c_int, C.c_double, C.c_uint64, C.c_float


def _ptr(a):
    return C.c_void_p(a.ctypes.data)


def aligned(shape, dtype, align=64):
    """numpy array whose data pointer is `align`-byte aligned (brisk::Hamming dereferences __m128i*)."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    raw = np.zeros(n + align, np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n].view(dtype).reshape(shape)


class Ref:
    """The compiled reference (brisk) behind C entry points."""

    def __init__(self):
        if not os.path.exists(REF_PATH):
            raise FileNotFoundError(REF_PATH + " missing: run `make -C oracle` where /root/reference exists")
        L = self.L = C.CDLL(REF_PATH)
        L.ref_harris.argtypes = [_vp, _i, _i, _vp]
        L.ref_maxima.argtypes = [_vp, _i, _i, _i, _vp, _i]
        L.ref_uniformity.argtypes = [_vp, _i, _d, _i, _i, _u64]
        L.ref_lut.argtypes = [_vp]
        L.ref_detect.argtypes = [_vp, _i, _i, _d, _u64, _d, _u64, _vp, _i]
        L.ref_extractor_create.argtypes = [_i, _i]
        L.ref_extractor_create.restype = _vp
        L.ref_extractor_destroy.argtypes = [_vp]
        L.ref_extractor_set_camera.argtypes = [_vp, _vp, _vp, _i, _i, _f]
        L.ref_extractor_set_direction.argtypes = [_vp, _vp]
        L.ref_describe.argtypes = [_vp, _vp, _i, _i, _vp, _i, _vp]
        L.ref_integral.argtypes = [_vp, _i, _i, _vp]
        L.ref_pattern_points.argtypes = [_vp, _i, _vp]
        L.ref_pattern_meta.argtypes = [_vp, _vp, _vp, _vp, _vp, _i]
        L.ref_hamming.argtypes = [_vp, _vp, _i]
        L.ref_hamming.restype = C.c_uint32
        L.ref_hamming_best.argtypes = [_vp, _i, _vp, _i, C.c_uint32, _vp]
        L.ref_std_sort.argtypes = [_vp, _i]
        L.shim_eigen2x2.argtypes = [_vp, _vp]
        L.shim_gemm_2x3_3x2.argtypes = [_vp, _vp, _vp]
        L.shim_normalize3.argtypes = [_vp, _vp]

    # the cv shim's restatement of OpenCV arithmetic (pinned against cv2 goldens in tests/test_oracle_pins.py)
    def shim_eigen2x2(self, a):
        a = np.ascontiguousarray(a, np.float32).reshape(4)
        out = np.zeros(2, np.float32)
        self.L.shim_eigen2x2(_ptr(a), _ptr(out))
        return out

    def shim_gemm(self, a, b):
        a = np.ascontiguousarray(a, np.float32).reshape(6)
        b = np.ascontiguousarray(b, np.float32).reshape(6)
        out = np.zeros(4, np.float32)
        self.L.shim_gemm_2x3_3x2(_ptr(a), _ptr(b), _ptr(out))
        return out.reshape(2, 2)

    def shim_normalize3(self, v):
        v = np.ascontiguousarray(v, np.float32).reshape(3)
        out = np.zeros(3, np.float32)
        self.L.shim_normalize3(_ptr(v), _ptr(out))
        return out

    @staticmethod
    def _img(img):
        img = np.ascontiguousarray(img, np.uint8)
        assert img.ndim == 2
        return img

    def harris(self, img):
        img = self._img(img)
        out = np.zeros(img.shape, np.int32)
        self.L.ref_harris(_ptr(img), img.shape[0], img.shape[1], _ptr(out))
        return out

    def maxima(self, img, abs_threshold, cap=1 << 20):
        img = self._img(img)
        out = np.zeros(cap, PWS_DTYPE)
        n = self.L.ref_maxima(_ptr(img), img.shape[0], img.shape[1], int(abs_threshold), _ptr(out), cap)
        assert n <= cap
        return out[:n].copy()

    def uniformity(self, points, radius, rows, cols, max_kpt):
        pts = np.ascontiguousarray(points, PWS_DTYPE).copy()
        n = self.L.ref_uniformity(_ptr(pts), len(pts), float(radius), rows, cols, int(max_kpt))
        return pts[:n].copy()

    def std_sort(self, points):
        """std::sort of PointWithScore (the reference's operator<, libstdc++ introsort) - tie order included."""
        pts = np.ascontiguousarray(points, PWS_DTYPE).copy()
        self.L.ref_std_sort(_ptr(pts), len(pts))
        return pts

    def lut(self):
        out = np.zeros((31, 31), np.float32)
        self.L.ref_lut(_ptr(out))
        return out

    def detect(self, img, radius, octaves, abs_threshold, max_kpt, cap=1 << 17):
        img = self._img(img)
        out = np.zeros(cap, KP_DTYPE)
        n = self.L.ref_detect(_ptr(img), img.shape[0], img.shape[1], float(radius), int(octaves), float(abs_threshold),
                              int(max_kpt), _ptr(out), cap)
        if n < 0:
            raise RuntimeError("reference detector threw")
        assert n <= cap
        return out[:n].copy()

    def integral(self, img):
        img = self._img(img)
        out = np.zeros((img.shape[0] + 1, img.shape[1] + 1), np.int32)
        self.L.ref_integral(_ptr(img), img.shape[0], img.shape[1], _ptr(out))
        return out

    def hamming(self, a, b):
        aa, bb = aligned(48, np.uint8), aligned(48, np.uint8)
        aa[:] = a
        bb[:] = b
        return int(self.L.ref_hamming(_ptr(aa), _ptr(bb), 3))

    def hamming_best(self, queries, mp, threshold):
        q = aligned(queries.shape, np.uint8)
        m = aligned(mp.shape, np.uint8)
        q[:] = queries
        m[:] = mp
        best = np.zeros(len(q), np.uint64)
        self.L.ref_hamming_best(_ptr(q), len(q), _ptr(m), len(m), int(threshold), _ptr(best))
        return best


class RefExtractor:
    """brisk::BriskDescriptorExtractor(rotationInvariant, scaleInvariant) of the compiled reference."""

    def __init__(self, ref, rotation_invariant=True, scale_invariant=False):
        self.ref = ref
        self.h = C.c_void_p(ref.L.ref_extractor_create(int(rotation_invariant), int(scale_invariant)))

    def close(self):
        if self.h:
            self.ref.L.ref_extractor_destroy(self.h)
            self.h = None

    def set_camera(self, rays, jacobians, focal):
        rays = np.ascontiguousarray(rays, np.float32)
        jacobians = np.ascontiguousarray(jacobians, np.float32)
        rows, cols = rays.shape[:2]
        assert rays.shape == (rows, cols, 3) and jacobians.shape == (rows, cols, 6)
        self.ref.L.ref_extractor_set_camera(self.h, _ptr(rays), _ptr(jacobians), rows, cols, float(focal))

    def set_direction(self, d):
        d = np.ascontiguousarray(d, np.float32)
        self.ref.L.ref_extractor_set_direction(self.h, _ptr(d))

    def describe(self, img, keypoints):
        img = Ref._img(img)
        kp = np.ascontiguousarray(keypoints, KP_DTYPE).copy()
        desc = aligned((max(len(kp), 1), 48), np.uint8)
        n = self.ref.L.ref_describe(self.h, _ptr(img), img.shape[0], img.shape[1], _ptr(kp), len(kp), _ptr(desc))
        if n < 0:
            raise RuntimeError("reference extractor threw")
        return kp[:n].copy(), desc[:n].copy()

    def pattern(self, scale):
        out = np.zeros((1024, 66, 3), np.float32)
        self.ref.L.ref_pattern_points(self.h, int(scale), _ptr(out))
        return out

    def meta(self):
        sl = np.zeros(64, np.float32)
        sz = np.zeros(64, np.uint32)
        sp = np.zeros((384, 2), np.uint32)
        lp = np.zeros((856, 4), np.int32)
        n = self.ref.L.ref_pattern_meta(self.h, _ptr(sl), _ptr(sz), _ptr(sp), _ptr(lp), 856)
        assert n == 856
        return sl, sz, sp, lp


class PortStereoView(C.Structure):
    _fields_ = [("desc", C.c_void_p), ("size", C.c_void_p), ("e_C", C.c_void_p), ("bp_valid", C.c_void_p),
                ("n", C.c_int), ("C", C.c_double * 9), ("r", C.c_double * 3), ("f", C.c_double)]


class Port:
    """Our CPU restatement of the Frontend.cpp matchers (oracle/port/matchers.cc)."""

    def __init__(self, variant=0):
        """variant != 0: a sensitivity build (PORT_EIGEN_VARIANT bits, oracle/port/matchers.cc)."""
        self.variant = int(variant)
        path = PORT_PATH if not variant else PORT_PATH.replace(".so", "_v%d.so" % variant)
        if not os.path.exists(path):
            raise FileNotFoundError(path + " missing: run `make -C oracle port variants`")
        L = self.L = C.CDLL(path)
        L.port_hamming.argtypes = [_vp, _vp]
        L.port_hamming.restype = C.c_uint32
        L.port_hamming_best.argtypes = [_vp, _i, _vp, C.c_int64, C.c_uint32, _u64, _vp, _i]
        L.port_match_to_map_thread.argtypes = [_i, _i, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _d, _vp, _vp, _vp,
                                               _vp]
        L.port_match_stereo.argtypes = [C.POINTER(PortStereoView), C.POINTER(PortStereoView), _d, _vp, _vp, _vp, _vp]
        L.port_match_motion_stereo.argtypes = [C.POINTER(PortStereoView), _vp, C.POINTER(PortStereoView), _d, _vp, _vp,
                                               _vp, _vp, _vp]
        L.port_match_uninit_thread.argtypes = [_i, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _d, _vp,
                                               _vp, _vp, _vp, _vp]

    def back_project(self, cam, keypoints):
        """cam: a ctypes structure with the okc_pinhole layout; keypoints: 28-byte cv::KeyPoint records."""
        kp = np.ascontiguousarray(keypoints)
        n = len(kp)
        e = np.zeros((max(n, 1), 3), np.float64)
        valid = np.zeros(max(n, 1), np.uint8)
        self.L.port_back_project.argtypes = [_vp, _vp, _i, _vp, _vp]
        nv = self.L.port_back_project(C.byref(cam), _ptr(kp), n, _ptr(e), _ptr(valid))
        return e[:n], valid[:n], int(nv)

    def project_landmarks(self, cam, T_CW, hp_W, repr_threshold):
        T = np.ascontiguousarray(T_CW, np.float64).reshape(12)
        h = np.ascontiguousarray(hp_W, np.float64).reshape(-1, 4)
        n = len(h)
        proj = np.zeros((max(n, 1), 2), np.float64)
        status = np.zeros(max(n, 1), np.uint8)
        vis = np.zeros(max(n, 1), np.uint8)
        self.L.port_project_landmarks.argtypes = [_vp, _vp, _vp, _i, _d, _vp, _vp, _vp]
        self.L.port_project_landmarks(C.byref(cam), _ptr(T), _ptr(h), n, float(repr_threshold), _ptr(proj), _ptr(status),
                                      _ptr(vis))
        return proj[:n], status[:n], vis[:n]

    def prepare_landmarks(self, views, hp_W, quality, visible, obs_offset, obs_view, obs_kp, r_WC1, focal_sum,
                          loop_closure=False):
        Cw = np.ascontiguousarray(views["C_WC"], np.float64).reshape(-1, 9)
        rw = np.ascontiguousarray(views["r_WC"], np.float64).reshape(-1, 3)
        row0 = np.ascontiguousarray(views["row0"], np.int64)
        desc = np.ascontiguousarray(views["desc"], np.uint8).reshape(-1, 48)
        eC = np.ascontiguousarray(views["e_C"], np.float64).reshape(-1, 3)
        hp = np.ascontiguousarray(hp_W, np.float64).reshape(-1, 4)
        L = len(hp)
        q = np.ascontiguousarray(quality, np.float64)
        vis = np.ascontiguousarray(visible, np.uint8)
        off = np.ascontiguousarray(obs_offset, np.int64)
        ov = np.ascontiguousarray(obs_view, np.int32)
        ok = np.ascontiguousarray(obs_kp, np.int32)
        r1 = np.ascontiguousarray(r_WC1, np.float64).reshape(3)
        cap = max(3 * L, 1)
        out = dict(is3d=np.zeros(max(L, 1), np.uint8), rows=np.zeros(max(L, 1), np.int32), pool=np.zeros((cap, 48), np.uint8),
                   owner=np.zeros(cap, np.int32), e_W=np.zeros((cap, 3)), r_W=np.zeros((cap, 3)),
                   filled=np.zeros(cap, np.uint8), row_obs=np.zeros(cap, np.int32))
        f = self.L.port_prepare_landmarks
        f.argtypes = [_i, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _d, _i, _vp, _vp, _vp, _vp, _vp,
                      _vp, _vp, _vp]
        n = f(len(Cw), _ptr(Cw), _ptr(rw), _ptr(row0), _ptr(desc), _ptr(eC), L, _ptr(hp), _ptr(q), _ptr(vis), _ptr(off),
              _ptr(ov), _ptr(ok), _ptr(r1), float(focal_sum), int(bool(loop_closure)), _ptr(out["is3d"]), _ptr(out["rows"]),
              _ptr(out["pool"]), _ptr(out["owner"]), _ptr(out["e_W"]), _ptr(out["r_W"]), _ptr(out["filled"]),
              _ptr(out["row_obs"]))
        res = {k: (v[:L] if k in ("is3d", "rows") else v[:n]) for k, v in out.items()}
        res["nm"] = int(n)
        return res

    def vocabulary_transform(self, node_id, parent_id, descriptors, weights, word_of_node, depth, features, levelsup=0):
        """Same inputs as okvis2x_b200.Vocabulary (file order); the tree is flattened here the way load() builds it."""
        nid = np.asarray(node_id, np.int64)
        pid = np.asarray(parent_id, np.int64)
        N = len(nid) + 1
        kids = [[] for _ in range(N)]
        for i in range(len(nid)):
            kids[pid[i]].append(int(nid[i]))
        off = np.zeros(N + 1, np.int32)
        off[1:] = np.cumsum([len(k) for k in kids])
        children = np.array([c for k in kids for c in k], np.int32)
        desc = np.zeros((N, 48), np.uint8)
        desc[nid] = np.asarray(descriptors, np.uint8).reshape(-1, 48)
        w = np.zeros(N, np.float64)
        w[nid] = weights
        word = np.full(N, -1, np.int32)
        word[nid] = word_of_node
        f = np.ascontiguousarray(features, np.uint8).reshape(-1, 48)
        n = len(f)
        wid = np.zeros(max(n, 1), np.uint32)
        wt = np.zeros(max(n, 1), np.float64)
        node = np.zeros(max(n, 1), np.uint32)
        fn = self.L.port_vocabulary_transform
        fn.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp]
        fn(_ptr(off), _ptr(children), _ptr(desc), _ptr(w), _ptr(word), _ptr(f), n, int(depth), int(levelsup), _ptr(wid),
           _ptr(wt), _ptr(node))
        return wid[:n], wt[:n], node[:n]

    def descriptors_to_hex(self, descriptors):
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 48)
        out = np.zeros((max(len(d), 1), 96), np.uint8)
        fn = self.L.port_descriptors_to_hex
        fn.argtypes = [_vp, _i, _vp]
        fn(_ptr(d), len(d), _ptr(out))
        return [bytes(r).decode("ascii") for r in out[:len(d)]]

    def descriptors_from_hex(self, strings):
        rows = [s.encode("latin-1") if isinstance(s, str) else bytes(s) for s in strings]
        h = np.frombuffer(b"".join(r[:96] for r in rows), np.uint8).copy() if rows else np.zeros(1, np.uint8)
        out = np.zeros((max(len(rows), 1), 48), np.uint8)
        fn = self.L.port_descriptors_from_hex
        fn.argtypes = [_vp, _i, _vp]
        fn(_ptr(h), len(rows), _ptr(out))
        return out[:len(rows)]

    def descriptor_mean(self, descriptors, group_start):
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 48)
        gs = np.ascontiguousarray(group_start, np.int32)
        n = len(gs) - 1
        out = np.zeros((max(n, 1), 48), np.uint8)
        fn = self.L.port_descriptor_mean
        fn.argtypes = [_vp, _vp, _i, _vp]
        fn(_ptr(d), _ptr(gs), n, _ptr(out))
        return out[:n]

    def keyframe_overlap(self, widths, heights, kp_offset, keypoints, matched):
        w = np.ascontiguousarray(widths, np.int32)
        h = np.ascontiguousarray(heights, np.int32)
        off = np.ascontiguousarray(kp_offset, np.int64)
        kp = np.ascontiguousarray(keypoints)
        m = np.ascontiguousarray(matched, np.uint8)
        inter = np.zeros(max(len(w), 1), np.int32)
        uni = np.zeros(max(len(w), 1), np.int32)
        self.L.port_keyframe_overlap.argtypes = [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
        self.L.port_keyframe_overlap(len(w), _ptr(w), _ptr(h), _ptr(off), _ptr(kp), _ptr(m), _ptr(inter), _ptr(uni))
        return inter[:len(w)], uni[:len(w)]

    def hamming_best(self, queries, mp, threshold, index_offset=0, nthreads=1):
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 48)
        m = np.ascontiguousarray(mp, np.uint8).reshape(-1, 48)
        best = np.zeros(len(q), np.uint64)
        self.L.port_hamming_best(_ptr(q), len(q), _ptr(m), len(m), int(threshold), int(index_offset), _ptr(best),
                                 int(nthreads))
        return best

    def match_to_map(self, queries, kp_xy, use, mp, owner, proj_xy, is3d, thresh_sq, threshold, num_slices):
        """All thread slices of matchToMapByThread, run one after another (they touch disjoint k)."""
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 48)
        kp = np.ascontiguousarray(kp_xy, np.float64).reshape(-1, 2)
        us = np.ascontiguousarray(use, np.uint8)
        m = np.ascontiguousarray(mp, np.uint8).reshape(-1, 48)
        pr = np.ascontiguousarray(proj_xy, np.float64).reshape(-1, 2)
        L = len(pr)
        owner = np.asarray(owner, np.int64)
        assert np.all(np.diff(owner) >= 0)
        row_start = np.searchsorted(owner, np.arange(L + 1)).astype(np.int32)
        i3 = np.ascontiguousarray(is3d, np.uint8) if is3d is not None else None
        nk = len(q)
        dist = np.full(nk, float(threshold), np.float64)
        ids = np.full(nk, -1, np.int32)
        ctrs = np.zeros(num_slices, np.uint64)
        errs = np.zeros(num_slices, np.float64)
        for t in range(num_slices):
            self.L.port_match_to_map_thread(t, num_slices, _ptr(q), _ptr(kp), _ptr(us), nk, _ptr(m), _ptr(row_start), L,
                                            _ptr(pr), _ptr(i3) if i3 is not None else None, float(thresh_sq),
                                            _ptr(dist), _ptr(ids), _ptr(ctrs), _ptr(errs))
        return dist, ids, ctrs, errs

    def match_to_map_uninitialised(self, queries, kp_e_W, use, previous_owner, mp, owner, map_e_W, map_r_W,
                                   num_landmarks, r_WC, focal, threshold, num_slices):
        """All thread slices of matchToMapByThreadUnitialised, one after another."""
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 48)
        ke = np.ascontiguousarray(kp_e_W, np.float64).reshape(-1, 3)
        us = np.ascontiguousarray(use, np.uint8)
        pv = np.ascontiguousarray(previous_owner, np.int32) if previous_owner is not None else None
        m = np.ascontiguousarray(mp, np.uint8).reshape(-1, 48)
        me = np.ascontiguousarray(map_e_W, np.float64).reshape(-1, 3)
        mr = np.ascontiguousarray(map_r_W, np.float64).reshape(-1, 3)
        rc = np.ascontiguousarray(r_WC, np.float64)
        owner = np.asarray(owner, np.int64)
        row_start = np.searchsorted(owner, np.arange(num_landmarks + 1)).astype(np.int32)
        nk = len(q)
        dist = np.full(nk, float(threshold), np.float64)
        ids = np.full(nk, -1, np.int32)
        hps = np.zeros((nk, 4), np.float64)
        hw = np.zeros(nk, np.uint8)
        ctrs = np.zeros(num_slices, np.uint64)
        for t in range(num_slices):
            self.L.port_match_uninit_thread(t, num_slices, _ptr(q), _ptr(ke), _ptr(us), _ptr(pv) if pv is not None else None,
                                            nk, _ptr(m), _ptr(row_start), int(num_landmarks), _ptr(me), _ptr(mr),
                                            _ptr(rc), float(focal), _ptr(dist), _ptr(ids), _ptr(hps), _ptr(hw),
                                            _ptr(ctrs))
        return dist, ids, hps, hw, ctrs

    @staticmethod
    def _view(v, keep):
        s = PortStereoView()
        desc = np.ascontiguousarray(v["desc"], np.uint8).reshape(-1, 48)
        size = np.ascontiguousarray(v["size"], np.float64)
        e_C = np.ascontiguousarray(v["e_C"], np.float64).reshape(-1, 3)
        bpv = np.ascontiguousarray(v["bp_valid"], np.uint8)
        keep.extend([desc, size, e_C, bpv])
        s.desc, s.size, s.e_C, s.bp_valid = desc.ctypes.data, size.ctypes.data, e_C.ctypes.data, bpv.ctypes.data
        s.n = len(desc)
        s.C[:] = list(np.asarray(v["C_WC"], np.float64).reshape(9))
        s.r[:] = list(np.asarray(v["r_WC"], np.float64).reshape(3))
        s.f = float(v["f"])
        return s

    def match_motion_stereo(self, view0, use0, view1, threshold):
        keep = []
        s0, s1 = self._view(view0, keep), self._view(view1, keep)
        u0 = np.ascontiguousarray(use0, np.uint8)
        n0 = s0.n
        dist = np.empty(n0, np.float64)
        k1 = np.empty(n0, np.int32)
        hp = np.empty((n0, 4), np.float64)
        init = np.empty(n0, np.uint8)
        qual = np.empty(n0, np.float64)
        self.L.port_match_motion_stereo(C.byref(s0), _ptr(u0), C.byref(s1), float(threshold), _ptr(dist), _ptr(k1),
                                        _ptr(hp), _ptr(init), _ptr(qual))
        return dist, k1, hp, init, qual

    def match_stereo(self, view0, view1, threshold):
        keep = []

        def mk(v):
            s = PortStereoView()
            desc = np.ascontiguousarray(v["desc"], np.uint8).reshape(-1, 48)
            size = np.ascontiguousarray(v["size"], np.float64)
            e_C = np.ascontiguousarray(v["e_C"], np.float64).reshape(-1, 3)
            bpv = np.ascontiguousarray(v["bp_valid"], np.uint8)
            keep.extend([desc, size, e_C, bpv])
            s.desc, s.size, s.e_C, s.bp_valid = desc.ctypes.data, size.ctypes.data, e_C.ctypes.data, bpv.ctypes.data
            s.n = len(desc)
            s.C[:] = list(np.asarray(v["C_WC"], np.float64).reshape(9))
            s.r[:] = list(np.asarray(v["r_WC"], np.float64).reshape(3))
            s.f = float(v["f"])
            return s

        s0, s1 = mk(view0), mk(view1)
        n0 = s0.n
        dist = np.empty(n0, np.float64)
        k1 = np.empty(n0, np.int32)
        hp = np.empty((n0, 4), np.float64)
        init = np.empty(n0, np.uint8)
        self.L.port_match_stereo(C.byref(s0), C.byref(s1), float(threshold), _ptr(dist), _ptr(k1), _ptr(hp), _ptr(init))
        return dist, k1, hp, init
