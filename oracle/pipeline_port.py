"""TEST INFRASTRUCTURE ONLY: the CPU frame pipeline the B200 path is checked and timed against.

Per camera image of a multiframe, in the reference's order (Frontend::detectAndDescribe, then
Frontend::dataAssociationAndInitialization, okvis_frontend/src/Frontend.cpp:204-262, 674-1143):
  detect + describe   external/brisk compiled unchanged (oracle/_ref)
  back-projections    Frame::computeBackProjections                      (oracle/port/camera.cc)
  landmark stage      projection + FoV test + pruning + pool packing     (oracle/port/camera.cc, Frontend.cpp:1376-1500)
  matchToMapByThread / matchToMapByThreadUnitialised                     (oracle/port/matchers.cc)
  matchMotionStereo worker, matchStereo loops                            (oracle/port/matchers.cc)
Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import threading

import numpy as np

import refbind
from refbind import _ptr


def normalized_rows(e, variant=0):
    """Eigen's v.normalized() row-wise: v / sqrt((x x + y y) + z z), element-wise division. variant: the
    PORT_EIGEN_VARIANT bits of the sensitivity builds (bit 0: other dot order, bit 1: multiply by the reciprocal)."""
    e = np.asarray(e, np.float64).reshape(-1, 3)
    if variant & 1:
        n2 = e[:, 0] * e[:, 0] + (e[:, 1] * e[:, 1] + e[:, 2] * e[:, 2])
    else:
        n2 = (e[:, 0] * e[:, 0] + e[:, 1] * e[:, 1]) + e[:, 2] * e[:, 2]
    out = e.copy()
    nz = n2 > 0.0
    if variant & 2:
        out[nz] = e[nz] * (1.0 / np.sqrt(n2[nz]))[:, None]
    else:
        out[nz] = e[nz] / np.sqrt(n2[nz])[:, None]
    return out


def rotate_rows(C, e, variant=0):
    """C * e per row with Eigen's 3x3 * 3x1 order ((c0 x + c1 y) + c2 z); variant bit 3: c0 x + (c1 y + c2 z)."""
    C = np.asarray(C, np.float64).reshape(3, 3)
    if variant & 8:
        return np.stack([C[r, 0] * e[:, 0] + (C[r, 1] * e[:, 1] + C[r, 2] * e[:, 2]) for r in range(3)], 1)
    return np.stack([(C[r, 0] * e[:, 0] + C[r, 1] * e[:, 1]) + C[r, 2] * e[:, 2] for r in range(3)], 1)


class CpuPipeline:
    def __init__(self, ref=None, port=None):
        self.ref = ref or refbind.Ref()
        self.port = port or refbind.Port()

    def detect_describe(self, ext, img, det):
        kp = self.ref.detect(img, *det)
        return ext.describe(img, kp)

    def landmark_stage(self, cam, pose, kp, desc, e_C, valid, pool_args, threshold=60.0, slices=4, uninitialised=True,
                       use_imu=True, loop_closure=False, threaded=False):
        """pose = (C_WC [3,3], r_WC [3], T_CW [12]); pool_args = (views, hp_W, quality, obs_offset, obs_view, obs_kp).
        Returns dict(prep=..., a14=(dist, owner, ctrs, errs), a15=(dist, owner, hps, hw, ctrs))."""
        C_WC, r_WC, T_CW = pose
        views, hp_W, quality, off, ov, ok = pool_args
        f = 0.5 * (cam.fu + cam.fv)
        repr_thr = 3.0 + f * 0.06 if use_imu else 3.0 + f * 0.34
        proj, status, vis = self.port.project_landmarks(cam, T_CW, hp_W, repr_thr)
        prep = self.port.prepare_landmarks(views, hp_W, quality, vis, off, ov, ok, r_WC, cam.fu + cam.fv, loop_closure)
        kp_xy = np.stack([kp["x"], kp["y"]], 1).astype(np.float64)
        n = len(kp)
        a14 = self._a14(desc, kp_xy, prep, proj, repr_thr * repr_thr, threshold, slices, threaded)
        out = dict(proj=proj, visible=vis, prep=prep, a14=a14)
        if uninitialised:
            vb = getattr(self.port, "variant", 0)
            e_W = rotate_rows(C_WC, normalized_rows(e_C, vb), vb)
            use = (np.asarray(valid) != 0) & (a14[1] < 0)
            # rows of 3D landmarks are skipped by the uninitialised matcher (Frontend.cpp:1878-1880)
            rows = np.nonzero(prep["is3d"][prep["owner"]] == 0)[0]
            out["a15"] = self._a15(desc, e_W, use.astype(np.uint8), prep, rows, len(hp_W), r_WC, f, threshold, slices,
                                   threaded)
        return out

    @staticmethod
    def _run(fn, slices, threaded):
        """The reference's num_matching_threads workers (Frontend.cpp:1519-1536, 1631-1645) or the slices in turn."""
        if threaded and slices > 1:
            th = [threading.Thread(target=fn, args=(t,)) for t in range(slices)]
            for x in th:
                x.start()
            for x in th:
                x.join()
        else:
            for t in range(slices):
                fn(t)

    def _a14(self, desc, kp_xy, prep, proj, thresh_sq, threshold, slices, threaded):
        q = np.ascontiguousarray(desc, np.uint8).reshape(-1, 48)
        kp = np.ascontiguousarray(kp_xy, np.float64)
        m = np.ascontiguousarray(prep["pool"], np.uint8).reshape(-1, 48)
        pr = np.ascontiguousarray(proj, np.float64)
        L = len(pr)
        row_start = np.searchsorted(prep["owner"].astype(np.int64), np.arange(L + 1)).astype(np.int32)
        i3 = np.ascontiguousarray(prep["is3d"], np.uint8)
        nk = len(q)
        us = np.ones(nk, np.uint8)
        dist = np.full(nk, float(threshold), np.float64)
        ids = np.full(nk, -1, np.int32)
        ctrs = np.zeros(slices, np.uint64)
        errs = np.zeros(slices, np.float64)
        self._run(lambda t: self.port.L.port_match_to_map_thread(
            t, slices, _ptr(q), _ptr(kp), _ptr(us), nk, _ptr(m), _ptr(row_start), L, _ptr(pr), _ptr(i3), float(thresh_sq),
            _ptr(dist), _ptr(ids), _ptr(ctrs), _ptr(errs)), slices, threaded)
        return dist, ids, ctrs, errs

    def _a15(self, desc, e_W, use, prep, rows, L, r_WC, focal, threshold, slices, threaded):
        q = np.ascontiguousarray(desc, np.uint8).reshape(-1, 48)
        ke = np.ascontiguousarray(e_W, np.float64)
        m = np.ascontiguousarray(prep["pool"][rows])
        me = np.ascontiguousarray(prep["e_W"][rows])
        mr = np.ascontiguousarray(prep["r_W"][rows])
        rc = np.ascontiguousarray(r_WC, np.float64)
        row_start = np.searchsorted(prep["owner"][rows].astype(np.int64), np.arange(L + 1)).astype(np.int32)
        nk = len(q)
        dist = np.full(nk, float(threshold), np.float64)
        ids = np.full(nk, -1, np.int32)
        hps = np.zeros((nk, 4), np.float64)
        hw = np.zeros(nk, np.uint8)
        ctrs = np.zeros(slices, np.uint64)
        self._run(lambda t: self.port.L.port_match_uninit_thread(
            t, slices, _ptr(q), _ptr(ke), _ptr(use), None, nk, _ptr(m), _ptr(row_start), int(L), _ptr(me), _ptr(mr),
            _ptr(rc), float(focal), _ptr(dist), _ptr(ids), _ptr(hps), _ptr(hw), _ptr(ctrs)), slices, threaded)
        return dist, ids, hps, hw, ctrs

    def stereo(self, v0, v1, threshold=60.0):
        return self.port.match_stereo(v0, v1, threshold)

    def motion(self, kf_view, use0, cur_view, unmatched, threshold=60.0):
        """cur_view restricted to the keypoints flagged in `unmatched` (the reference packs them); k1 is mapped back to
        the index inside the current image."""
        idx = np.nonzero(unmatched)[0]
        sub = dict(cur_view, desc=cur_view["desc"][idx], size=cur_view["size"][idx], e_C=cur_view["e_C"][idx],
                   bp_valid=cur_view["bp_valid"][idx])
        dist, k1, hp, init, q = self.port.match_motion_stereo(kf_view, use0, sub, threshold)
        k1 = np.where(k1 >= 0, idx[np.clip(k1, 0, max(len(idx) - 1, 0))] if len(idx) else -1, -1).astype(np.int32)
        return dist, k1, hp, init, q
