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
import numpy as np

import refbind


def normalized_rows(e):
    """Eigen's v.normalized() row-wise: v / sqrt((x x + y y) + z z), element-wise division."""
    e = np.asarray(e, np.float64).reshape(-1, 3)
    n2 = (e[:, 0] * e[:, 0] + e[:, 1] * e[:, 1]) + e[:, 2] * e[:, 2]
    out = e.copy()
    nz = n2 > 0.0
    out[nz] = e[nz] / np.sqrt(n2[nz])[:, None]
    return out


def rotate_rows(C, e):
    """C * e per row with Eigen's 3x3 * 3x1 order ((c0 x + c1 y) + c2 z)."""
    C = np.asarray(C, np.float64).reshape(3, 3)
    return np.stack([(C[r, 0] * e[:, 0] + C[r, 1] * e[:, 1]) + C[r, 2] * e[:, 2] for r in range(3)], 1)


class CpuPipeline:
    def __init__(self, ref=None, port=None):
        self.ref = ref or refbind.Ref()
        self.port = port or refbind.Port()

    def detect_describe(self, ext, img, det):
        kp = self.ref.detect(img, *det)
        return ext.describe(img, kp)

    def landmark_stage(self, cam, pose, kp, desc, e_C, valid, pool_args, threshold=60.0, slices=4, uninitialised=True,
                       use_imu=True, loop_closure=False):
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
        a14 = self.port.match_to_map(desc, kp_xy, np.ones(n, np.uint8), prep["pool"], prep["owner"], proj, prep["is3d"],
                                     repr_thr * repr_thr, threshold, slices)
        out = dict(proj=proj, visible=vis, prep=prep, a14=a14)
        if uninitialised:
            e_W = rotate_rows(C_WC, normalized_rows(e_C))
            use = (np.asarray(valid) != 0) & (a14[1] < 0)
            # rows of 3D landmarks are skipped by the uninitialised matcher (Frontend.cpp:1878-1880)
            sel = prep["is3d"][prep["owner"]] == 0
            rows = np.nonzero(sel)[0]
            out["a15"] = self.port.match_to_map_uninitialised(
                desc, e_W, use.astype(np.uint8), None, prep["pool"][rows], prep["owner"][rows], prep["e_W"][rows],
                prep["r_W"][rows], len(hp_W), r_WC, f, threshold, slices)
        return out

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
