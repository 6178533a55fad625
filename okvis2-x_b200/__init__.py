"""okvis2-x_b200: host-side mirror (Python, ctypes) of the okvis_cuda C ABI.

The product is `libokvis_cuda.so` (hand-written sm_100a kernels behind include/okvis_cuda.h). This package
only loads it and mirrors the reference's call surface for the hot path:

  okvis::MultiFrame::detect / describe (okvis_cv/include/okvis/MultiFrame.hpp:158-164)   -> Camera.detect / describe
  Frontend::detectAndDescribe (okvis_frontend/src/Frontend.cpp:204-262)                  -> Camera.detect_describe
  Frontend matching loops (Frontend.cpp:316-345, 1744-1956, 2259-2316)                   -> Context.match_*

There is no CPU fallback: a missing library or a missing GPU raises.
The directory name has a hyphen, so import it through `load()` in __graft_entry__.py (alias `okvis2x_b200`).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libokvis_cuda.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
PWS_DTYPE = np.dtype([("score", "<i4"), ("x", "<u2"), ("y", "<u2")])
NO_MATCH = np.uint64(0xFFFFFFFFFFFFFFFF)

class JobView(C.Structure):
    _fields_ = [("base", C.c_void_p), ("stride", C.c_size_t), ("off_dist", C.c_size_t), ("off_k1", C.c_size_t),
                ("off_hp", C.c_size_t), ("off_initialisable", C.c_size_t), ("off_quality", C.c_size_t), ("n", C.c_int),
                ("rows", C.c_int)]


class BatchView(C.Structure):
    """okc_batch_view (include/okvis_cuda.h)."""
    _fields_ = [("n_slots", C.c_int), ("max_keypoints", C.c_int), ("num_slices", C.c_int), ("counters", C.c_void_p),
                ("slots", C.c_void_p), ("slot_stride", C.c_size_t), ("off_keypoints", C.c_size_t),
                ("off_descriptors", C.c_size_t), ("off_best_dist", C.c_size_t), ("off_best_owner", C.c_size_t),
                ("off_stats", C.c_size_t), ("map_match", C.c_int), ("uninit", C.c_void_p), ("uninit_stride", C.c_size_t),
                ("u_off_dist", C.c_size_t), ("u_off_owner", C.c_size_t), ("u_off_hps", C.c_size_t),
                ("u_off_written", C.c_size_t), ("uninit_ctrs", C.c_void_p), ("uninit_ctrs_stride", C.c_int),
                ("stereo", JobView), ("motion", JobView), ("motion_n0", C.c_void_p)]


EXPORTS = [
    "okc_last_error_string", "okc_version", "okc_context_create", "okc_context_destroy", "okc_context_device",
    "okc_camera_create", "okc_camera_destroy", "okc_detect", "okc_describe", "okc_detect_describe",
    "okc_detect_describe_enqueue", "okc_detect_describe_finish", "okc_debug_scores", "okc_debug_candidates",
    "okc_match_batched", "okc_descriptor_set_create", "okc_descriptor_set_destroy", "okc_match_sets",
    "okc_match_to_map", "okc_match_stereo", "okc_match_to_map_uninitialised", "okc_match_motion_stereo",
    "okc_detect_describe_enqueue_device", "okc_map_window_create", "okc_map_window_destroy",
    "okc_camera_match_to_map_enqueue", "okc_camera_frame_finish", "okc_camera_stream", "okc_debug_kernel_timing",
    "okc_debug_kernel_times", "okc_batch_create", "okc_batch_destroy", "okc_batch_size", "okc_batch_set_graph", "okc_batch_view_results", "okc_batch_transfer_bytes", "okc_match_device_no_alloc", "okc_batch_enqueue",
    "okc_batch_finish", "okc_batch_slot_frame", "okc_batch_slot_matches", "okc_batch_read_all", "okc_batch_stream",
    "okc_debug_batch_kernel_timing", "okc_debug_batch_kernel_times",
    "okc_camera_set_geometry", "okc_batch_set_poses", "okc_batch_slot_back_projections", "okc_landmark_pool_create",
    "okc_landmark_pool_update", "okc_landmark_pool_destroy", "okc_batch_match_landmarks_enqueue",
    "okc_batch_slot_uninitialised", "okc_batch_slot_landmarks", "okc_batch_match_stereo_enqueue",
    "okc_batch_pair_stereo", "okc_keyframe_create", "okc_keyframe_destroy", "okc_batch_retain",
    "okc_keyframe_set_view", "okc_batch_match_motion_enqueue", "okc_batch_job_motion", "okc_batch_read_stages",
    "okc_packed_view_bytes", "okc_batch_pack_views", "okc_batch_match_stereo_packed_enqueue", "okc_match_sets_ordered",
]
STAGE_UNINITIALISED, STAGE_NO_IMU, STAGE_LOOP_CLOSURE = 1, 2, 4


class OkcError(RuntimeError):
    """A non-zero status of the C ABI; `code` is the OKC_ERR_* value (1 invalid, 2 CUDA, 3 capacity)."""
    def __init__(self, message, code=1):
        super().__init__(message)
        self.code = code


class DetectParams(C.Structure):
    _fields_ = [("uniformity_radius", C.c_double), ("octaves", C.c_uint64), ("absolute_threshold", C.c_double),
                ("max_num_keypoints", C.c_uint64)]


class Pinhole(C.Structure):
    """okc_pinhole: cameras::PinholeCamera<D> intrinsics. distortion: 0 none, 1 radial-tangential (k1, k2, p1, p2),
    2 equidistant (k1..k4)."""
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("fu", C.c_double), ("fv", C.c_double), ("cu", C.c_double),
                ("cv", C.c_double), ("distortion", C.c_int), ("k", C.c_double * 8)]

    @staticmethod
    def make(width, height, f, c, distortion, k):
        p = Pinhole(int(width), int(height), float(f[0]), float(f[1]), float(c[0]), float(c[1]),
                    {"none": 0, "radialtangential": 1, "equidistant": 2}.get(distortion, distortion))
        for i, v in enumerate(k):
            p.k[i] = float(v)
        return p


class LandmarkViews(C.Structure):
    _fields_ = [("n_views", C.c_int), ("C_WC", C.c_void_p), ("r_WC", C.c_void_p), ("row0", C.c_void_p),
                ("desc", C.c_void_p), ("e_C", C.c_void_p), ("n_rows", C.c_int64)]


class Vocabulary:
    """DBoW2 vocabulary tree on the device (okc_vocabulary_*). Nodes in file order: node_id (1..n), parent_id,
    descriptors [n,48], weights [n], word_of_node [n] (-1 for inner nodes); depth = L of the file."""

    def __init__(self, ctx, node_id, parent_id, descriptors, weights, word_of_node, depth):
        nid = np.ascontiguousarray(node_id, np.int32)
        pid = np.ascontiguousarray(parent_id, np.int32)
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 48)
        w = np.ascontiguousarray(weights, np.float64)
        wo = np.ascontiguousarray(word_of_node, np.int32)
        self.h = C.c_void_p()
        self.depth = int(depth)
        _check(lib().okc_vocabulary_create(ctx.h, len(nid), _p(nid), _p(pid), _p(d), _p(w), _p(wo), self.depth,
                                           C.byref(self.h)), "okc_vocabulary_create")

    def transform(self, features, levelsup=0):
        f = np.ascontiguousarray(features, np.uint8).reshape(-1, 48)
        n = len(f)
        word = np.zeros(max(n, 1), np.uint32)
        weight = np.zeros(max(n, 1), np.float64)
        node = np.zeros(max(n, 1), np.uint32)
        _check(lib().okc_vocabulary_transform(self.h, _p(f), n, self.depth, int(levelsup), _p(word), _p(weight), _p(node)),
               "okc_vocabulary_transform")
        return word[:n], weight[:n], node[:n]

    def close(self):
        if self.h:
            lib().okc_vocabulary_destroy(self.h)
            self.h = C.c_void_p()


class DescribeParams(C.Structure):
    _fields_ = [("rotation_invariant", C.c_int), ("scale_invariant", C.c_int), ("extraction_direction", C.c_float * 3)]


class StereoView(C.Structure):
    _fields_ = [("desc", C.c_void_p), ("size", C.c_void_p), ("e_C", C.c_void_p), ("bp_valid", C.c_void_p),
                ("n", C.c_int), ("C_WC", C.c_double * 9), ("r_WC", C.c_double * 3), ("f", C.c_double)]


_lib = None


def lib():
    """Load libokvis_cuda.so (fails loudly when it has not been built: run __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OkcError("%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        vp, i, u32, u64, i64, d, f = C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_int64, C.c_double, C.c_float
        L.okc_last_error_string.restype = C.c_char_p
        L.okc_context_create.argtypes = [i, C.POINTER(vp)]
        L.okc_context_destroy.argtypes = [vp]
        L.okc_camera_create.argtypes = [vp, i, i, vp, vp, f, i, C.POINTER(vp)]
        L.okc_camera_destroy.argtypes = [vp]
        L.okc_detect.argtypes = [vp, vp, C.POINTER(DetectParams), vp, i, C.POINTER(i)]
        L.okc_describe.argtypes = [vp, vp, C.POINTER(DescribeParams), vp, C.POINTER(i), vp]
        L.okc_detect_describe.argtypes = [vp, vp, C.POINTER(DetectParams), C.POINTER(DescribeParams), vp, i,
                                          C.POINTER(i), vp]
        L.okc_detect_describe_enqueue.argtypes = [vp, vp, C.POINTER(DetectParams), C.POINTER(DescribeParams)]
        L.okc_detect_describe_finish.argtypes = [vp, vp, i, C.POINTER(i), vp]
        L.okc_debug_scores.argtypes = [vp, vp, vp]
        L.okc_debug_candidates.argtypes = [vp, vp, i, vp, i, C.POINTER(i)]
        L.okc_debug_host_tables.argtypes = [vp, C.POINTER(i), C.POINTER(i), vp, vp, vp]
        L.okc_debug_set_match_prefix.argtypes = [i]
        L.okc_debug_set_select_wave.argtypes = [i]
        L.okc_debug_set_harris_rows.argtypes = [i]
        L.okc_back_project.argtypes = [vp, vp, vp, i, vp, vp, C.POINTER(i)]
        L.okc_keyframe_overlap.argtypes = [vp, i, vp, vp, vp, vp, vp, vp, vp]
        L.okc_vocabulary_create.argtypes = [vp, i, vp, vp, vp, vp, vp, i, C.POINTER(vp)]
        L.okc_vocabulary_destroy.argtypes = [vp]
        L.okc_vocabulary_transform.argtypes = [vp, vp, i, i, i, vp, vp, vp]
        L.okc_descriptors_to_hex.argtypes = [vp, vp, i, vp]
        L.okc_descriptors_from_hex.argtypes = [vp, vp, i, vp]
        L.okc_descriptor_mean.argtypes = [vp, vp, vp, i, vp]
        L.okc_prepare_landmarks.argtypes = [vp, vp, i, vp, vp, vp, vp, vp, vp, vp, C.c_double, i, vp, vp, C.POINTER(i), vp,
                                            vp, vp, vp, vp, vp]
        L.okc_project_landmarks.argtypes = [vp, vp, vp, vp, i, C.c_double, vp, vp, vp]
        L.okc_debug_set_match_binned.argtypes = [i]
        L.okc_debug_batch_counter.argtypes = [vp, i, i, C.POINTER(i)]
        L.okc_debug_alloc_counts.argtypes = [C.POINTER(u64), C.POINTER(u64)]
        L.okc_debug_std_sort.argtypes = [vp, vp, i, vp]
        L.okc_match_batched.argtypes = [vp, vp, i, vp, i64, u32, u64, vp]
        L.okc_descriptor_set_create.argtypes = [vp, vp, i64, C.POINTER(vp)]
        L.okc_descriptor_set_destroy.argtypes = [vp]
        L.okc_descriptor_set_device_ptr.argtypes = [vp]
        L.okc_descriptor_set_device_ptr.restype = vp
        L.okc_match_sets.argtypes = [vp, vp, vp, u32, u64, vp, vp, vp]
        L.okc_match_to_map.argtypes = [vp, vp, vp, vp, i, vp, vp, i, vp, i, d, d, i, vp, vp, vp, vp]
        L.okc_match_stereo.argtypes = [vp, C.POINTER(StereoView), C.POINTER(StereoView), d, vp, vp, vp, vp]
        L.okc_match_motion_stereo.argtypes = [vp, C.POINTER(StereoView), vp, C.POINTER(StereoView), d, vp, vp, vp, vp,
                                              vp]
        L.okc_match_to_map_uninitialised.argtypes = [vp, vp, vp, vp, vp, i, vp, vp, vp, vp, i, vp, d, d, i, vp, vp, vp,
                                                     vp, vp]
        L.okc_detect_describe_enqueue_device.argtypes = [vp, vp, C.POINTER(DetectParams), C.POINTER(DescribeParams)]
        L.okc_map_window_create.argtypes = [vp, vp, vp, i, i, C.POINTER(vp)]
        L.okc_map_window_destroy.argtypes = [vp]
        L.okc_camera_match_to_map_enqueue.argtypes = [vp, vp, vp, i, d, d, i]
        L.okc_camera_frame_finish.argtypes = [vp, vp, vp, i, C.POINTER(i), vp, vp, vp, vp, vp]
        L.okc_camera_stream.argtypes = [vp]
        L.okc_camera_stream.restype = vp
        L.okc_debug_kernel_timing.argtypes = [vp, i]
        L.okc_debug_kernel_times.argtypes = [vp, vp, i, vp, i, C.POINTER(i)]
        L.okc_batch_create.argtypes = [vp, vp, i, C.POINTER(vp)]
        L.okc_batch_destroy.argtypes = [vp]
        L.okc_batch_size.argtypes = [vp]
        L.okc_batch_set_graph.argtypes = [vp, C.c_int]
        L.okc_batch_view_results.argtypes = [vp, vp]
        L.okc_batch_transfer_bytes.argtypes = [vp, vp, vp]
        L.okc_match_device_no_alloc.argtypes = [C.c_int, vp, C.c_int, vp, C.c_int64, C.c_uint32, C.c_uint64, vp, vp]
        L.okc_batch_enqueue.argtypes = [vp, vp, i, C.POINTER(DetectParams), vp, vp, vp, i, vp, d, i]
        L.okc_batch_read_all.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp]
        L.okc_batch_finish.argtypes = [vp]
        L.okc_batch_slot_frame.argtypes = [vp, i, vp, i, C.POINTER(i), vp]
        L.okc_batch_slot_matches.argtypes = [vp, i, vp, vp, vp, vp]
        L.okc_batch_stream.argtypes = [vp]
        L.okc_batch_stream.restype = vp
        L.okc_debug_batch_kernel_timing.argtypes = [vp, i]
        L.okc_debug_batch_kernel_times.argtypes = [vp, vp, i, vp, i, C.POINTER(i)]
        L.okc_camera_set_geometry.argtypes = [vp, vp]
        L.okc_batch_set_poses.argtypes = [vp, vp]
        L.okc_batch_slot_back_projections.argtypes = [vp, i, vp, vp, C.POINTER(i)]
        L.okc_landmark_pool_create.argtypes = [vp, vp, i, vp, vp, vp, vp, vp, C.POINTER(vp)]
        L.okc_landmark_pool_update.argtypes = [vp, vp, vp, vp, vp]
        L.okc_landmark_pool_destroy.argtypes = [vp]
        L.okc_batch_match_landmarks_enqueue.argtypes = [vp, vp, d, i, i]
        L.okc_batch_slot_uninitialised.argtypes = [vp, i, vp, vp, vp, vp, vp]
        L.okc_batch_slot_landmarks.argtypes = [vp, i, vp, vp, vp, vp, C.POINTER(i), vp, vp]
        L.okc_batch_match_stereo_enqueue.argtypes = [vp, i, vp, vp, d]
        L.okc_batch_pair_stereo.argtypes = [vp, i, vp, vp, vp, vp]
        L.okc_keyframe_create.argtypes = [vp, i, i, C.POINTER(vp)]
        L.okc_keyframe_destroy.argtypes = [vp]
        L.okc_batch_retain.argtypes = [vp, i, vp, vp]
        L.okc_keyframe_set_view.argtypes = [vp, i, vp, vp, d, vp, i]
        L.okc_batch_match_motion_enqueue.argtypes = [vp, i, vp, vp, vp, d, i]
        L.okc_batch_job_motion.argtypes = [vp, i, C.POINTER(i), vp, vp, vp, vp, vp]
        L.okc_batch_read_stages.argtypes = [vp] + [vp] * 15
        L.okc_packed_view_bytes.argtypes = [i]
        L.okc_packed_view_bytes.restype = C.c_size_t
        L.okc_batch_pack_views.argtypes = [vp, i, vp, vp]
        L.okc_batch_match_stereo_packed_enqueue.argtypes = [vp, i, vp, vp, vp, vp, vp, d]
        L.okc_match_sets_ordered.argtypes = [vp, vp, vp, u32, u64, vp, vp]
        _lib = L
    return _lib


def _check(rc, what):
    if rc != 0:
        raise OkcError("%s failed (code %d): %s" % (what, rc, lib().okc_last_error_string().decode()), rc)


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


def host_tables():
    """Pattern tables as built on the host at context creation (no GPU needed)."""
    pts = np.zeros((1024, 66, 3), np.float32)
    lp = np.zeros((856, 4), np.int32)
    sp = np.zeros((384, 2), np.uint8)
    lut = np.zeros((31, 31), np.float32)
    bs, bd = C.c_int(0), C.c_int(0)
    _check(lib().okc_debug_host_tables(_p(pts), C.byref(bs), C.byref(bd), _p(lp), _p(sp), _p(lut)), "host_tables")
    return dict(points=pts, basicscale=bs.value, border=bd.value, long_pairs=lp, short_pairs=sp, lut=lut)


class Context:
    def __init__(self, device=-1):
        self.h = C.c_void_p()
        _check(lib().okc_context_create(device, C.byref(self.h)), "okc_context_create")

    def close(self):
        if self.h:
            lib().okc_context_destroy(self.h)
            self.h = C.c_void_p()

    # -- camera geometry (Frame::computeBackProjections; FoV test of Frontend::matchToMap) --
    def back_project(self, cam, keypoints):
        kp = np.ascontiguousarray(keypoints, KP_DTYPE)
        n = len(kp)
        e = np.zeros((max(n, 1), 3), np.float64)
        valid = np.zeros(max(n, 1), np.uint8)
        nv = C.c_int(0)
        _check(lib().okc_back_project(self.h, C.byref(cam), _p(kp), n, _p(e), _p(valid), C.byref(nv)), "okc_back_project")
        return e[:n], valid[:n], nv.value

    def project_landmarks(self, cam, T_CW, hp_W, repr_threshold):
        T = np.ascontiguousarray(T_CW, np.float64).reshape(12)
        h = np.ascontiguousarray(hp_W, np.float64).reshape(-1, 4)
        n = len(h)
        proj = np.zeros((max(n, 1), 2), np.float64)
        status = np.zeros(max(n, 1), np.uint8)
        vis = np.zeros(max(n, 1), np.uint8)
        _check(lib().okc_project_landmarks(self.h, C.byref(cam), _p(T), _p(h), n, float(repr_threshold), _p(proj),
                                           _p(status), _p(vis)), "okc_project_landmarks")
        return proj[:n], status[:n], vis[:n]

    def prepare_landmarks(self, views, hp_W, quality, visible, obs_offset, obs_view, obs_kp, r_WC1, focal_sum,
                          loop_closure=False):
        """views: dict(C_WC [V,9], r_WC [V,3], row0 [V], desc [N,48], e_C [N,3]). Returns dict(is3d, rows, nm, pool,
        owner, e_W, r_W, filled, row_obs) (Frontend.cpp:1409-1500)."""
        Cw = np.ascontiguousarray(views["C_WC"], np.float64).reshape(-1, 9)
        rw = np.ascontiguousarray(views["r_WC"], np.float64).reshape(-1, 3)
        row0 = np.ascontiguousarray(views["row0"], np.int64)
        desc = np.ascontiguousarray(views["desc"], np.uint8).reshape(-1, 48)
        eC = np.ascontiguousarray(views["e_C"], np.float64).reshape(-1, 3)
        lv = LandmarkViews(len(Cw), Cw.ctypes.data, rw.ctypes.data, row0.ctypes.data, desc.ctypes.data, eC.ctypes.data,
                           len(desc))
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
        nm = C.c_int(0)
        _check(lib().okc_prepare_landmarks(self.h, C.byref(lv), L, _p(hp), _p(q), _p(vis), _p(off), _p(ov), _p(ok), _p(r1),
                                           float(focal_sum), int(bool(loop_closure)), _p(out["is3d"]), _p(out["rows"]),
                                           C.byref(nm), _p(out["pool"]), _p(out["owner"]), _p(out["e_W"]), _p(out["r_W"]),
                                           _p(out["filled"]), _p(out["row_obs"])), "okc_prepare_landmarks")
        n = nm.value
        res = {k: (v[:L] if k in ("is3d", "rows") else v[:n]) for k, v in out.items()}
        res["nm"] = n
        return res

    def keyframe_overlap(self, widths, heights, kp_offset, keypoints, matched):
        """Frontend::doWeNeedANewKeyframe masks: (intersection[v], union[v]) pixel counts per view."""
        w = np.ascontiguousarray(widths, np.int32)
        h = np.ascontiguousarray(heights, np.int32)
        off = np.ascontiguousarray(kp_offset, np.int64)
        kp = np.ascontiguousarray(keypoints, KP_DTYPE)
        m = np.ascontiguousarray(matched, np.uint8)
        n = len(w)
        inter = np.zeros(max(n, 1), np.int32)
        uni = np.zeros(max(n, 1), np.int32)
        _check(lib().okc_keyframe_overlap(self.h, n, _p(w), _p(h), _p(off), _p(kp), _p(m), _p(inter), _p(uni)),
               "okc_keyframe_overlap")
        return inter[:n], uni[:n]

    # -- formats either side of the descriptor (Component.cpp:228-236, 434-437; FBrisk.cpp:25-60) --
    def descriptors_to_hex(self, descriptors):
        """The 96-character text form of each descriptor in a saved map -> list of str."""
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 48)
        out = np.zeros((max(len(d), 1), 96), np.uint8)
        _check(lib().okc_descriptors_to_hex(self.h, _p(d), len(d), _p(out)), "okc_descriptors_to_hex")
        return [bytes(r).decode("ascii") for r in out[:len(d)]]

    def descriptors_from_hex(self, strings):
        """Inverse, with the reference reader's tolerance (see include/okvis_cuda.h). Every string must hold >= 96 characters."""
        rows = [s.encode("latin-1") if isinstance(s, str) else bytes(s) for s in strings]
        for r in rows:
            if len(r) < 96:
                raise IndexError("descriptor string shorter than 96 characters (std::out_of_range in the reference)")
        h = np.frombuffer(b"".join(r[:96] for r in rows), np.uint8) if rows else np.zeros(0, np.uint8)
        out = np.zeros((max(len(rows), 1), 48), np.uint8)
        _check(lib().okc_descriptors_from_hex(self.h, _p(np.ascontiguousarray(h)), len(rows), _p(out)),
               "okc_descriptors_from_hex")
        return out[:len(rows)]

    def descriptor_mean(self, descriptors, group_start):
        """FBrisk::meanValue of every group of rows (group g = rows group_start[g]..group_start[g+1]-1)."""
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 48)
        gs = np.ascontiguousarray(group_start, np.int32)
        n = len(gs) - 1
        out = np.zeros((max(n, 1), 48), np.uint8)
        _check(lib().okc_descriptor_mean(self.h, _p(d), _p(gs), n, _p(out)), "okc_descriptor_mean")
        return out[:n]

    # -- brute force (Frontend::verifyRecognisedPlace, Frontend.cpp:316-345; the 10k x 1M sweep) --
    def match_batched(self, queries, mp, threshold=60, index_offset=0):
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 48)
        m = np.ascontiguousarray(mp, np.uint8).reshape(-1, 48)
        best = np.empty(len(q), np.uint64)
        _check(lib().okc_match_batched(self.h, _p(q), len(q), _p(m), len(m), int(threshold), int(index_offset),
                                       _p(best)), "okc_match_batched")
        return best

    def descriptor_set(self, rows):
        return DescriptorSet(self, rows)

    def match_sets(self, qset, mset, threshold=60, index_offset=0, best_device=None, stream=None, want_host=True):
        best = np.empty(qset.n, np.uint64) if want_host else None
        _check(lib().okc_match_sets(self.h, qset.h, mset.h, int(threshold), int(index_offset), _p(best),
                                    C.c_void_p(best_device) if best_device else None,
                                    C.c_void_p(stream) if stream else None), "okc_match_sets")
        return best

    # -- Frontend::matchToMapByThread, Frontend.cpp:1744-1827 --
    def match_sets_ordered(self, qset, mset, threshold, index_offset, best_device_ptr, stream=None):
        """Per-shard best keys as int64 with INT64_MAX for "no match" (ready for an NCCL min all-reduce)."""
        _check(lib().okc_match_sets_ordered(self.h, qset.h, mset.h, int(threshold), int(index_offset),
                                            C.c_void_p(int(best_device_ptr)), C.c_void_p(stream) if stream else None),
               "okc_match_sets_ordered")

    def match_to_map(self, queries, kp_xy, use, mp, map_owner, proj_xy, thresh_sq, hamming_threshold=60.0,
                     num_slices=4):
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 48)
        kp = np.ascontiguousarray(kp_xy, np.float64).reshape(-1, 2)
        us = np.ascontiguousarray(use, np.uint8)
        m = np.ascontiguousarray(mp, np.uint8).reshape(-1, 48)
        ow = np.ascontiguousarray(map_owner, np.int32)
        pr = np.ascontiguousarray(proj_xy, np.float64).reshape(-1, 2)
        nk = len(q)
        bd = np.empty(nk, np.float64)
        bo = np.empty(nk, np.int32)
        ctrs = np.zeros(num_slices, np.uint64)
        errs = np.zeros(num_slices, np.float64)
        _check(lib().okc_match_to_map(self.h, _p(q), _p(kp), _p(us), nk, _p(m), _p(ow), len(m), _p(pr), len(pr),
                                      float(thresh_sq), float(hamming_threshold), int(num_slices), _p(bd), _p(bo),
                                      _p(ctrs), _p(errs)), "okc_match_to_map")
        return bd, bo, ctrs, errs

    # -- Frontend::matchToMapByThreadUnitialised, Frontend.cpp:1830-1956 --
    def match_to_map_uninitialised(self, queries, kp_e_W, use, previous_owner, mp, map_owner, map_e_W, map_r_W, r_WC,
                                   focal, hamming_threshold=60.0, num_slices=4):
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 48)
        ke = np.ascontiguousarray(kp_e_W, np.float64).reshape(-1, 3)
        us = np.ascontiguousarray(use, np.uint8)
        pv = np.ascontiguousarray(previous_owner, np.int32) if previous_owner is not None else None
        m = np.ascontiguousarray(mp, np.uint8).reshape(-1, 48)
        ow = np.ascontiguousarray(map_owner, np.int32)
        me = np.ascontiguousarray(map_e_W, np.float64).reshape(-1, 3)
        mr = np.ascontiguousarray(map_r_W, np.float64).reshape(-1, 3)
        rc = np.ascontiguousarray(r_WC, np.float64)
        nk = len(q)
        bd = np.empty(max(nk, 1), np.float64)
        bo = np.empty(max(nk, 1), np.int32)
        hps = np.zeros((max(nk, 1), 4), np.float64)
        hw = np.zeros(max(nk, 1), np.uint8)
        ctrs = np.zeros(num_slices, np.uint64)
        _check(lib().okc_match_to_map_uninitialised(self.h, _p(q), _p(ke), _p(us), _p(pv), nk, _p(m), _p(ow), _p(me),
                                                    _p(mr), len(m), _p(rc), float(focal), float(hamming_threshold),
                                                    int(num_slices), _p(bd), _p(bo), _p(hps), _p(hw), _p(ctrs)),
               "okc_match_to_map_uninitialised")
        return bd[:nk], bo[:nk], hps[:nk], hw[:nk], ctrs

    # -- worker lambda of Frontend::matchMotionStereo, Frontend.cpp:2053-2137 --
    def match_motion_stereo(self, view0, use0, view1, hamming_threshold=60.0):
        keep = []
        s0, s1 = self._view(view0, keep), self._view(view1, keep)
        u0 = np.ascontiguousarray(use0, np.uint8)
        n0 = s0.n
        dist = np.empty(n0, np.float64)
        k1 = np.empty(n0, np.int32)
        hp = np.empty((n0, 4), np.float64)
        init = np.empty(n0, np.uint8)
        qual = np.empty(n0, np.float64)
        _check(lib().okc_match_motion_stereo(self.h, C.byref(s0), _p(u0), C.byref(s1), float(hamming_threshold),
                                             _p(dist), _p(k1), _p(hp), _p(init), _p(qual)), "okc_match_motion_stereo")
        return dist, k1, hp, init, qual

    @staticmethod
    def _view(v, keep):
        s = StereoView()
        desc = np.ascontiguousarray(v["desc"], np.uint8).reshape(-1, 48)
        size = np.ascontiguousarray(v["size"], np.float64)
        e_C = np.ascontiguousarray(v["e_C"], np.float64).reshape(-1, 3)
        bpv = np.ascontiguousarray(v["bp_valid"], np.uint8)
        keep.extend([desc, size, e_C, bpv])
        s.desc, s.size, s.e_C, s.bp_valid = desc.ctypes.data, size.ctypes.data, e_C.ctypes.data, bpv.ctypes.data
        s.n = len(desc)
        s.C_WC[:] = list(np.asarray(v["C_WC"], np.float64).reshape(9))
        s.r_WC[:] = list(np.asarray(v["r_WC"], np.float64).reshape(3))
        s.f = float(v["f"])
        return s

    # -- Frontend::matchStereo inner loops, Frontend.cpp:2259-2316 --
    def match_stereo(self, view0, view1, hamming_threshold=60.0):
        keep = []

        def mk(v):
            s = StereoView()
            desc = np.ascontiguousarray(v["desc"], np.uint8).reshape(-1, 48)
            size = np.ascontiguousarray(v["size"], np.float64)
            e_C = np.ascontiguousarray(v["e_C"], np.float64).reshape(-1, 3)
            bpv = np.ascontiguousarray(v["bp_valid"], np.uint8)
            keep.extend([desc, size, e_C, bpv])
            s.desc, s.size, s.e_C, s.bp_valid = desc.ctypes.data, size.ctypes.data, e_C.ctypes.data, bpv.ctypes.data
            s.n = len(desc)
            s.C_WC[:] = list(np.asarray(v["C_WC"], np.float64).reshape(9))
            s.r_WC[:] = list(np.asarray(v["r_WC"], np.float64).reshape(3))
            s.f = float(v["f"])
            return s

        s0, s1 = mk(view0), mk(view1)
        n0 = s0.n
        dist = np.empty(n0, np.float64)
        k1 = np.empty(n0, np.int32)
        hp = np.empty((n0, 4), np.float64)
        init = np.empty(n0, np.uint8)
        _check(lib().okc_match_stereo(self.h, C.byref(s0), C.byref(s1), float(hamming_threshold), _p(dist), _p(k1),
                                      _p(hp), _p(init)), "okc_match_stereo")
        return dist, k1, hp, init


def packed_view_bytes(max_keypoints):
    return int(lib().okc_packed_view_bytes(int(max_keypoints)))


def slot_pose(C_WC, r_WC):
    """okc_slot_pose row (24 doubles: T_CW[12], C_WC[9], r_WC[3]) of a camera pose T_WC = (C_WC, r_WC); the inverse
    as okvis::kinematics::Transformation::inverse() forms it: C^T and -(C^T r)."""
    Cm = np.asarray(C_WC, np.float64).reshape(3, 3)
    r = np.asarray(r_WC, np.float64).reshape(3)
    Ct = np.ascontiguousarray(Cm.T)
    t = np.empty(3)
    for a in range(3):   # coefficient-wise (x0 y0 + x1 y1) + x2 y2, then negated
        t[a] = -((Ct[a, 0] * r[0] + Ct[a, 1] * r[1]) + Ct[a, 2] * r[2])
    return np.concatenate([Ct.reshape(9), t, Cm.reshape(9), r])


class LandmarkPool:
    """Device-resident landmarks of the keyframe window (okc_landmark_pool): arguments as Context.prepare_landmarks."""

    def __init__(self, ctx, views, hp_W, quality, obs_offset, obs_view, obs_kp):
        Cw = np.ascontiguousarray(views["C_WC"], np.float64).reshape(-1, 9)
        rw = np.ascontiguousarray(views["r_WC"], np.float64).reshape(-1, 3)
        row0 = np.ascontiguousarray(views["row0"], np.int64)
        desc = np.ascontiguousarray(views["desc"], np.uint8).reshape(-1, 48)
        eC = np.ascontiguousarray(views["e_C"], np.float64).reshape(-1, 3)
        lv = LandmarkViews(len(Cw), Cw.ctypes.data, rw.ctypes.data, row0.ctypes.data, desc.ctypes.data, eC.ctypes.data,
                           len(desc))
        hp = np.ascontiguousarray(hp_W, np.float64).reshape(-1, 4)
        self.L = len(hp)
        q = np.ascontiguousarray(quality, np.float64)
        off = np.ascontiguousarray(obs_offset, np.int64)
        ov = np.ascontiguousarray(obs_view, np.int32)
        ok = np.ascontiguousarray(obs_kp, np.int32)
        self.h = C.c_void_p()
        _check(lib().okc_landmark_pool_create(ctx.h, C.byref(lv), self.L, _p(hp), _p(q), _p(off), _p(ov), _p(ok),
                                              C.byref(self.h)), "okc_landmark_pool_create")

    def update(self, hp_W=None, quality=None, C_WC=None, r_WC=None):
        a = [None if x is None else np.ascontiguousarray(x, np.float64) for x in (hp_W, quality, C_WC, r_WC)]
        _check(lib().okc_landmark_pool_update(self.h, *[_p(x) for x in a]), "okc_landmark_pool_update")

    def close(self):
        if self.h:
            lib().okc_landmark_pool_destroy(self.h)
            self.h = C.c_void_p()


class Keyframe:
    """An older multiframe retained on the device for matchMotionStereo (okc_keyframe)."""

    def __init__(self, ctx, n_views, max_keypoints):
        self.n_views, self.max_kp = n_views, max_keypoints
        self.h = C.c_void_p()
        _check(lib().okc_keyframe_create(ctx.h, n_views, max_keypoints, C.byref(self.h)), "okc_keyframe_create")

    def set_view(self, view, C_WC, r_WC, f, use0=None):
        Cw = np.ascontiguousarray(C_WC, np.float64).reshape(9)
        rw = np.ascontiguousarray(r_WC, np.float64).reshape(3)
        u = None if use0 is None else np.ascontiguousarray(use0, np.uint8)
        _check(lib().okc_keyframe_set_view(self.h, view, _p(Cw), _p(rw), float(f), _p(u), 0 if u is None else len(u)),
               "okc_keyframe_set_view")

    def close(self):
        if self.h:
            lib().okc_keyframe_destroy(self.h)
            self.h = C.c_void_p()


class DescriptorSet:
    def __init__(self, ctx, rows):
        r = np.ascontiguousarray(rows, np.uint8).reshape(-1, 48)
        self.n = len(r)
        self.h = C.c_void_p()
        _check(lib().okc_descriptor_set_create(ctx.h, _p(r), self.n, C.byref(self.h)), "okc_descriptor_set_create")

    def close(self):
        if self.h:
            lib().okc_descriptor_set_destroy(self.h)
            self.h = C.c_void_p()


class MapWindow:
    """Descriptor pool of the landmarks to match (Frontend.cpp:1365-1411), resident on the device."""

    def __init__(self, ctx, rows, owner, num_landmarks):
        r = np.ascontiguousarray(rows, np.uint8).reshape(-1, 48)
        o = np.ascontiguousarray(owner, np.int32)
        assert len(r) == len(o)
        self.nm, self.num_landmarks = len(r), int(num_landmarks)
        self.h = C.c_void_p()
        _check(lib().okc_map_window_create(ctx.h, _p(r), _p(o), self.nm, self.num_landmarks, C.byref(self.h)),
               "okc_map_window_create")

    def close(self):
        if self.h:
            lib().okc_map_window_destroy(self.h)
            self.h = C.c_void_p()


class Camera:
    """One camera of a MultiFrame: owns a stream, device scratch and (optionally) the camera-awareness maps."""

    def __init__(self, ctx, width, height, rays=None, jacobians=None, focal=500.0, max_keypoints=4096):
        self.ctx, self.W, self.H, self.max_kp = ctx, width, height, max_keypoints
        self.h = C.c_void_p()
        if rays is not None:
            rays = np.ascontiguousarray(rays, np.float32)
            jacobians = np.ascontiguousarray(jacobians, np.float32)
            assert rays.shape == (height, width, 3) and jacobians.shape == (height, width, 6)
        _check(lib().okc_camera_create(ctx.h, width, height, _p(rays), _p(jacobians), float(focal), max_keypoints,
                                       C.byref(self.h)), "okc_camera_create")

    def close(self):
        if self.h:
            lib().okc_camera_destroy(self.h)
            self.h = C.c_void_p()

    def set_geometry(self, pinhole):
        """Pinhole.make(...) of this camera: back-projections after describe, focal lengths of the landmark stage."""
        _check(lib().okc_camera_set_geometry(self.h, C.byref(pinhole)), "okc_camera_set_geometry")
        self.pinhole = pinhole

    def _img(self, image):
        img = np.asarray(image)
        # the reference throws std::runtime_error for non-continuous / non-8-bit / multi-channel input
        # (scale-space-feature-detector.h:95-97)
        if img.dtype != np.uint8 or img.ndim != 2 or not img.flags["C_CONTIGUOUS"]:
            raise OkcError("BRISK requires continuous 1-channel 8-bit images")
        if img.shape != (self.H, self.W):
            raise OkcError("image shape %s does not match the camera (%d, %d)" % (img.shape, self.H, self.W))
        return img

    @staticmethod
    def _dp(radius, octaves, abs_threshold, max_kpt):
        return DetectParams(float(radius), int(octaves), float(abs_threshold), int(max_kpt))

    @staticmethod
    def _xp(rotation_invariant, scale_invariant, direction):
        x = DescribeParams(int(rotation_invariant), int(scale_invariant))
        x.extraction_direction[:] = [float(v) for v in direction]
        return x

    def detect(self, image, radius, octaves, abs_threshold, max_kpt):
        img = self._img(image)
        kp = np.zeros(self.max_kp, KP_DTYPE)
        n = C.c_int(0)
        dp = self._dp(radius, octaves, abs_threshold, max_kpt)
        _check(lib().okc_detect(self.h, _p(img), C.byref(dp), _p(kp), self.max_kp, C.byref(n)), "okc_detect")
        return kp[:n.value].copy()

    def describe(self, image, keypoints, rotation_invariant=True, scale_invariant=False, direction=(0.0, 1.0, 0.0)):
        img = self._img(image)
        kp = np.zeros(max(len(keypoints), 1), KP_DTYPE)
        kp[:len(keypoints)] = keypoints
        n = C.c_int(len(keypoints))
        desc = np.zeros((max(len(keypoints), 1), 48), np.uint8)
        xp = self._xp(rotation_invariant, scale_invariant, direction)
        _check(lib().okc_describe(self.h, _p(img), C.byref(xp), _p(kp), C.byref(n), _p(desc)), "okc_describe")
        return kp[:n.value].copy(), desc[:n.value].copy()

    def detect_describe(self, image, radius, octaves, abs_threshold, max_kpt, rotation_invariant=True,
                        scale_invariant=False, direction=(0.0, 1.0, 0.0)):
        self.enqueue(image, radius, octaves, abs_threshold, max_kpt, rotation_invariant, scale_invariant, direction)
        return self.finish()

    def enqueue(self, image, radius, octaves, abs_threshold, max_kpt, rotation_invariant=True, scale_invariant=False,
                direction=(0.0, 1.0, 0.0)):
        img = self._img(image)
        dp = self._dp(radius, octaves, abs_threshold, max_kpt)
        xp = self._xp(rotation_invariant, scale_invariant, direction)
        _check(lib().okc_detect_describe_enqueue(self.h, _p(img), C.byref(dp), C.byref(xp)),
               "okc_detect_describe_enqueue")

    def finish(self):
        kp = np.zeros(self.max_kp, KP_DTYPE)
        desc = np.zeros((self.max_kp, 48), np.uint8)
        n = C.c_int(0)
        _check(lib().okc_detect_describe_finish(self.h, _p(kp), self.max_kp, C.byref(n), _p(desc)),
               "okc_detect_describe_finish")
        return kp[:n.value].copy(), desc[:n.value].copy()

    # -- frame pipeline: detect -> describe -> matchToMap with everything resident in HBM --
    def enqueue_device(self, image_ptr, radius, octaves, abs_threshold, max_kpt, rotation_invariant=True,
                       scale_invariant=False, direction=(0.0, 1.0, 0.0)):
        """image_ptr: device address of a u8[H*W] image (e.g. torch_tensor.data_ptr())."""
        dp = self._dp(radius, octaves, abs_threshold, max_kpt)
        xp = self._xp(rotation_invariant, scale_invariant, direction)
        _check(lib().okc_detect_describe_enqueue_device(self.h, C.c_void_p(image_ptr), C.byref(dp), C.byref(xp)),
               "okc_detect_describe_enqueue_device")

    def match_to_map_enqueue(self, window, proj_xy, thresh_sq, hamming_threshold=60.0, proj_device_ptr=None,
                             num_slices=4):
        self._slices = int(num_slices)
        if proj_device_ptr is not None:
            _check(lib().okc_camera_match_to_map_enqueue(self.h, window.h, C.c_void_p(proj_device_ptr), 1,
                                                         float(thresh_sq), float(hamming_threshold), self._slices),
                   "okc_camera_match_to_map_enqueue")
            return
        pr = np.ascontiguousarray(proj_xy, np.float64).reshape(-1, 2)
        assert len(pr) == window.num_landmarks
        _check(lib().okc_camera_match_to_map_enqueue(self.h, window.h, _p(pr), 0, float(thresh_sq),
                                                     float(hamming_threshold), self._slices),
               "okc_camera_match_to_map_enqueue")

    def frame_finish(self, window):
        kp = np.zeros(self.max_kp, KP_DTYPE)
        desc = np.zeros((self.max_kp, 48), np.uint8)
        n = C.c_int(0)
        bd = np.empty(self.max_kp, np.float64)
        bo = np.empty(self.max_kp, np.int32)
        ctrs = np.zeros(self._slices, np.uint64)
        errs = np.zeros(self._slices, np.float64)
        _check(lib().okc_camera_frame_finish(self.h, window.h, _p(kp), self.max_kp, C.byref(n), _p(desc), _p(bd),
                                             _p(bo), _p(ctrs), _p(errs)), "okc_camera_frame_finish")
        k = n.value
        return kp[:k].copy(), desc[:k].copy(), bd[:k].copy(), bo[:k].copy(), ctrs, errs

    def stream(self):
        return lib().okc_camera_stream(self.h)

    def kernel_timing(self, enable):
        _check(lib().okc_debug_kernel_timing(self.h, int(bool(enable))), "okc_debug_kernel_timing")

    def kernel_times(self):
        names = C.create_string_buffer(512)
        ms = np.zeros(16, np.float32)
        n = C.c_int(0)
        _check(lib().okc_debug_kernel_times(self.h, names, 512, _p(ms), 16, C.byref(n)), "okc_debug_kernel_times")
        return list(zip(names.value.decode().split("\n"), [float(v) for v in ms[:n.value]])) if n.value else []

    def debug_std_sort(self, scores):
        """Permutation the device's std::sort reproduction gives to `scores` (descending, libstdc++ tie order)."""
        sc = np.ascontiguousarray(scores, np.int32)
        order = np.zeros(max(len(sc), 1), np.uint32)
        _check(lib().okc_debug_std_sort(self.h, _p(sc), len(sc), _p(order)), "okc_debug_std_sort")
        return order[:len(sc)].copy()

    def debug_scores(self, image):
        img = self._img(image)
        out = np.zeros((self.H, self.W), np.int32)
        _check(lib().okc_debug_scores(self.h, _p(img), _p(out)), "okc_debug_scores")
        return out

    def debug_candidates(self, image, abs_threshold, cap=1 << 17):
        img = self._img(image)
        out = np.zeros(cap, PWS_DTYPE)
        n = C.c_int(0)
        _check(lib().okc_debug_candidates(self.h, _p(img), int(abs_threshold), _p(out), cap, C.byref(n)),
               "okc_debug_candidates")
        return out[:n.value].copy()


class Batch:
    """n_slots images per launch chain (cameras of a multiframe and/or frames of a sequence); see okvis_cuda.h."""

    def __init__(self, ctx, cams):
        self.ctx, self.cams, self.n = ctx, list(cams), len(cams)
        self.max_kp = cams[0].max_kp
        arr = (C.c_void_p * self.n)(*[c.h for c in cams])
        self.h = C.c_void_p()
        _check(lib().okc_batch_create(ctx.h, arr, self.n, C.byref(self.h)), "okc_batch_create")
        self._xp = (DescribeParams * self.n)()
        self._img = (C.c_void_p * self.n)()
        self._win = (C.c_void_p * self.n)()
        self._proj = (C.c_void_p * self.n)()
        self._gate = (C.c_double * self.n)()
        self._keep = None

    def close(self):
        if self.h:
            lib().okc_batch_destroy(self.h)
            self.h = C.c_void_p()

    def enqueue(self, images, radius, octaves, abs_threshold, max_kpt, directions, rotation_invariant=True,
                images_on_device=False, windows=None, proj=None, gates=None, hamming_threshold=60.0,
                proj_on_device=False, num_slices=4):
        """images: n arrays (host) or n device addresses; directions: n 3-vectors; windows/proj/gates: per slot."""
        keep = []
        for k in range(self.n):
            if images_on_device:
                self._img[k] = int(images[k])
            else:
                a = images[k]
                if isinstance(a, np.ndarray):
                    a = self.cams[k]._img(a)
                    keep.append(a)
                    self._img[k] = a.ctypes.data
                else:
                    self._img[k] = int(a)  # raw host address (e.g. a pinned torch tensor's data_ptr())
            self._xp[k].rotation_invariant = int(rotation_invariant)
            self._xp[k].scale_invariant = 0
            self._xp[k].extraction_direction[:] = [float(v) for v in directions[k]]
        dp = DetectParams(float(radius), int(octaves), float(abs_threshold), int(max_kpt))
        if windows is None:
            _check(lib().okc_batch_enqueue(self.h, self._img, int(images_on_device), C.byref(dp), self._xp, None, None,
                                           0, None, float(hamming_threshold), 1), "okc_batch_enqueue")
        else:
            for k in range(self.n):
                self._win[k] = windows[k].h if windows[k] is not None else None
                if proj_on_device or not isinstance(proj[k], np.ndarray):
                    self._proj[k] = int(proj[k])
                else:
                    p = np.ascontiguousarray(proj[k], np.float64)
                    keep.append(p)
                    self._proj[k] = p.ctypes.data
                self._gate[k] = float(gates[k])
            _check(lib().okc_batch_enqueue(self.h, self._img, int(images_on_device), C.byref(dp), self._xp, self._win,
                                           self._proj, int(proj_on_device), self._gate, float(hamming_threshold),
                                           int(num_slices)), "okc_batch_enqueue")
        self._slices = int(num_slices)
        self._matching = windows is not None
        self._keep = keep  # host buffers stay alive until finish()

    # -- low-overhead form for pipelines: per-slot settings once, then only address arrays per enqueue --
    def configure(self, radius, octaves, abs_threshold, max_kpt, directions, rotation_invariant=True, windows=None,
                  gates=None, hamming_threshold=60.0, num_slices=4):
        self._dp = DetectParams(float(radius), int(octaves), float(abs_threshold), int(max_kpt))
        for k in range(self.n):
            self._xp[k].rotation_invariant = int(rotation_invariant)
            self._xp[k].scale_invariant = 0
            self._xp[k].extraction_direction[:] = [float(v) for v in directions[k]]
            if windows is not None:
                self._win[k] = windows[k].h if windows[k] is not None else None
                self._gate[k] = float(gates[k])
        self._matching = windows is not None
        self._slices = int(num_slices) if windows is not None else 1
        self._thr = float(hamming_threshold)

    def enqueue_addrs(self, image_addrs, images_on_device, proj_addrs=None, proj_on_device=False):
        """image_addrs / proj_addrs: contiguous np.uint64 arrays of n host or device addresses (u8[H*W] images,
        double[2*L] projections). The arrays themselves are only read during the call."""
        m = self._matching
        _check(lib().okc_batch_enqueue(self.h, C.c_void_p(image_addrs.ctypes.data), int(images_on_device),
                                       C.byref(self._dp), self._xp, self._win if m else None,
                                       C.c_void_p(proj_addrs.ctypes.data) if m else None, int(proj_on_device),
                                       self._gate if m else None, self._thr, self._slices), "okc_batch_enqueue")

    def finish(self):
        _check(lib().okc_batch_finish(self.h), "okc_batch_finish")
        self._keep = None

    def slot_frame(self, slot):
        kp = np.empty(self.max_kp, KP_DTYPE)
        desc = np.empty((self.max_kp, 48), np.uint8)
        n = C.c_int(0)
        _check(lib().okc_batch_slot_frame(self.h, slot, _p(kp), self.max_kp, C.byref(n), _p(desc)),
               "okc_batch_slot_frame")
        return kp[:n.value], desc[:n.value]

    def set_graph(self, enable):
        """CUDA-graph replay of the launch chains (default on); the results do not depend on it."""
        _check(lib().okc_batch_set_graph(self.h, int(bool(enable))), "okc_batch_set_graph")

    def slot_matches(self, slot, nk):
        bd = np.empty(self.max_kp, np.float64)
        bo = np.empty(self.max_kp, np.int32)
        ctrs = np.zeros(self._slices, np.uint64)
        errs = np.zeros(self._slices, np.float64)
        _check(lib().okc_batch_slot_matches(self.h, slot, _p(bd), _p(bo), _p(ctrs), _p(errs)),
               "okc_batch_slot_matches")
        return bd[:nk], bo[:nk], ctrs, errs

    def read_all(self, out=None):
        """All slots at once into (reusable) arrays: dict(kp [n,max_kp], n [n], desc [n,max_kp,48], dist, owner,
        ctrs [n,slices], errs [n,slices])."""
        if out is None:
            out = dict(kp=np.empty((self.n, self.max_kp), KP_DTYPE), n=np.zeros(self.n, np.int32),
                       desc=np.empty((self.n, self.max_kp, 48), np.uint8),
                       dist=np.empty((self.n, self.max_kp), np.float64), owner=np.empty((self.n, self.max_kp), np.int32),
                       ctrs=np.zeros((self.n, 64), np.uint64), errs=np.zeros((self.n, 64), np.float64))
        m = self._matching or getattr(self, "_pool_matching", False)
        ctrs = out["ctrs"].reshape(-1)
        errs = out["errs"].reshape(-1)
        _check(lib().okc_batch_read_all(self.h, _p(out["kp"]), _p(out["n"]), _p(out["desc"]),
                                        _p(out["dist"]) if m else None, _p(out["owner"]) if m else None,
                                        _p(ctrs) if m else None, _p(errs) if m else None), "okc_batch_read_all")
        return out

    # -- frame pipeline beyond matchToMapByThread --
    def set_poses(self, poses):
        """poses: [n, 24] rows of slot_pose()."""
        p = np.ascontiguousarray(poses, np.float64).reshape(self.n, 24)
        _check(lib().okc_batch_set_poses(self.h, _p(p)), "okc_batch_set_poses")

    def slot_back_projections(self, slot, nk):
        e = np.zeros((self.max_kp, 3), np.float64)
        v = np.zeros(self.max_kp, np.uint8)
        nv = C.c_int(0)
        _check(lib().okc_batch_slot_back_projections(self.h, slot, _p(e), _p(v), C.byref(nv)),
               "okc_batch_slot_back_projections")
        return e[:nk], v[:nk], nv.value

    def match_landmarks_enqueue(self, pool, hamming_threshold=60.0, num_slices=4, flags=0):
        _check(lib().okc_batch_match_landmarks_enqueue(self.h, pool.h, float(hamming_threshold), int(num_slices),
                                                       int(flags)), "okc_batch_match_landmarks_enqueue")
        self._slices = int(num_slices)
        self._pool_matching = True   # read_all returns the matchToMapByThread results of the landmark stage
        self._pool_L = pool.L

    def slot_uninitialised(self, slot, nk):
        bd = np.empty(self.max_kp, np.float64)
        bo = np.empty(self.max_kp, np.int32)
        hp = np.empty((self.max_kp, 4), np.float64)
        hw = np.empty(self.max_kp, np.uint8)
        ctrs = np.zeros(self._slices, np.uint64)
        _check(lib().okc_batch_slot_uninitialised(self.h, slot, _p(bd), _p(bo), _p(hp), _p(hw), _p(ctrs)),
               "okc_batch_slot_uninitialised")
        return bd[:nk], bo[:nk], hp[:nk], hw[:nk], ctrs

    def slot_landmarks(self, slot):
        L = self._pool_L
        out = dict(proj=np.zeros((max(L, 1), 2)), visible=np.zeros(max(L, 1), np.uint8), is3d=np.zeros(max(L, 1), np.uint8),
                   rows=np.zeros(max(L, 1), np.int32), owner=np.zeros(max(3 * L, 1), np.int32),
                   pool=np.zeros((max(3 * L, 1), 48), np.uint8))
        nm = C.c_int(0)
        _check(lib().okc_batch_slot_landmarks(self.h, slot, _p(out["proj"]), _p(out["visible"]), _p(out["is3d"]),
                                              _p(out["rows"]), C.byref(nm), _p(out["owner"]), _p(out["pool"])),
               "okc_batch_slot_landmarks")
        n = nm.value
        return dict(proj=out["proj"][:L], visible=out["visible"][:L], is3d=out["is3d"][:L], rows=out["rows"][:L], nm=n,
                    owner=out["owner"][:n], pool=out["pool"][:n])

    def match_stereo_enqueue(self, pairs, hamming_threshold=60.0):
        """pairs: [(slot0, slot1)]; the marshalled arrays are cached per list object (pipelines pass the same list)."""
        c = getattr(self, "_pairs_cache", None)
        if c is None or c[0] is not pairs:
            c = (pairs, np.ascontiguousarray([p[0] for p in pairs], np.int32),
                 np.ascontiguousarray([p[1] for p in pairs], np.int32))
            self._pairs_cache = c
        s0, s1 = c[1], c[2]
        _check(lib().okc_batch_match_stereo_enqueue(self.h, len(pairs), _p(s0), _p(s1), float(hamming_threshold)),
               "okc_batch_match_stereo_enqueue")

    def pair_stereo(self, pair, n0):
        dist = np.empty(self.max_kp, np.float64)
        k1 = np.empty(self.max_kp, np.int32)
        hp = np.empty((self.max_kp, 4), np.float64)
        init = np.empty(self.max_kp, np.uint8)
        _check(lib().okc_batch_pair_stereo(self.h, pair, _p(dist), _p(k1), _p(hp), _p(init)), "okc_batch_pair_stereo")
        return dist[:n0], k1[:n0], hp[:n0], init[:n0]

    def retain(self, slots, keyframe):
        s = np.ascontiguousarray(slots, np.int32)
        _check(lib().okc_batch_retain(self.h, len(s), _p(s), keyframe.h), "okc_batch_retain")

    def match_motion_enqueue(self, jobs, hamming_threshold=60.0, only_unmatched=True):
        """jobs: [(keyframe, view, slot)]; the marshalled arrays are cached per list object."""
        c = getattr(self, "_jobs_cache", None)
        if c is None or c[0] is not jobs:
            c = (jobs, (C.c_void_p * len(jobs))(*[j[0].h for j in jobs]), np.ascontiguousarray([j[1] for j in jobs], np.int32),
                 np.ascontiguousarray([j[2] for j in jobs], np.int32), max([self.max_kp] + [j[0].max_kp for j in jobs]))
            self._jobs_cache = c
        kf, v, s = c[1], c[2], c[3]
        self._motion_cap = c[4]
        _check(lib().okc_batch_match_motion_enqueue(self.h, len(jobs), kf, _p(v), _p(s), float(hamming_threshold),
                                                    int(bool(only_unmatched))), "okc_batch_match_motion_enqueue")

    def job_motion(self, job):
        cap = self._motion_cap
        dist = np.empty(cap, np.float64)
        k1 = np.empty(cap, np.int32)
        hp = np.empty((cap, 4), np.float64)
        init = np.empty(cap, np.uint8)
        q = np.empty(cap, np.float64)
        n0 = C.c_int(0)
        _check(lib().okc_batch_job_motion(self.h, job, C.byref(n0), _p(dist), _p(k1), _p(hp), _p(init), _p(q)),
               "okc_batch_job_motion")
        n = n0.value
        return dist[:n], k1[:n], hp[:n], init[:n], q[:n]

    def pack_views(self, slots, packed_device_ptr):
        """Views of `slots` into consecutive records of packed_view_bytes(max_kp) bytes of caller device memory."""
        s = np.ascontiguousarray(slots, np.int32)
        _check(lib().okc_batch_pack_views(self.h, len(s), _p(s), C.c_void_p(int(packed_device_ptr))),
               "okc_batch_pack_views")

    def match_stereo_packed_enqueue(self, slot0, packed_device_ptr, packed_index, poses1, f1, hamming_threshold=60.0):
        s0 = np.ascontiguousarray(slot0, np.int32)
        pi = np.ascontiguousarray(packed_index, np.int32)
        ps = np.ascontiguousarray(poses1, np.float64).reshape(len(s0), 24)
        ff = np.ascontiguousarray(f1, np.float64)
        _check(lib().okc_batch_match_stereo_packed_enqueue(self.h, len(s0), _p(s0), C.c_void_p(int(packed_device_ptr)),
                                                           _p(pi), _p(ps), _p(ff), float(hamming_threshold)),
               "okc_batch_match_stereo_packed_enqueue")

    def read_stages(self, out=None, n_pairs=0, n_jobs=0, uninitialised=False):
        """All stage results of a finished frame into (reusable) arrays: dict(u_dist, u_owner, u_hps, u_written, u_ctrs,
        s_dist, s_k1, s_hp, s_init, m_n0, m_dist, m_k1, m_hp, m_init, m_quality)."""
        if out is None:
            mk, rows = self.max_kp, getattr(self, "_motion_cap", self.max_kp)
            out = dict(u_dist=np.empty((self.n, mk)), u_owner=np.empty((self.n, mk), np.int32), u_hps=np.empty((self.n, mk, 4)),
                       u_written=np.empty((self.n, mk), np.uint8), u_ctrs=np.zeros((self.n, 64), np.uint64),
                       s_dist=np.empty((max(n_pairs, 1), mk)), s_k1=np.empty((max(n_pairs, 1), mk), np.int32),
                       s_hp=np.empty((max(n_pairs, 1), mk, 4)), s_init=np.empty((max(n_pairs, 1), mk), np.uint8),
                       m_n0=np.zeros(max(n_jobs, 1), np.int32), m_dist=np.empty((max(n_jobs, 1), rows)),
                       m_k1=np.empty((max(n_jobs, 1), rows), np.int32), m_hp=np.empty((max(n_jobs, 1), rows, 4)),
                       m_init=np.empty((max(n_jobs, 1), rows), np.uint8), m_quality=np.empty((max(n_jobs, 1), rows)))
        u = [out[k] if uninitialised else None for k in ("u_dist", "u_owner", "u_hps", "u_written")]
        uc = out["u_ctrs"].reshape(-1) if uninitialised else None
        s_ = [out[k] if n_pairs else None for k in ("s_dist", "s_k1", "s_hp", "s_init")]
        m_ = [out[k] if n_jobs else None for k in ("m_n0", "m_dist", "m_k1", "m_hp", "m_init", "m_quality")]
        _check(lib().okc_batch_read_stages(self.h, *[_p(x) for x in u], _p(uc), *[_p(x) for x in s_], *[_p(x) for x in m_]),
               "okc_batch_read_stages")
        return out

    def transfer_bytes(self):
        """(host->device, device->host) bytes the last frame's copies moved."""
        a, b = C.c_uint64(0), C.c_uint64(0)
        _check(lib().okc_batch_transfer_bytes(self.h, C.byref(a), C.byref(b)), "okc_batch_transfer_bytes")
        return int(a.value), int(b.value)

    def view(self):
        """okc_batch_view_results: the results of a finished frame as numpy arrays that ALIAS the batch's pinned host
        memory (no copy; valid until the next enqueue on this batch). dict(n [slots], kp [slots, max_kp], desc
        [slots, max_kp, 48], dist / owner [slots, max_kp], stats (count, sum) [slots, slices], u_dist / u_owner /
        u_hps / u_written [slots, max_kp(,4)], u_ctrs [slots, slices], s_* [pairs, rows(,4)], m_* [jobs, rows(,4)],
        m_n0 [jobs]); groups of stages that did not run are absent."""
        v = BatchView()
        _check(lib().okc_batch_view_results(self.h, C.byref(v)), "okc_batch_view_results")
        key = bytes(v)
        cached = getattr(self, "_view_cache", None)
        if cached is not None and cached[0] == key:
            return cached[1]
        B, mk, S = v.n_slots, v.max_keypoints, max(v.num_slices, 1)

        def arr(base, nbytes, off, shape, strides, dtype):
            raw = np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(base))
            return np.ndarray(shape, dtype, raw, off, strides)

        out = {"n": np.ctypeslib.as_array((C.c_int32 * (8 * B)).from_address(v.counters)).reshape(B, 8)[:, 2]}
        tot = v.slot_stride * B
        out["kp"] = arr(v.slots, tot, v.off_keypoints, (B, mk), (v.slot_stride, 28), KP_DTYPE)
        out["desc"] = arr(v.slots, tot, v.off_descriptors, (B, mk, 48), (v.slot_stride, 48, 1), np.uint8)
        if v.map_match:
            out["dist"] = arr(v.slots, tot, v.off_best_dist, (B, mk), (v.slot_stride, 8), np.float64)
            out["owner"] = arr(v.slots, tot, v.off_best_owner, (B, mk), (v.slot_stride, 4), np.int32)
            out["stats"] = arr(v.slots, tot, v.off_stats, (B, S), (v.slot_stride, 16),
                               np.dtype([("count", "<u8"), ("sum", "<f8")]))
        if v.uninit:
            ut = v.uninit_stride * B
            out["u_dist"] = arr(v.uninit, ut, v.u_off_dist, (B, mk), (v.uninit_stride, 8), np.float64)
            out["u_owner"] = arr(v.uninit, ut, v.u_off_owner, (B, mk), (v.uninit_stride, 4), np.int32)
            out["u_hps"] = arr(v.uninit, ut, v.u_off_hps, (B, mk, 4), (v.uninit_stride, 32, 8), np.float64)
            out["u_written"] = arr(v.uninit, ut, v.u_off_written, (B, mk), (v.uninit_stride, 1), np.uint8)
            out["u_ctrs"] = np.ctypeslib.as_array((C.c_uint64 * (v.uninit_ctrs_stride * B)).from_address(v.uninit_ctrs)) \
                .reshape(B, v.uninit_ctrs_stride)[:, :S]
        for tag, jv in (("s", v.stereo), ("m", v.motion)):
            if not jv.base:
                continue
            jt, J, R = jv.stride * jv.n, jv.n, jv.rows
            out[tag + "_dist"] = arr(jv.base, jt, jv.off_dist, (J, R), (jv.stride, 8), np.float64)
            out[tag + "_k1"] = arr(jv.base, jt, jv.off_k1, (J, R), (jv.stride, 4), np.int32)
            out[tag + "_hp"] = arr(jv.base, jt, jv.off_hp, (J, R, 4), (jv.stride, 32, 8), np.float64)
            out[tag + "_init"] = arr(jv.base, jt, jv.off_initialisable, (J, R), (jv.stride, 1), np.uint8)
            if tag == "m":
                out["m_quality"] = arr(jv.base, jt, jv.off_quality, (J, R), (jv.stride, 8), np.float64)
                out["m_n0"] = np.ctypeslib.as_array((C.c_int32 * J).from_address(v.motion_n0))
        self._view_cache = (key, out)
        return out

    def stream(self):
        return lib().okc_batch_stream(self.h)

    def kernel_timing(self, enable):
        _check(lib().okc_debug_batch_kernel_timing(self.h, int(bool(enable))), "okc_debug_batch_kernel_timing")

    def kernel_times(self):
        names = C.create_string_buffer(1024)
        ms = np.zeros(32, np.float32)
        n = C.c_int(0)
        _check(lib().okc_debug_batch_kernel_times(self.h, names, 1024, _p(ms), 32, C.byref(n)),
               "okc_debug_batch_kernel_times")
        return list(zip(names.value.decode().split("\n"), [float(v) for v in ms[:n.value]])) if n.value else []
