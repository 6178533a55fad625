#!/usr/bin/env python3
"""bench.py - BASELINE.json's metric on BASELINE.json's configs, one JSON line on stdout.

Headline workload (configs[1]): EuRoC-config frontend on a synthetic 752x480 stereo sequence. One "step" = one pass of
the hot path over one batch of --frames-per-batch stereo frames; per frame and camera: BRISK detect (Harris + NMS +
uniformity + sub-pixel) -> camera-aware 48-byte describe -> back-projections -> Frontend::matchToMap against a
2 000-landmark keyframe window kept in HBM (landmark preparation for the frame's pose, matchToMapByThread on the 3D
landmarks, matchToMapByThreadUnitialised on the others). The metric is frames/s (= steps x frames per step / time).

  value  frames/s with every input already in HBM (image pool larger than L2, landmark pool), results left on the
         device; CUDA events around the K steps; max over ranks.
  e2e    frames/s through the drop-in C ABI with HOST buffers: H2D of the images and poses, D2H of keypoints,
         descriptors and match lists inside the timed region, every step.
  extra.configs   the same three numbers (+ CPU baseline + in-run parity count) for every BASELINE.json config:
         [0] stereo pair + matchStereo, [1] the headline, [2] Hilti22 5 x 720x540 multiframe (cameras sharded over ranks
         under --gpus N, keypoints + descriptors all-gathered over NCCL for the cross-camera matchStereo),
         [3] D455 848x480 at a 2 500-keypoint budget with matchMotionStereo against 5 keyframes, [4] the 10k x 1M sweep.
  --impl reference   the reference's own CPU code (oracle/_ref = external/brisk compiled unchanged, + the restated
         Frontend.cpp / okvis_cv loops of oracle/port) on all host cores, headline workload.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

E2E_REPEATS = 3
DIRECTION = (0.05, 0.99, 0.1)   # gravity in the camera frame (Frontend.cpp:234-238); any fixed unit-ish vector
Z0, DISPARITY, STEP = 3.0, 24, (2, 1)   # the synthetic world: see okvis2-x_b200/synthetic.py (rig_sequence)
HAMMING = 60.0

# name -> (rig, frames in the image pool (> L2), keyframes, stages, dense scene)
CONFIGS = {
    "stereo_pair": dict(index=0, rig="euroc", frames=200, keyframes=0, landmarks=False, stereo=True, motion=False,
                        what="configs[0]: 752x480 stereo pair, 2x(detect+describe+back-project) + matchStereo"),
    "stereo_pair_noise": dict(index=0, rig="euroc", frames=64, keyframes=0, landmarks=False, stereo=True, motion=False,
                              noise=True,
                              what="configs[0], variant S-noise: i.i.d. uniform u8 images (the reference's own test input, "
                                   "okvis_cv/test/TestFrame.cpp:62-65) - the worst case for the number of NMS candidates; "
                                   "2x(detect+describe+back-project) + matchStereo"),
    "euroc_sequence": dict(index=1, rig="euroc", frames=200, keyframes=5, landmarks=True, stereo=False, motion=False,
                           what="configs[1]: EuRoC 752x480 stereo sequence, per frame 2x(detect+describe+back-project) + "
                                "matchToMap (landmark preparation, matchToMapByThread, matchToMapByThreadUnitialised) vs a "
                                "2000-landmark keyframe window"),
    "hilti_multiframe": dict(index=2, rig="hilti", frames=80, keyframes=5, landmarks=True, stereo=True, motion=False,
                             what="configs[2]: Hilti22 5 x 720x540 multiframe, per camera detect+describe+back-project + "
                                  "matchToMap, matchStereo over the 4 adjacent camera pairs"),
    "d455_sequence": dict(index=3, rig="d455", frames=180, keyframes=5, landmarks=True, stereo=False, motion=True,
                          what="configs[3]: D455 848x480 stereo at max_num_keypoints 2500, per frame 2x(detect+describe+"
                               "back-project) + matchToMap + matchMotionStereo vs 5 keyframes"),
}
HEADLINE = "euroc_sequence"
NUM_LANDMARKS = 2000


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300, help="batches of --frames-per-batch frames (headline config)")
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames-per-batch", type=int, default=32, help="frames (multiframes) per launch chain")
    ap.add_argument("--in-flight", type=int, default=4, help="batches in flight (one stream each)")
    ap.add_argument("--no-sweep", action="store_true", help="skip the 10k x 1M matcher sweep in `extra`")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--only", default="", help="comma-separated config names (default: all; the headline always runs)")
    ap.add_argument("--sparse", action="store_true", help="round-1 scene (400 large rectangles) instead of the dense one")
    ap.add_argument("--no-graph", action="store_true", help="plain kernel launches instead of CUDA-graph replay of the chains")
    return ap.parse_args(argv)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, smax, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                smax.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(smax)), samples=len(sm))
        out["reasons"] = sorted(reasons)
        return out


# ------------------------------------------------------------------------------------------------------------
# Workload: everything a step reads, as host arrays (pure functions of the seeds + the keyframe views, which are
# produced by whichever implementation builds the world)
# ------------------------------------------------------------------------------------------------------------
def build_images(syn, name, sparse=False):
    conf = CONFIGS[name]
    rig = syn.RIGS[conf["rig"]]
    W, H, n_cams = rig["width"], rig["height"], rig["n_cams"]
    if conf.get("noise"):
        imgs = np.random.default_rng(77).integers(0, 256, (n_cams, conf["frames"], H, W), dtype=np.uint8)
    else:
        imgs = syn.rig_sequence(conf["frames"], n_cams, W, H, seed=2 + conf["index"], step=STEP, disparity=DISPARITY,
                                dense=not sparse, rects=400)
    rays, jac = syn.rig_maps(rig)
    r_all = syn.rig_poses(conf["frames"], n_cams, rig["f"][0], rig["f"][1], STEP, DISPARITY, Z0)
    kf_frames = [int(round((i + 0.5) * conf["frames"] / conf["keyframes"])) for i in range(conf["keyframes"])]
    return dict(name=name, conf=conf, rig=rig, cfg=rig["cfg"], W=W, H=H, n_cams=n_cams, images=imgs, rays=rays, jac=jac,
                r_all=r_all, kf_frames=kf_frames, f_mean=0.5 * (rig["f"][0] + rig["f"][1]))


def build_world(syn, wl, keyframe_views):
    """keyframe_views[(frame, cam)] = dict(kp, desc, e_C, valid): the landmark pool of the keyframe window and the
    eligibility flags of the keyframes' keypoints (estimator state in the reference: a seeded 80 %)."""
    views = []
    for f in wl["kf_frames"]:
        for c in range(wl["n_cams"]):
            v = dict(keyframe_views[(f, c)])
            v["r_WC"] = wl["r_all"][f, c]
            views.append(v)
    wl["views"] = views
    if wl["conf"]["landmarks"]:
        wl["pool_args"] = syn.plane_landmarks(views, Z0, wl["rig"]["f"][0], NUM_LANDMARKS, seed=5)
    rng = np.random.default_rng(11)
    wl["use0"] = [(rng.random(len(v["kp"])) < 0.8).astype(np.uint8) for v in views]
    return wl


# ------------------------------------------------------------------------------------------------------------
# CPU arm: the reference's own code (oracle/_ref) + the restated okvis loops (oracle/port). Checker / baseline only.
# ------------------------------------------------------------------------------------------------------------
class CpuFrontend:
    def __init__(self, mod, syn, wl):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import pipeline_port
        import refbind
        self.rb, self.mod, self.syn, self.wl = refbind, mod, syn, wl
        self.ref = refbind.Ref()
        self.port = refbind.Port()
        self.pipe = pipeline_port.CpuPipeline(self.ref, self.port)
        rig = wl["rig"]
        self.model = mod.Pinhole.make(wl["W"], wl["H"], rig["f"], rig["c"], rig["distortion"], rig["k"])
        self.kind = "reference"  # detect/describe/Hamming = compiled reference; okvis loops = restated port
        c = wl["cfg"]
        self.det = (c["radius"], 0, c["abs_threshold"], c["max_kpt"])

    def extractors(self):
        """One camera-aware extractor per camera (not thread-safe: one set per worker)."""
        out = []
        for _ in range(self.wl["n_cams"]):
            e = self.rb.RefExtractor(self.ref, True, False)
            e.set_camera(self.wl["rays"], self.wl["jac"], np.float32(self.wl["rig"]["f"][0]))
            e.set_direction(np.asarray(DIRECTION, np.float32))
            out.append(e)
        return out

    def view_of(self, f, c, ext):
        kp, desc = self.pipe.detect_describe(ext, self.wl["images"][c, f], self.det)
        e_C, valid, _ = self.port.back_project(self.model, kp)
        return dict(kp=kp, desc=desc, e_C=e_C, valid=valid)

    def keyframe_views(self, exts):
        return {(f, c): self.view_of(f, c, exts[c]) for f in self.wl["kf_frames"] for c in range(self.wl["n_cams"])}

    def stereo_view(self, v, r_WC):
        return dict(desc=v["desc"], size=v["kp"]["size"].astype(np.float64), e_C=v["e_C"], bp_valid=v["valid"],
                    C_WC=np.eye(3).reshape(9), r_WC=np.asarray(r_WC, np.float64), f=self.wl["f_mean"])

    def stages(self, f, cur, threaded):
        """Everything after describe for the multiframe `f` whose per-camera views are `cur`."""
        wl, conf = self.wl, self.wl["conf"]
        out = {"lm": [], "stereo": [], "motion": []}
        n_cams = wl["n_cams"]
        if conf["landmarks"]:
            for c in range(n_cams):
                r = wl["r_all"][f, c]
                pose = (np.eye(3), r, self.mod.slot_pose(np.eye(3), r)[:12])
                v = cur[c]
                out["lm"].append(self.pipe.landmark_stage(self.model, pose, v["kp"], v["desc"], v["e_C"], v["valid"],
                                                          wl["pool_args"], HAMMING, wl["cfg"]["slices"], threaded=threaded))
        if conf["motion"]:
            nk = len(wl["kf_frames"])
            for k in range(nk):
                for c in range(n_cams):
                    kv = wl["views"][k * n_cams + c]
                    v0 = self.stereo_view(kv, kv["r_WC"])
                    v1 = self.stereo_view(cur[c], wl["r_all"][f, c])
                    lm = out["lm"][c]
                    unmatched = (lm["a14"][1] < 0) & (lm["a15"][1] < 0)
                    out["motion"].append(self.pipe.motion(v0, wl["use0"][k * n_cams + c], v1, unmatched, HAMMING))
        if conf["stereo"]:
            for c in range(n_cams - 1):
                out["stereo"].append(self.port.match_stereo(self.stereo_view(cur[c], wl["r_all"][f, c]),
                                                            self.stereo_view(cur[c + 1], wl["r_all"][f, c + 1]), HAMMING))
        return out

    def frame_like_reference(self, exts, f):
        """Threading as the reference does it: one thread per camera for detect+describe (ThreadedSlam.cpp:676-693),
        num_matching_threads slices per camera for the two map matchers (Frontend.cpp:1519-1536, 1631-1645), the rest
        on the calling thread."""
        n_cams = self.wl["n_cams"]
        cur = [None] * n_cams

        def dd(c):
            cur[c] = self.view_of(f, c, exts[c])
        th = [threading.Thread(target=dd, args=(c,)) for c in range(n_cams)]
        for x in th:
            x.start()
        for x in th:
            x.join()
        return cur, self.stages(f, cur, True)

    def frame_sequential(self, exts, f):
        cur = [self.view_of(f, c, exts[c]) for c in range(self.wl["n_cams"])]
        return cur, self.stages(f, cur, False)

    def throughput_all_cores(self, frames, cores):
        """One whole frame per worker at a time, `cores` workers: the most the host can do with this code.
        Extractors (a 52 MB pattern table each) are built once per worker, outside the timed region."""
        if not hasattr(self, "_pool") or len(self._pool) < cores:
            self._pool = [self.extractors() for _ in range(cores)]
        frames = list(frames)
        lock = threading.Lock()
        it = iter(frames)

        def worker(exts):
            while True:
                with lock:
                    f = next(it, None)
                if f is None:
                    break
                self.frame_sequential(exts, f)
        th = [threading.Thread(target=worker, args=(self._pool[i],)) for i in range(cores)]
        t0 = time.perf_counter()
        for x in th:
            x.start()
        for x in th:
            x.join()
        return len(frames) / (time.perf_counter() - t0)


# ------------------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------------------
def stage_bytes(kernel, W, H, K, M, L):
    """ALGORITHMIC bytes of the stage a kernel belongs to, per image (DESIGN.md 'Roofline')."""
    if kernel in ("harris_nms", "order_candidates", "select_keypoints"):
        return W * H + K * 28                       # detect: image in, keypoints out
    if kernel == "border_filter+describe":
        return W * H + K * (28 + 4 * 12 + 24) + K * (48 + 28)   # image + kp + 4 rays + 1 Jacobian in; desc + kp out
    if kernel == "back_project":
        return K * (28 + 24 + 1)
    if kernel == "prepare_landmarks":
        return L * (32 + 8) + M * (48 + 48 + 4 + 48)            # points in; rows, rays, centres, owners out
    return (K + M) * 48 + K * 16                    # any matcher: descriptors in, (dist, id) out


def wl_ptr(arr):
    """Address of a (C-contiguous) numpy image stack."""
    assert arr.flags["C_CONTIGUOUS"]
    return arr.ctypes.data


class GpuConfig:
    """One BASELINE.json frame config on this rank's GPU: handles, pools, batches and the step loops."""

    def __init__(self, args, mod, syn, torch, ctx, wl, rank, world, shard_cameras=False):
        self.args, self.mod, self.syn, self.torch, self.ctx, self.wl = args, mod, syn, torch, ctx, wl
        self.rank, self.world = rank, world
        conf, rig = wl["conf"], wl["rig"]
        self.conf = conf
        W, H = wl["W"], wl["H"]
        # camera-sharded multiframe (configs[2] under --gpus N): camera c lives on rank c mod N
        self.cams_here = [c for c in range(wl["n_cams"]) if (not shard_cameras) or c % world == rank]
        self.shard_cameras = shard_cameras
        self.nc = len(self.cams_here)
        self.max_kp = rig["max_kp"]
        self.model = mod.Pinhole.make(W, H, rig["f"], rig["c"], rig["distortion"], rig["k"])
        self.cams = [mod.Camera(ctx, W, H, wl["rays"], wl["jac"], focal=rig["f"][0], max_keypoints=self.max_kp)
                     for _ in self.cams_here]
        for cam in self.cams:
            cam.set_geometry(self.model)
        c = wl["cfg"]
        self.det = (c["radius"], 0, c["abs_threshold"], c["max_kpt"])
        self.slices = c["slices"]
        self.F = args.frames_per_batch if conf["index"] != 2 else max(1, (args.frames_per_batch * 2) // wl["n_cams"])
        self.F_shard = self.F      # block size of the frame sharding (latency() temporarily runs with F = 1)
        self.D = args.in_flight
        self.n_frames = conf["frames"]
        self.pool = None
        self.kfs = []
        self.gather = None

    # ---- the world: keyframes through the product itself (their data also goes to the CPU arm as input) ----
    def build_world(self):
        mod, wl = self.mod, self.wl
        if not wl["kf_frames"]:
            return build_world(self.syn, wl, {})
        all_cams = [mod.Camera(self.ctx, wl["W"], wl["H"], wl["rays"], wl["jac"], focal=wl["rig"]["f"][0],
                               max_keypoints=self.max_kp) for _ in range(wl["n_cams"])]
        for cam in all_cams:
            cam.set_geometry(self.model)
        b = mod.Batch(self.ctx, all_cams)
        views = {}
        for k, f in enumerate(wl["kf_frames"]):
            b.enqueue([wl["images"][c, f] for c in range(wl["n_cams"])], *self.det, [DIRECTION] * wl["n_cams"])
            b.set_poses(np.stack([mod.slot_pose(np.eye(3), wl["r_all"][f, c]) for c in range(wl["n_cams"])]))
            kf = mod.Keyframe(self.ctx, wl["n_cams"], self.max_kp)
            b.retain(list(range(wl["n_cams"])), kf)
            b.finish()
            for c in range(wl["n_cams"]):
                kp, desc = b.slot_frame(c)
                e_C, valid, _ = b.slot_back_projections(c, len(kp))
                views[(f, c)] = dict(kp=kp.copy(), desc=desc.copy(), e_C=e_C.copy(), valid=valid.copy())
            self.kfs.append(kf)
        b.close()
        for cam in all_cams:
            cam.close()
        build_world(self.syn, wl, views)
        if self.conf["landmarks"]:
            self.pool = mod.LandmarkPool(self.ctx, *wl["pool_args"])
        for k in range(len(wl["kf_frames"])):
            for c in range(wl["n_cams"]):
                v = wl["views"][k * wl["n_cams"] + c]
                self.kfs[k].set_view(c, np.eye(3), v["r_WC"], wl["f_mean"], wl["use0"][k * wl["n_cams"] + c])
        return wl

    # ---- batches ----
    def setup(self):
        mod, torch, wl = self.mod, self.torch, self.wl
        F, nc = self.F, self.nc
        # slot order inside a batch: camera-major (slot j * F + i = frame i of this rank's j-th camera), so that
        # consecutive frames of one camera are adjacent in the pinned host pool and travel in one copy
        self.batches = [mod.Batch(self.ctx, [self.cams[s // F] for s in range(nc * F)]) for _ in range(self.D)]
        for b in self.batches:
            b.configure(*self.det, [DIRECTION] * (nc * F))
            b.set_graph(not self.args.no_graph)
        self.d_img = [torch.from_numpy(wl["images"][c]).cuda() for c in self.cams_here]
        self.h_img = [torch.from_numpy(wl["images"][c]).pin_memory() for c in self.cams_here]
        self.img_stride = wl["W"] * wl["H"]
        self.streams = [torch.cuda.ExternalStream(b.stream()) for b in self.batches]
        self.pairs = []
        if self.conf["stereo"] and not self.shard_cameras:
            self.pairs = [(j * F + i, (j + 1) * F + i) for i in range(F) for j in range(nc - 1)]
        self.jobs = []
        if self.conf["motion"]:
            self.jobs = [(self.kfs[k], self.cams_here[j], j * F + i) for i in range(F) for k in range(len(self.kfs))
                         for j in range(nc)]
        self.outs, self.stage_outs = {}, {}

    def frame_of(self, k):
        """Frame index of this rank's k-th frame. Frames are sharded in blocks of F consecutive frames (one batch),
        blocks round-robin over the ranks: a batch's images are then adjacent in the host pool and travel in one copy
        per camera, as they do on one GPU (frame-wise round-robin made that 64 copies of 361 KB per batch and cost a
        third of the 2-GPU end-to-end throughput). Cameras sharded: every rank sees every frame."""
        if self.shard_cameras:
            return k % self.n_frames
        F = self.F_shard
        return ((k // F) * self.world * F + self.rank * F + k % F) % self.n_frames

    def slot_args(self, k0, device, pageable=False):
        """(image addresses, poses) of the batch of frames k0 .. k0 + F - 1: images in HBM (device), in pinned host
        memory, or in ordinary pageable host memory (numpy arrays: the library stages them through its pinned block)."""
        F, nc = self.F, self.nc
        f = np.asarray([self.frame_of(k) for k in range(k0, k0 + F)])
        imgs = np.empty(nc * F, np.uint64)
        poses = np.empty((nc * F, 24))
        for j, c in enumerate(self.cams_here):
            base = wl_ptr(self.wl["images"][c]) if pageable else (self.d_img[j] if device else self.h_img[j]).data_ptr()
            imgs[j * F:(j + 1) * F] = base + f.astype(np.uint64) * np.uint64(self.img_stride)
            poses[j * F:(j + 1) * F] = self.pose_rows[f, c]
        return imgs, poses

    def enqueue(self, b, imgs, poses, device, k0=0):
        if self.conf["landmarks"] or self.pairs or self.jobs:
            b.set_poses(poses)   # (before the enqueue: the slot table then travels once per frame)
        b.enqueue_addrs(imgs, device)
        if self.conf["landmarks"]:
            b.match_landmarks_enqueue(self.pool, HAMMING, self.slices, self.mod.STAGE_UNINITIALISED)
        if self.pairs:
            b.match_stereo_enqueue(self.pairs, HAMMING)
        if self.jobs:
            b.match_motion_enqueue(self.jobs, HAMMING, only_unmatched=True)
        if self.gather is not None:   # camera-sharded multiframe: view exchange + cross-camera matchStereo
            self.gather.step(self.batches.index(b), k0)

    def launches_per_batch(self):
        """Kernels of this library launched per batch (memsets and copies not counted)."""
        n = 6                                        # harris_nms, order, select, sub-pixel, border_filter, describe
        n += 1                                       # back-projections
        if self.conf["landmarks"]:
            n += 4 + 3 + 2                           # prepare (4), bin + match + stats, uninitialised: filter + resolve
        n += 2 if self.pairs else 0                  # matchStereo: filter + resolve
        n += 3 if self.jobs else 0                   # unmatched lists, matchMotionStereo: filter + resolve
        n += 3 if self.gather is not None else 0     # pack views, cross-camera matchStereo (+ one NCCL all-gather)
        return n

    def read(self, b):
        """The results of a finished batch: zero-copy views of the batch's pinned arena (okc_batch_view_results) in the
        key layout of read_all / read_stages; (r, s) as the copying readers returned them."""
        v = b.view()
        r = {"n": v["n"], "kp": v["kp"], "desc": v["desc"]}
        if "dist" in v:
            with np.errstate(invalid="ignore", divide="ignore"):
                r.update(dist=v["dist"], owner=v["owner"], ctrs=v["stats"]["count"],
                         errs=v["stats"]["sum"] / v["stats"]["count"].astype(np.float64))
        self.outs[id(b)] = r
        s = {k: v[k] for k in v if k[:2] in ("u_", "s_", "m_")} or None
        self.stage_outs[id(b)] = s
        return r, s

    def precompute_poses(self):
        wl = self.wl
        self.pose_rows = np.empty((self.n_frames, wl["n_cams"], 24))
        for f in range(self.n_frames):
            for c in range(wl["n_cams"]):
                self.pose_rows[f, c] = self.mod.slot_pose(np.eye(3), wl["r_all"][f, c])

    # ---- timed loops ----
    def run_value(self, n_steps, n_warm, barrier, max_over_ranks):
        torch, F, D = self.torch, self.F, self.D
        # set-up, untimed and apart from the W warm-up steps: every batch in flight runs its launch chain three times
        # (plain launches at first sight of a configuration, graph capture at the second, first replay), so that a short
        # run (--steps 20 --warmup 5 covers batch 0 twice and the others once) has no graph capture in its timed steps
        for rep in range(3):
            for j in range(D):
                self.enqueue(self.batches[j], *self.slot_args(j * F, True), True, j * F)
            for b in self.batches:
                b.finish()
        for j in range(n_warm):
            self.enqueue(self.batches[j % D], *self.slot_args(j * F, True), True, j * F)
        work = [self.slot_args((n_warm + j) * F, True) for j in range(n_steps)]
        for b in self.batches:
            b.finish()
        barrier()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        cur = torch.cuda.current_stream()
        start.record(cur)
        for s in self.streams:
            s.wait_event(start)
        # a batch is finished (results in pinned host memory) before it is enqueued again: D batches in flight, as a
        # caller that consumes the results runs it (deeper queues per stream measured 0-10 % slower)
        t_host = time.perf_counter()
        for j, (imgs, poses) in enumerate(work):
            if j >= D:
                self.batches[j % D].finish()
            self.enqueue(self.batches[j % D], imgs, poses, True, (n_warm + j) * F)
        t_host = time.perf_counter() - t_host
        for s in self.streams:
            e = torch.cuda.Event()
            e.record(s)
            cur.wait_event(e)
        end.record(cur)
        end.synchronize()
        ms = max_over_ranks(start.elapsed_time(end))
        barrier()
        for b in self.batches:
            b.finish()
        return ms, t_host / max(n_steps, 1)

    def run_e2e(self, k_begin, k_end, keep=None, pageable=False):
        F, D = self.F, self.D
        pending = []
        for j, k in enumerate(range(k_begin, k_end, F)):
            b = self.batches[j % D]
            if len(pending) == D:
                pb, pk = pending.pop(0)
                pb.finish()
                r = self.read(pb)
                if keep is not None:
                    keep(pk, r)
            self.enqueue(b, *self.slot_args(k, False, pageable), False, k)
            pending.append((b, k))
        for pb, pk in pending:
            pb.finish()
            r = self.read(pb)
            if keep is not None:
                keep(pk, r)

    def time_e2e(self, n_steps, n_warm, barrier, max_over_ranks, pageable=False):
        F = self.F
        self.run_e2e(0, max(3 * self.D, n_warm) * F, None, pageable)   # (3 uses per batch: plain, capture, replay)
        runs = []
        for _ in range(E2E_REPEATS):
            barrier()
            t0 = time.perf_counter()
            self.run_e2e(n_warm * F, (n_warm + n_steps) * F, None, pageable)
            self.torch.cuda.synchronize()
            runs.append(max_over_ranks((time.perf_counter() - t0) * 1e3))
        return float(np.median(runs)), runs

    def bytes_per_frame(self):
        """(H2D, D2H) bytes one multiframe moves in the e2e arm (this rank's cameras): counted by the library at the
        copies it issues (okc_batch_transfer_bytes), mean over the batches of the last e2e run."""
        t = [b.transfer_bytes() for b in self.batches]
        return (sum(x[0] for x in t) / len(t) / self.F, sum(x[1] for x in t) / len(t) / self.F)

    def kernel_times(self, n_batches):
        b0 = self.batches[0]
        b0.kernel_timing(True)
        acc, cnt = {}, 0
        for j in range(n_batches):
            self.enqueue(b0, *self.slot_args(j * self.F, True), True, j * self.F)
            for name, ms in b0.kernel_times():
                acc[name] = acc.get(name, 0.0) + ms
            cnt += 1
        b0.kernel_timing(False)
        b0.finish()
        return {n: v / cnt for n, v in acc.items()}

    def host_enqueue_us(self, n=40):
        """Host time of enqueueing one batch (every stage call; device-resident inputs, the batch idle before each
        enqueue so that nothing blocks): median wall clock."""
        b = self.batches[0]
        t = []
        for i in range(n + 5):
            imgs, poses = self.slot_args(i * self.F, True)
            t0 = time.perf_counter()
            self.enqueue(b, imgs, poses, True, i * self.F)
            t.append((time.perf_counter() - t0) * 1e6)
            b.finish()
        return float(np.median(t[5:]))

    def latency(self, n=60, graph=True):
        """One multiframe alone through the C ABI with host buffers: wall clock around enqueue / finish / read."""
        F_save, batches_save, streams_save, pairs_save, jobs_save = self.F, self.batches, self.streams, self.pairs, self.jobs
        gather_save, self.gather = self.gather, None   # (the latency of this rank's cameras alone)
        self.F = 1
        self.setup_single()
        self.batches[0].set_graph(graph)
        lat = []
        for i in range(n + 10):
            imgs, poses = self.slot_args(i, False)
            t1 = time.perf_counter()
            self.enqueue(self.batches[0], imgs, poses, False)
            self.batches[0].finish()
            self.read(self.batches[0])
            lat.append((time.perf_counter() - t1) * 1e3)
        self.batches[0].close()
        self.F, self.batches, self.streams, self.pairs, self.jobs = F_save, batches_save, streams_save, pairs_save, jobs_save
        self.gather = gather_save
        self.outs, self.stage_outs = {}, {}
        return float(np.median(lat[10:]))

    def setup_single(self):
        nc = self.nc
        b = self.mod.Batch(self.ctx, [self.cams[j] for j in range(nc)])
        b.configure(*self.det, [DIRECTION] * nc)
        self.batches = [b]
        self.pairs = [(j, j + 1) for j in range(nc - 1)] if (self.conf["stereo"] and not self.shard_cameras) else []
        self.jobs = [(self.kfs[k], self.cams_here[j], j) for k in range(len(self.kfs)) for j in range(nc)] \
            if self.conf["motion"] else []
        self.outs, self.stage_outs = {}, {}

    def close(self):
        for b in self.batches:
            b.close()
        for k in self.kfs:
            k.close()
        if self.pool:
            self.pool.close()
        for c in self.cams:
            c.close()
        self.d_img = self.h_img = None


def parity_check(gc, cpu, n_frames):
    """The product's e2e results on the first frames against the CPU pipeline fed with the SAME keypoints and
    descriptors (so that every later stage sees bit-identical inputs); detect / describe against the reference's own
    output. Returns (multiframes checked, mismatching items, items)."""
    wl, F, nc = gc.wl, gc.F, gc.nc
    got = {}

    bp = {}
    bp_tol = 1e-12 if wl["rig"]["distortion"] == "equidistant" else 0.0

    def keep(k0, rs):
        r, s = rs
        b = [x for x in gc.batches if id(x) in gc.outs and gc.outs[id(x)] is r][0]
        rc = {k: v.copy() for k, v in r.items()}     # (the views alias the batch's arena: the next enqueue overwrites it)
        sc = None if s is None else {k: v.copy() for k, v in s.items()}
        for i in range(F):
            got[k0 + i] = (i, rc, sc)
            if k0 + i < n_frames:
                for j in range(nc):
                    e_C, valid, _ = b.slot_back_projections(j * F + i, int(r["n"][j * F + i]))
                    bp[(k0 + i, j)] = (e_C.copy(), valid.copy())
    gc.run_e2e(0, ((n_frames + F - 1) // F) * F, keep)
    exts = cpu.extractors()
    bad = items = 0

    def check(ok):
        nonlocal bad, items
        items += 1
        bad += 0 if ok else 1
    slices = gc.slices
    for k in range(n_frames):
        i, r, s = got[k]
        f = gc.frame_of(k)
        cur = []
        for j, c in enumerate(gc.cams_here):
            slot = j * F + i
            n = int(r["n"][slot])
            kp, desc = r["kp"][slot, :n], r["desc"][slot, :n]
            want = cpu.view_of(f, c, exts[c])
            check(len(want["kp"]) == n and np.array_equal(want["desc"], desc)
                  and np.array_equal(want["kp"]["response"], kp["response"])
                  and float(np.abs(want["kp"]["x"] - kp["x"]).max(initial=0)) <= 1e-4
                  and float(np.abs(want["kp"]["y"] - kp["y"]).max(initial=0)) <= 1e-4
                  and float(np.abs(want["kp"]["angle"] - kp["angle"]).max(initial=0)) <= 1e-4)
            pe, pvalid, _ = cpu.port.back_project(cpu.model, kp)
            e_C, valid = bp[(k, j)]
            # radial-tangential: +, -, *, / only -> bit-exact; equidistant: atan / sqrt of glibc vs the device, last ulp
            check(np.array_equal(valid, pvalid) and ((e_C.tobytes() == pe.tobytes()) if bp_tol == 0.0
                                                     else float(np.abs(e_C - pe).max(initial=0)) <= bp_tol))
            cur.append(dict(kp=kp, desc=desc, e_C=e_C, valid=valid))
        st = cpu.stages(f, cur, False) if gc.cams_here == list(range(wl["n_cams"])) else None
        if st is None:
            continue
        for j in range(nc):
            slot = j * F + i
            n = int(r["n"][slot])
            if gc.conf["landmarks"]:
                lm = st["lm"][j]
                check(np.array_equal(r["dist"][slot, :n], lm["a14"][0]) and np.array_equal(r["owner"][slot, :n], lm["a14"][1]))
                ctrs = r["ctrs"].reshape(-1)[slot * slices:(slot + 1) * slices]
                errs = r["errs"].reshape(-1)[slot * slices:(slot + 1) * slices]
                # (equidistant cameras: the projections go through atan, so the summed reprojection errors agree to 1e-9)
                check(np.array_equal(ctrs, lm["a14"][2]) and
                      (np.array_equal(errs, lm["a14"][3], equal_nan=True) if bp_tol == 0.0
                       else np.allclose(errs, lm["a14"][3], rtol=1e-9, atol=0.0, equal_nan=True)))
                uc = s["u_ctrs"].reshape(-1)[slot * slices:(slot + 1) * slices]
                check(np.array_equal(s["u_dist"][slot, :n], lm["a15"][0]) and np.array_equal(s["u_owner"][slot, :n], lm["a15"][1])
                      and np.array_equal(s["u_written"][slot, :n], lm["a15"][3]) and np.array_equal(uc, lm["a15"][4])
                      and np.array_equal(s["u_hps"][slot, :n, :3], lm["a15"][2][:, :3]))
        for p, (s0, s1) in enumerate(gc.pairs):
            if s0 % F != i:
                continue
            n0 = int(r["n"][s0])
            w = st["stereo"][s0 // F]
            check(np.array_equal(s["s_dist"][p, :n0], w[0]) and np.array_equal(s["s_k1"][p, :n0], w[1])
                  and np.array_equal(s["s_hp"][p, :n0], w[2]) and np.array_equal(s["s_init"][p, :n0], w[3]))
        nk = len(gc.kfs)
        for q, (kf, c, slot) in enumerate(gc.jobs):
            if slot % F != i:
                continue
            k_idx = (q // nc) % nk
            w = st["motion"][k_idx * nc + slot // F]
            n0 = int(s["m_n0"][q])
            check(n0 == len(w[0]) and np.array_equal(s["m_dist"][q, :n0], w[0]) and np.array_equal(s["m_k1"][q, :n0], w[1])
                  and np.array_equal(s["m_hp"][q, :n0], w[2]) and np.array_equal(s["m_init"][q, :n0], w[3])
                  and float(np.abs(s["m_quality"][q, :n0] - w[4]).max(initial=0)) <= 1e-12)
    for e in exts:
        e.close()
    return n_frames, bad, items


def cpu_baseline(cpu, n_frames_pool, frame_of, budget_s):
    """The CPU code with the reference's own threading on a bounded sample of the same workload."""
    exts = cpu.extractors()
    for f in range(2):
        cpu.frame_like_reference(exts, frame_of(f))
    n, t_cpu = 0, 0.0
    while n < n_frames_pool and t_cpu < budget_s:
        t0 = time.perf_counter()
        cpu.frame_like_reference(exts, frame_of(n))
        t_cpu += time.perf_counter() - t0
        n += 1
    for e in exts:
        e.close()
    wl = cpu.wl
    return {"value": round(n / t_cpu, 2), "unit": "frames/s", "cores": max(wl["n_cams"], wl["cfg"]["slices"]),
            "kind": cpu.kind,
            "sample": "%d multiframes of the same sequence; threading as the reference: %d threads (one per camera) for "
                      "detect+describe, %d matching slices per camera for the map matchers, the 2D-2D matchers on one "
                      "thread; host has %d cores" % (n, wl["n_cams"], wl["cfg"]["slices"], os.cpu_count())}


def run_frame_config(args, name, mod, syn, torch, dist, ctx, rank, world, headline, sampler_box):
    local = int(os.environ.get("LOCAL_RANK", 0))
    wl = build_images(syn, name, args.sparse)
    shard_cameras = CONFIGS[name]["index"] == 2 and world > 1
    group = None
    if shard_cameras and world > wl["n_cams"]:
        # more GPUs than cameras: camera c -> rank c, the other ranks sit this config out (a multiframe has no more
        # independent units); collectives of the config run on the members' group
        group = dist.new_group(list(range(wl["n_cams"])))
        if rank >= wl["n_cams"]:
            dist.barrier()
            return None, None
        world = wl["n_cams"]
    gc = GpuConfig(args, mod, syn, torch, ctx, wl, rank, world, shard_cameras)
    gc.group = group
    gc.build_world()
    gc.precompute_poses()
    gc.setup()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group=group)
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        return float(t.item())

    n_steps = max(args.steps, 1) if headline else max(min(args.steps, 60), 1)
    n_warm = max(args.warmup, 3)
    if headline and rank == 0:
        # the clock sampler starts before the warm-up so that nvidia-smi's start-up (which takes driver locks) is over
        # when the timed region begins; it keeps sampling every 100 ms until after the e2e arm
        sampler_box.append(ClockSampler(local))
    F = gc.F
    without_exchange = None
    if shard_cameras and gc.conf["stereo"]:
        # first the step without the exchange (how the per-camera work scales), then the complete config
        ms0, _ = gc.run_value(min(n_steps, 20), n_warm, barrier, max_over_ranks)
        without_exchange = {"value": round(min(n_steps, 20) * F / (ms0 * 1e-3), 2), "unit": "frames/s",
                            "what": "the per-camera stages only (no view exchange, no cross-camera matchStereo)"}
        sh = __import__("importlib").import_module("okvis2x_b200.sharding")
        gc.gather = sh.CameraGather(mod, torch, dist, ctx, gc, HAMMING)
    dev_ms, t_host = gc.run_value(n_steps, n_warm, barrier, max_over_ranks)
    e2e_ms, e2e_runs = gc.time_e2e(n_steps, n_warm, barrier, max_over_ranks)
    h2d, d2h = gc.bytes_per_frame()     # (of the e2e run's last batches)
    pageable_ms = None
    if headline and world == 1:   # the same arm with the images in pageable host memory (staged by the library)
        pageable_ms, _ = gc.time_e2e(min(n_steps, 60), n_warm, barrier, max_over_ranks, pageable=True)
    barrier()
    clocks = sampler_box.pop().stop() if (headline and rank == 0 and sampler_box) else None
    lat = gc.latency(graph=not args.no_graph) if (rank == 0 or world == 1) else None
    lat_plain = gc.latency(graph=False) if (rank == 0 or world == 1) and not args.no_graph else None
    t_enq = gc.host_enqueue_us() if gc.gather is None else None
    kms = gc.kernel_times(min(n_steps, 20))
    frames_per_step = F * (1 if shard_cameras else world)
    total = n_steps * frames_per_step
    gc.run_e2e(0, F)   # one batch so that the arena of batch 0 holds a frame to take K from
    r0 = gc.outs[id(gc.batches[0])]
    K = float(np.mean(r0["n"]))
    L = len(wl["pool_args"][1]) if gc.conf["landmarks"] else 0
    M = 3.0 * L
    peak, peak_src = peaks()
    imgs_per_launch = gc.nc * F
    per_kernel = {}
    for kname, ms in kms.items():
        alg = stage_bytes(kname, wl["W"], wl["H"], K, M, L) * imgs_per_launch
        per_kernel[kname] = {"ms": round(ms, 5), "algorithmic_bytes": int(alg), "gbs": round(alg / (ms * 1e-3) / 1e9, 2),
                             "frac": round(alg / (ms * 1e-3) / 1e9 / peak, 6)}
    dom = max(kms, key=kms.get)
    roofline = {"bound": "hbm", "kernel": dom, "achieved": per_kernel[dom]["gbs"], "peak": peak, "unit": "GB/s",
                "frac": per_kernel[dom]["frac"], "traffic": ncu_traffic(dom, imgs_per_launch), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": per_kernel[dom]["algorithmic_bytes"], "images_per_launch": imgs_per_launch,
                "dominant": "the longest kernel of the chain (every kernel is listed in `kernels`)", "kernels": per_kernel,
                "note": "per-launch times of one batch running alone; in the timed loop %d batches overlap" % gc.D}
    out = {
        "workload": gc.conf["what"], "value": round(total / (dev_ms * 1e-3), 2), "unit": "frames/s",
        "ms_per_step": round(dev_ms / n_steps, 5), "steps": n_steps,
        "step": "one batch = %d multiframes (%d images on this rank) through the whole launch chain" % (F, imgs_per_launch),
        "frames_in_pool": gc.n_frames, "keypoints_per_image": round(K, 1), "landmarks": L,
        "sharding": ("cameras over ranks (camera c -> rank c mod N), keypoints + descriptors all-gathered for matchStereo"
                     if shard_cameras else "batches of consecutive frames round-robin over ranks"),
        "host_enqueue_us_per_batch": round(t_enq, 1) if t_enq is not None else None,
        "launch_chains": "plain launches" if args.no_graph else "CUDA graphs (one per chain, captured at the second use)",
        "e2e": {"value": round(total / (e2e_ms * 1e-3), 2), "unit": "frames/s", "ms_per_step": round(e2e_ms / n_steps, 5),
                "h2d_bytes_per_step": int(h2d * F), "d2h_bytes_per_step": int(d2h * F),
                "timed": "median of %d runs of K steps (ms: %s)" % (E2E_REPEATS, ", ".join("%.1f" % v for v in e2e_runs)),
                "pcie_gbs_per_gpu": round(n_steps * F * (h2d + d2h) / (e2e_ms * 1e-3) / 1e9, 1),
                "pcie_ceiling_gbs_per_gpu": pcie_ceiling(world if not shard_cameras else min(world, wl["n_cams"])),
                "pcie_ceiling_source": "profiles/r02_pcie_ceiling.json (copies only, same block sizes, same N)",
                "host_memory": "pinned (cudaMallocHost ring of frames); pageable: see e2e_pageable"},
        "gpu_launches": gc.launches_per_batch() * n_steps,
        "roofline": roofline,
    }
    if pageable_ms is not None:
        out["e2e_pageable"] = {"value": round(min(n_steps, 60) * frames_per_step / (pageable_ms * 1e-3), 2), "unit": "frames/s",
                               "what": "the e2e arm with the caller's images in pageable host memory: the library "
                                       "copies them into its pinned staging block (host memcpy) before the DMA"}
    if without_exchange is not None:
        out["without_exchange"] = without_exchange
        out["exchange"] = gc.gather.describe()
    if lat is not None:
        out["single_multiframe_latency_ms"] = round(lat, 4)
        if lat_plain is not None:
            out["single_multiframe_latency_ms_plain_launches"] = round(lat_plain, 4)
    if clocks is not None:
        out["clocks"] = clocks
    st = dict(gc=gc, wl=wl, F=F)
    if group is not None:
        dist.barrier()     # (the ranks without a camera wait here)
        out["sharding"] += "; %d of the ranks hold a camera" % world
    return out, st


def pcie_ceiling(n_gpus):
    """The committed copy-only ceiling of the box for n_gpus ranks (tools/pcie_ceiling.py), GB/s per GPU with both
    directions active; None when no run for that N is committed."""
    p = os.path.join(ROOT, "profiles", "r02_pcie_ceiling.json")
    if not os.path.exists(p):
        return None
    for r in json.load(open(p))["runs"]:
        if r["n_gpus"] == n_gpus:
            return r["both_gbs_per_gpu"]
    return None


def ncu_traffic(kernel, images_per_launch):
    """DRAM bytes per launch of `kernel` from the newest committed ncu capture (profiles/*/traffic.json), or None when
    there is none for this batch size. Never measured under this run: it is the committed profile's number."""
    import glob
    key = {"border_filter+describe": "describe_kernel", "match_to_map": "match_to_map_binned",
           "select_keypoints": "select_keypoints_wave"}.get(kernel, kernel)
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r02*", "traffic.json")), reverse=True):
        try:
            d = json.load(open(path))
            if d.get("images_per_launch") != images_per_launch:
                continue
            for name, v in d.get("kernels", {}).items():
                if name.startswith(key):
                    return v["dram_bytes_per_launch"]
        except (OSError, ValueError, KeyError):
            continue
    return None


def run_sweep(torch, dist, mod, ctx, syn, rank, world, variant="iid", iters=5):
    """BASELINE config 5 (the metric's second half): 10k x 1M Hamming sweep, map rows sharded over ranks, per-shard
    bests min-reduced over NCCL. L2 is flushed between iterations (the 48 MB map would fit in it)."""
    nq, nm = 10000, 1000000
    import importlib
    sh = importlib.import_module("okvis2x_b200.sharding")
    q, m, note = sh.sweep_inputs(syn, variant, nq, nm, mod, ctx)
    lo, hi = sh.map_row_shard(rank, world, nm)
    qs, ms = ctx.descriptor_set(q), ctx.descriptor_set(m[lo:hi])
    best = torch.empty(nq, dtype=torch.int64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    # a non-default stream: handle 0 would select the library's own stream, which torch events do not see
    cur = torch.cuda.Stream()
    torch.cuda.set_stream(cur)
    times, ktimes = [], []
    res = None
    for it in range(iters + 2):
        flush.fill_(it & 0xff)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record(cur)
        ctx.match_sets_ordered(qs, ms, 60, lo, best.data_ptr(), cur.cuda_stream)
        e1.record(cur)
        res = sh.merge_best(best, world, dist)   # one NCCL all-reduce(min) on the keys (identity at world == 1)
        e2.record(cur)
        e2.synchronize()
        if it >= 2:
            times.append(e0.elapsed_time(e2))
            ktimes.append(e0.elapsed_time(e1))
    torch.cuda.set_stream(torch.cuda.default_stream())
    t = float(np.median(times))
    if world > 1:
        tt = torch.tensor([t], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t = float(tt.item())
    peak, _ = peaks()
    pairs = float(nq) * nm
    alg = (nq + nm) * 48 + nq * 8
    matched = int((res != sh.ORDERED_NO_MATCH).sum().item())
    # fraction of the pairs whose 6-word lower bound is below the threshold, i.e. that take the exact 12-word path
    # (a property of the input, computed on a sample on the host)
    exact = sh.exact_path_fraction(q, m, 60)
    for s in (qs, ms):
        s.close()
    del flush
    return {"workload": "configs[4]: 10k x 1M descriptors (48 B), threshold 60, map rows sharded over ranks",
            "input": note, "gpairs_per_s": round(pairs / (t * 1e-3) / 1e9, 2), "ms": round(t, 4),
            "kernel_ms_rank0": round(float(np.median(ktimes)), 4), "matched": matched,
            "pairs_on_exact_path_frac": exact,
            "hbm_algorithmic_gbs": round(alg / (t * 1e-3) / 1e9, 2), "hbm_frac": round(alg / (t * 1e-3) / 1e9 / peak, 5),
            "l2": "flushed between iterations (256 MB write)", "scaling": "strong"}


def run_b200(args, rank, world):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (okvis_cuda has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        # keep stdout to the one JSON line: NCCL prints its version banner there when NCCL_DEBUG=VERSION
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    mod = g.load()
    mod.lib()
    import importlib
    syn = importlib.import_module("okvis2x_b200.synthetic")
    ctx = mod.Context(local)
    only = [x for x in args.only.split(",") if x]
    names = [n for n in CONFIGS if (not only or n in only or n == HEADLINE)]
    names.sort(key=lambda n: (n != HEADLINE, CONFIGS[n]["index"]))   # the headline first (clock sampler, fresh box)
    results, sampler_box = {}, []
    cpu_done = {}
    for name in names:
        headline = name == HEADLINE
        out, st = run_frame_config(args, name, mod, syn, torch, dist, ctx, rank, world, headline, sampler_box)
        if st is None:
            continue
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            cpu = CpuFrontend(mod, syn, st["wl"])
            n_check = 3 if CONFIGS[name]["index"] in (2, 3) else 6
            checked, bad, items = parity_check(st["gc"], cpu, n_check)
            out["parity"] = {"multiframes_checked": checked, "items": items, "mismatching": bad,
                             "what": "keypoints (1e-4) / descriptors vs the compiled reference; back-projections, map "
                                     "matches incl. ctrs / reprErrors, uninitialised matches, matchStereo, "
                                     "matchMotionStereo bit-exact vs the CPU restatement on the same keypoints"}
            out["cpu_baseline"] = cpu_baseline(cpu, st["gc"].n_frames, st["gc"].frame_of, 12.0 if headline else 6.0)
            cpu_done[name] = True
        st["gc"].close()
        del st
        torch.cuda.empty_cache()
        results[name] = out
    head = results[HEADLINE]
    out = {
        "metric": "frontend_frames_per_s", "value": head["value"], "unit": "frames/s", "n_gpus": world,
        "steps": head["steps"], "warmup": max(args.warmup, 3), "ms_per_step": head["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32 (f32+f64 gates)",
        "data": "synthetic",
        "config": {"workload": head["workload"], "frames_in_pool": head["frames_in_pool"],
                   "keypoints_per_image": head["keypoints_per_image"], "landmarks": head["landmarks"],
                   "step": head["step"], "frames_per_batch": args.frames_per_batch, "batches_in_flight": args.in_flight,
                   "host_enqueue_us_per_batch": head["host_enqueue_us_per_batch"],
                   "scene": "sparse (400 rectangles)" if args.sparse else "dense (corner-rich texture: the keypoint budget binds)",
                   "l2": "inputs larger than L2 (144 MB image pool cycled)", "sharding": head["sharding"]},
        "e2e": head["e2e"],
        "gpu_launches": head["gpu_launches"],
        "roofline": head["roofline"],
        "clocks": head.get("clocks"),
    }
    if "cpu_baseline" in head:
        out["cpu_baseline"] = dict(head["cpu_baseline"], parity=head.get("parity"))
    extra = {"configs": {}}
    for name, r in results.items():
        r = dict(r)
        r.pop("clocks", None)
        extra["configs"]["configs[%d] %s" % (CONFIGS[name]["index"], name)] = r
    if not args.no_sweep:
        sweeps = {}
        for variant in ("iid", "natural", "adversarial"):
            sweeps[variant] = run_sweep(torch, dist, mod, ctx, syn, rank, world, variant)
        extra["configs"]["configs[4] matcher_sweep"] = sweeps["iid"]
        extra["matcher_sweep_inputs"] = sweeps
    extra["single_frame_latency"] = {
        "ms": head.get("single_multiframe_latency_ms"),
        "ms_plain_launches": head.get("single_multiframe_latency_ms_plain_launches"),
        "what": "one stereo frame alone (batch of 2 images, nothing else in flight): host images in, keypoints + "
                "descriptors + matches out, wall clock around enqueue/finish/read, median of 60"}
    if "cpu_baseline" in head and head["cpu_baseline"].get("value"):
        extra["single_frame_latency"]["reference_ms"] = round(1e3 / head["cpu_baseline"]["value"], 3)
    out["extra"] = extra
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation on all host cores (rank 0 only), headline workload."""
    if rank != 0:
        return None
    import __graft_entry__ as g
    mod = g.load()
    import importlib
    syn = importlib.import_module("okvis2x_b200.synthetic")
    wl = build_images(syn, HEADLINE, args.sparse)
    cpu = CpuFrontend(mod, syn, wl)
    exts = cpu.extractors()
    build_world(syn, wl, cpu.keyframe_views(exts))
    for e in exts:
        e.close()
    cores = os.cpu_count() or 1
    sample = max(cores * 2, 32)  # frames per step: a bounded sample of the sequence
    n_frames = CONFIGS[HEADLINE]["frames"]
    warm = max(args.warmup, 0)
    steps = max(args.steps, 1)
    # every step is one bounded sample; the whole run is kept within a few minutes by shrinking the sample, never the
    # step count: the first warm-up step calibrates the rate
    t0 = time.perf_counter()
    cpu.throughput_all_cores(range(sample), cores)
    rate = sample / (time.perf_counter() - t0)
    budget_s = 150.0
    sample = int(max(cores, min(sample, rate * budget_s / (steps + warm))))
    for _ in range(max(warm - 1, 0)):
        cpu.throughput_all_cores(range(sample), cores)
    t0 = time.perf_counter()
    for s in range(steps):
        cpu.throughput_all_cores([(s * sample + i) % n_frames for i in range(sample)], cores)
    dt = time.perf_counter() - t0
    fps = steps * sample / dt
    return {"impl": "reference", "metric": "frontend_frames_per_s", "value": round(fps, 2), "unit": "frames/s",
            "n_gpus": world, "steps": steps, "warmup": warm, "ms_per_step": round(dt / steps * 1e3, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32 (f32+f64 gates)",
            "data": "synthetic",
            "config": {"workload": CONFIGS[HEADLINE]["what"], "frames_per_step": sample,
                       "scene": "sparse (400 rectangles)" if args.sparse else "dense (corner-rich texture: the keypoint budget binds)"},
            "cpu_baseline": {"value": round(fps, 2), "unit": "frames/s", "cores": cores, "kind": cpu.kind,
                             "sample": "%d steps x %d frames, one whole frame per worker thread, %d workers"
                                       % (steps, sample, cores)},
            "e2e": {"value": round(fps, 2), "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def main():
    args = parse_args()
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner, ...) are sent to stderr
    json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        out = run_reference(args, rank, world)
        if out is not None:
            print(json.dumps(out), file=json_out, flush=True)
        return
    out = run_b200(args, rank, world)
    if rank == 0:
        print(json.dumps(out), file=json_out, flush=True)


if __name__ == "__main__":
    main()
