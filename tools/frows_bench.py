"""GPU-box tool: wall-clock of the SURVEY 8(f) entry points (host buffers in and out, through the C ABI) next to the
CPU restatement in oracle/port on the same inputs. Prints one JSON object."""
import os, sys, json, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
import importlib, refbind
mod = ge.build()
syn = importlib.import_module("okvis2x_b200.synthetic")
ctx, port = mod.Context(0), refbind.Port()
from test_gpu_geometry import _landmark_scene, _keypoints
from test_gpu_vocabulary import make_tree


def timed(f, reps=20):
    f(); ts = []
    for _ in range(reps):
        t = time.perf_counter(); f(); ts.append(time.perf_counter() - t)
    return round(float(np.median(ts)) * 1e3, 4)

out = {}
e = syn.EUROC["cams"][0]
cam = mod.Pinhole.make(752, 480, e["f"], e["c"], "radialtangential", e["d"])
kp = _keypoints(mod, 752, 480, 700, 1)
out["back_project_700kp"] = {"gpu_ms": timed(lambda: ctx.back_project(cam, kp)), "cpu_port_ms": timed(lambda: port.back_project(cam, kp))}
views, hp_W, quality, visible, off, ov, ok = _landmark_scene(21, L=20000)
T = np.concatenate([np.eye(3).reshape(9), [0.1, 0.0, 0.2]])
out["project_landmarks_20k"] = {"gpu_ms": timed(lambda: ctx.project_landmarks(cam, T, hp_W, 30.5)),
                                "cpu_port_ms": timed(lambda: port.project_landmarks(cam, T, hp_W, 30.5))}
args = (views, hp_W, quality, visible, off, ov, ok, np.array([1.0, 0.0, -0.1]), 916.0)
out["prepare_landmarks_20k"] = {"gpu_ms": timed(lambda: ctx.prepare_landmarks(*args), 10),
                                "cpu_port_ms": timed(lambda: port.prepare_landmarks(*args), 10), "observations": int(off[-1])}
rng = np.random.default_rng(1)
w, h, offs, kps, ms = [], [], [0], [], []
for v in range(16):   # 8 keyframes x 2 cameras
    k = _keypoints(mod, 752, 480, 700, 10 + v)
    w.append(752); h.append(480); offs.append(offs[-1] + 700); kps.append(k); ms.append((rng.uniform(0, 1, 700) < 0.5).astype(np.uint8))
oa = (w, h, offs, np.concatenate(kps), np.concatenate(ms))
out["keyframe_overlap_16views"] = {"gpu_ms": timed(lambda: ctx.keyframe_overlap(*oa)), "cpu_port_ms": timed(lambda: port.keyframe_overlap(*oa))}
tree = make_tree(9, 3, 5)
voc = mod.Vocabulary(ctx, *tree, depth=3)
feats = rng.integers(0, 256, (1400, 48), dtype=np.uint8)
out["vocabulary_transform_1400"] = {"gpu_ms": timed(lambda: voc.transform(feats, 1)),
                                    "cpu_port_ms": timed(lambda: port.vocabulary_transform(*tree, 3, feats, 1), 5),
                                    "note": "the port's time includes flattening the tree in numpy"}

# ---- the same stages as DEVICE-RESIDENT pipeline stages (SURVEY 8 f1 / f2 as built in round 2): the landmark pool
# lives in HBM (okc_landmark_pool), the pose is the only per-frame input, projections / pruned pool / back-projections
# never leave the device. Per-stage device time (CUDA events around the stage's launches on the batch's stream) for a
# batch of B images, next to the CPU port doing the same work for B images one after the other.
pool = mod.LandmarkPool(ctx, views, hp_W, quality, off, ov, ok)
for B in (2, 32):
    cams = [mod.Camera(ctx, 752, 480, max_keypoints=768) for _ in range(2)]
    for c in cams:
        c.set_geometry(cam)
    batch = mod.Batch(ctx, [cams[i % 2] for i in range(B)])
    imgs = [syn.dense_image(752, 480, 40 + i) for i in range(B)]
    r_WC = [np.array([0.01 * i, 0.0, -0.2]) for i in range(B)]
    poses = np.stack([mod.slot_pose(np.eye(3), r) for r in r_WC])
    batch.kernel_timing(True)
    acc = {}
    for rep in range(12):
        batch.enqueue(imgs, 38.0, 0, 150.0, 700, [(0.0, 1.0, 0.0)] * B)
        batch.set_poses(poses)
        batch.match_landmarks_enqueue(pool, 60.0, 4, mod.STAGE_UNINITIALISED)
        batch.finish()
        if rep >= 2:
            for name, ms_ in batch.kernel_times():
                acc.setdefault(name, []).append(ms_)
    batch.kernel_timing(False)
    kp0, _ = batch.slot_frame(0)
    n_kp = len(kp0)
    lm = batch.slot_landmarks(0)

    def cpu_landmarks():
        for i in range(B):
            T = mod.slot_pose(np.eye(3), r_WC[i])[:12]
            proj, status, vis = port.project_landmarks(cam, T, hp_W, 30.5)
            port.prepare_landmarks(views, hp_W, quality, vis, off, ov, ok, r_WC[i], cam.fu + cam.fv)

    def cpu_bp():
        for i in range(B):
            port.back_project(cam, kp0)
    g_prep, g_bp = float(np.median(acc["prepare_landmarks"])), float(np.median(acc["back_project"]))
    c_prep, c_bp = timed(cpu_landmarks, 3), timed(cpu_bp, 5)
    out["device_resident_B%d" % B] = {
        "landmarks": len(hp_W), "visible_landmarks_slot0": int(lm["visible"].sum()), "pool_rows_slot0": int(lm["nm"]),
        "project+prepare_landmarks": {"gpu_stage_ms": round(g_prep, 4), "cpu_port_ms": c_prep, "speedup": round(c_prep / g_prep, 1)},
        "back_project_%dkp" % n_kp: {"gpu_stage_ms": round(g_bp, 4), "cpu_port_ms": c_bp, "speedup": round(c_bp / g_bp, 1)}}
    batch.close()
    for c in cams:
        c.close()
pool.close()
out["_note"] = "host-buffer calls: cudaMalloc + H2D + kernel + D2H + cudaFree per call dominate the GPU column at these sizes"
print(json.dumps(out))
