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
out["_note"] = "host-buffer calls: cudaMalloc + H2D + kernel + D2H + cudaFree per call dominate the GPU column at these sizes"
print(json.dumps(out))
