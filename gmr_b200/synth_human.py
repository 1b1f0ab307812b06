"""Second synthetic workload: motion authored on a HUMAN skeleton and pushed through human forward kinematics
(SURVEY.md 8d's generator), so that - unlike gmr_b200.synth, which authors the motion on the robot and maps it back -
the targets are not reachable by construction: other proportions, other joint ranges, limbs that ask for more than
the robot can do.  The IK then ends with the residuals real retargeting leaves (the reference's own error log sits
at |e| ~ 3, test_errors.csv) and more frames run into max_iter.

Per clip: root = smooth random walk (xy speed <= 1.5 m/s, yaw rate <= 1 rad/s, hips height 0.92 +- 0.03 m before
scaling); every joint = a rotation vector whose components are sums of 3 sinusoids (0.2-2 Hz; +-0.6 rad arms and legs,
+-0.2 rad spine, neck, feet) about a standing base pose with the arms down; body height ~ U(1.55, 1.90) scales the
skeleton and is passed as actual_human_height.  Skeletons: LAFAN1 topology and bone names (22 joints + the two FootMod
rows of utils/lafan1.py:30-39) or the 55 SMPL-X joints, bone offsets of a 1.75 m person, written directly in the
loaders' output space (metres, Z-up: utils/lafan1.py:20,31-32).
Bone-frame convention: a mocap rig's bone axes are arbitrary and the IK config's rotation offsets exist to absorb them
(ik_match_table*[...][4]); here each configured body's constant frame is chosen so that the BASE pose maps onto the
robot's qpos0 orientation through its table-1 offset - what a config tuned on real data does for a standing human.
Everything is float64 numpy, seeded, deterministic, independent of the number of worker processes.
"""
from __future__ import annotations

import numpy as np

from .synth import _axis_angle, _qconj, _qmul, _qrot, _smooth_noise, batch_fk, human_body_names

# name, parent, offset (m) in the loader's output space: x = left, y = back (-forward), z = up  [(x,y,z) Y-up cm -> (x,-z,y)/100]
LAFAN_SKELETON = [
    ("Hips", -1, (0, 0, 0)),
    ("LeftUpLeg", 0, (.103, -.005, -.018)), ("LeftLeg", 1, (0, 0, -.435)), ("LeftFoot", 2, (0, 0, -.419)), ("LeftToe", 3, (0, -.140, -.075)),
    ("RightUpLeg", 0, (-.103, -.005, -.018)), ("RightLeg", 5, (0, 0, -.435)), ("RightFoot", 6, (0, 0, -.419)), ("RightToe", 7, (0, -.140, -.075)),
    ("Spine", 0, (0, .016, .069)), ("Spine1", 9, (0, 0, .126)), ("Spine2", 10, (0, 0, .123)), ("Neck", 11, (0, 0, .258)), ("Head", 12, (0, -.013, .118)),
    ("LeftShoulder", 11, (.038, 0, .196)), ("LeftArm", 14, (.113, 0, 0)), ("LeftForeArm", 15, (.330, 0, 0)), ("LeftHand", 16, (.252, 0, 0)),
    ("RightShoulder", 11, (-.038, 0, .196)), ("RightArm", 18, (-.113, 0, 0)), ("RightForeArm", 19, (-.330, 0, 0)), ("RightHand", 20, (-.252, 0, 0)),
]
_SMPLX_BODY = [
    ("pelvis", -1, (0, 0, 0)), ("left_hip", 0, (.06, 0, -.09)), ("right_hip", 0, (-.06, 0, -.09)), ("spine1", 0, (0, .01, .11)),
    ("left_knee", 1, (.04, 0, -.38)), ("right_knee", 2, (-.04, 0, -.38)), ("spine2", 3, (0, -.01, .13)),
    ("left_ankle", 4, (-.01, .02, -.40)), ("right_ankle", 5, (.01, .02, -.40)), ("spine3", 6, (0, 0, .06)),
    ("left_foot", 7, (.02, -.12, -.06)), ("right_foot", 8, (-.02, -.12, -.06)), ("neck", 9, (0, .02, .21)),
    ("left_collar", 9, (.05, .02, .12)), ("right_collar", 9, (-.05, .02, .12)), ("head", 12, (0, -.03, .09)),
    ("left_shoulder", 13, (.12, .01, .04)), ("right_shoulder", 14, (-.12, .01, .04)), ("left_elbow", 16, (.26, 0, -.01)),
    ("right_elbow", 17, (-.26, 0, -.01)), ("left_wrist", 18, (.25, 0, 0)), ("right_wrist", 19, (-.25, 0, 0)),
    ("jaw", 15, (0, -.04, -.02)), ("left_eye_smplhf", 15, (.03, -.08, .03)), ("right_eye_smplhf", 15, (-.03, -.08, .03)),
]


def smplx_skeleton():
    sk = list(_SMPLX_BODY)
    for side, wrist, sx in (("left", 20, 1.0), ("right", 21, -1.0)):
        for f, dy in (("index", -.03), ("middle", -.01), ("pinky", .03), ("ring", .01), ("thumb", -.04)):
            par = wrist
            for i, dx in ((1, .09), (2, .03), (3, .025)):
                sk.append((f"{side}_{f}{i}", par, (sx * dx, dy if i == 1 else 0.0, 0.0)))
                par = len(sk) - 1
    return sk


SKELETONS = {"bvh": LAFAN_SKELETON, "smplx": smplx_skeleton()}
# joints that swing +-0.6 rad (limbs); the rest (spine, neck, head, feet, collars, fingers) +-0.2 rad
_LIMB = ("UpLeg", "Leg", "Arm", "ForeArm", "_hip", "_knee", "_shoulder", "_elbow")
_ARM_DOWN = {"LeftArm": -1.35, "RightArm": 1.35, "left_shoulder": -1.2, "right_shoulder": 1.2}      # about y (back axis): T-pose -> arms down


def _human_fk(sk, rootp, rootq, jq, scale):
    """rootp [N,3], rootq [N,4], jq [N,J,4] local rotations (joint 0 unused), scale [N] -> pos [N,J,3], quat [N,J,4]."""
    N, J = rootp.shape[0], len(sk)
    pos, quat = np.zeros((N, J, 3)), np.zeros((N, J, 4))
    pos[:, 0], quat[:, 0] = rootp, rootq
    for j in range(1, J):
        p = sk[j][1]
        pos[:, j] = pos[:, p] + _qrot(quat[:, p], np.asarray(sk[j][2], np.float64)[None] * scale[:, None])
        quat[:, j] = _qmul(quat[:, p], jq[:, j])
    return pos, quat


def _base_local(sk):
    base = np.tile([1.0, 0, 0, 0], (len(sk), 1))
    for j, (name, _, _) in enumerate(sk):
        if name in _ARM_DOWN:
            base[j] = _axis_angle(np.array([0.0, 1.0, 0.0]), np.array(_ARM_DOWN[name]))
    return base


def bone_frames(model, tables, src_human):
    """Constant frame C_b of every body of the stream (global orientation = FK rotation x C_b): for the bodies the
    config reads, the one that maps the base pose onto the robot's qpos0 orientation through the table-1 rotation
    offset; identity for the rest."""
    sk = SKELETONS[src_human]
    names = [n for n, _, _ in sk]
    base = _base_local(sk)
    _, gq = _human_fk(sk, np.zeros((1, 3)), np.array([[1.0, 0, 0, 0]]), base[None], np.ones(1))
    _, xquat = batch_fk(model, model.qpos0[None])
    # heading: the skeleton faces -y, the robots face +x: turn the robot side by -90 deg about z
    turn = _axis_angle(np.array([0.0, 0.0, 1.0]), np.array(-np.pi / 2))
    C = np.tile([1.0, 0, 0, 0], (len(sk), 1))
    for i, b in enumerate(tables.scaled_names):
        src = b.replace("FootMod", "Toe") if b.endswith("FootMod") else b      # the joint whose orientation the row carries
        if src not in names:
            continue
        body = None
        for k in range(2):
            hit = np.nonzero(np.asarray(tables.task_src[k]) == i)[0]
            if len(hit):
                body = int(tables.task_body[k][hit[0]])
                break
        if body is None:
            continue
        j = names.index(src)
        want = _qmul(_qmul(turn, xquat[0, body]), _qconj(tables.rot_off[i]))       # q_human with q_human x rot_off = robot link
        C[j] = _qmul(_qconj(gq[0, j]), want)
    return C


def synth_human_fk_clips(model, tables, src_human: str, B: int, T: int, seed: int = 0xB210, fps: float = 30.0,
                         dtype=np.float32, chunk: int = 256):
    """-> (human [B,T,h,7] dtype, heights [B], human_names): same contract as synth.synth_human_clips."""
    sk = SKELETONS[src_human]
    J = len(sk)
    jn = [n for n, _, _ in sk]
    names = human_body_names(src_human, tables)
    C = bone_frames(model, tables, src_human)
    base = _base_local(sk)
    amp = np.array([0.6 if any(n.endswith(s) or s in n for s in _LIMB) and "Shoulder" not in n else 0.2 for n in jn])
    amp[0] = 0.0
    out = np.zeros((B, T, len(names), 7), dtype)
    heights = np.zeros(B)
    t = np.arange(T) / fps
    for c0 in range(0, B, chunk):
        nb = min(B, c0 + chunk) - c0
        rng = np.random.default_rng([seed, c0])
        hts = rng.uniform(1.55, 1.90, size=nb)
        scale = np.repeat(hts / 1.75, T)
        # root: heading random walk, speed along the heading, hips height
        yaw_rate = 1.0 * _smooth_noise(rng, (nb, T, 1), 3, fps, 0.05, 0.5)[..., 0].clip(-1, 1)
        yaw = rng.uniform(-np.pi, np.pi, size=(nb, 1)) + np.cumsum(yaw_rate, 1) / fps
        speed = 0.75 * (1.0 + _smooth_noise(rng, (nb, T, 1), 3, fps, 0.05, 0.5)[..., 0].clip(-1, 1))
        fwd = np.stack([np.sin(yaw), -np.cos(yaw)], -1)                              # the skeleton faces -y at yaw 0
        xy = np.cumsum(fwd * speed[..., None], 1) / fps
        z = (0.92 + 0.03 * _smooth_noise(rng, (nb, T, 1), 2, fps, 0.5, 2.0)[..., 0].clip(-1, 1)) * (hts / 1.75)[:, None]
        rootp = np.concatenate([xy, z[..., None]], -1).reshape(nb * T, 3)
        lean = 0.1 * _smooth_noise(rng, (nb, T, 2), 2, fps, 0.2, 1.0)
        rootq = _qmul(_axis_angle(np.array([0.0, 0, 1.0]), yaw.reshape(-1)),
                      _qmul(_axis_angle(np.array([1.0, 0, 0]), lean[..., 0].reshape(-1)), _axis_angle(np.array([0.0, 1.0, 0]), lean[..., 1].reshape(-1))))
        rv = _smooth_noise(rng, (nb, T, J * 3), 3, fps, 0.2, 2.0).reshape(nb * T, J, 3) * amp[None, :, None]
        ang = np.linalg.norm(rv, axis=-1)
        jq = _qmul(np.broadcast_to(base, (nb * T, J, 4)), _axis_angle(rv / np.maximum(ang, 1e-12)[..., None], ang))
        pos, quat = _human_fk(sk, rootp, rootq, jq, scale)
        quat = _qmul(quat, np.broadcast_to(C, quat.shape))
        arr = np.zeros((nb * T, len(names), 7))
        for k, n in enumerate(names):
            if n.endswith("FootMod"):                                                 # utils/lafan1.py:35-39: Foot position, Toe orientation
                arr[:, k, :3] = pos[:, jn.index(n.replace("Mod", ""))]
                arr[:, k, 3:] = quat[:, jn.index(n.replace("FootMod", "Toe"))]
            else:
                arr[:, k, :3], arr[:, k, 3:] = pos[:, jn.index(n)], quat[:, jn.index(n)]
        arr = arr.reshape(nb, T, len(names), 7)
        dots = np.sum(arr[:, 1:, :, 3:] * arr[:, :-1, :, 3:], axis=-1)
        flip = np.cumprod(np.where(dots < 0, -1.0, 1.0), axis=1)
        arr[:, 1:, :, 3:] *= flip[..., None]
        out[c0:c0 + nb] = arr
        heights[c0:c0 + nb] = hts
    return out, heights, list(names)


def canonical_clips(B: int, seed: int = 0xB220, dtype=np.float32, path=None):
    """Third workload: the reference's one real motion (out/test_canonical.bvh through its own loader and synonym /
    FootMod / Spine2 / orientation helpers: tools/make_canonical_clip.py -> tests/golden/canonical_clip.npz, 250 frames),
    tiled B times with a per-clip heading (rotation about z of every position and orientation), a per-clip start
    phase (the clip played cyclically from another frame would jump, so the phase only reverses the playing direction
    for odd clips) and a per-clip actual_human_height ~ U(1.55, 1.90).  -> (human [B,250,24,7], heights [B], names)."""
    import os
    if path is None:
        path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "canonical_clip.npz")
    z = np.load(path, allow_pickle=False)
    base = z["frames"]                                     # [T,h,7] float64
    names = [str(n) for n in z["names"]]
    T, h = base.shape[:2]
    rng = np.random.default_rng(seed)
    yaw = rng.uniform(-np.pi, np.pi, size=B)
    heights = rng.uniform(1.55, 1.90, size=B)
    out = np.zeros((B, T, h, 7), dtype)
    for c0 in range(0, B, 256):
        n = min(B, c0 + 256) - c0
        qz = _axis_angle(np.array([0.0, 0.0, 1.0]), yaw[c0:c0 + n])[:, None, None, :]          # [n,1,1,4]
        clip = np.broadcast_to(base, (n, T, h, 7)).copy()
        rev = (np.arange(c0, c0 + n) % 2) == 1
        clip[rev] = clip[rev][:, ::-1]
        out[c0:c0 + n, ..., :3] = _qrot(np.broadcast_to(qz, (n, T, h, 4)), clip[..., :3])
        out[c0:c0 + n, ..., 3:] = _qmul(np.broadcast_to(qz, (n, T, h, 4)), clip[..., 3:])
    return out, heights, names
