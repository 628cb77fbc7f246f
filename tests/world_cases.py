"""Seeded synthetic SLAM worlds for the matcher tests that start from poses and 3-D map points (the
SearchByProjection overloads, Fuse, SearchBySim3): a cloud of map points with descriptors, scale-invariance ranges and
normals, and the keypoints two cameras see of it (pixel noise, descriptor bit flips, distractors, partial stereo)."""
import numpy as np

from oracle import KP_DTYPE

EUROC = dict(K=(435.2046959714599, 435.2046959714599, 367.4517211914062, 252.2008514404297), size=(752, 480), bf=47.90639384423901)
KB8 = dict(K=(190.97847715128717, 190.9733070521226, 254.93170605935475, 256.8974428996504), size=(512, 512), bf=0.0,
           kb=(0.003482389402, 0.000715034845, -0.002053236141, 0.000202936736))   # Examples/Monocular/TUM_512.yaml


def rodrigues(a):
    th = np.linalg.norm(a)
    if th < 1e-12:
        return np.eye(3)
    k = a / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K


def T_of(R, t):
    T = np.eye(4, dtype=np.float32)
    T[:3, :3] = R.astype(np.float32)
    T[:3, 3] = t.astype(np.float32)
    return T


def project(cam, Pc):
    """pixel coordinates of camera-frame points (float64; only decides what a camera sees)"""
    fx, fy, cx, cy = cam["K"]
    x, y, z = Pc[:, 0], Pc[:, 1], Pc[:, 2]
    if "kb" in cam:
        k = cam["kb"]
        r = np.hypot(x, y)
        th = np.arctan2(r, z)
        d = th * (1 + k[0] * th**2 + k[1] * th**4 + k[2] * th**6 + k[3] * th**8)
        s = np.where(r > 1e-9, d / np.maximum(r, 1e-9), 1.0)
        return fx * s * x + cx, fy * s * y + cy
    return fx * x / z + cx, fy * y / z + cy


class World:
    def __init__(self, seed=0, n_pts=1200, cam=EUROC, nlevels=8, scale=1.2, n_dup=0):
        rng = np.random.default_rng(seed)
        self.rng, self.cam, self.nlevels, self.scale = rng, cam, nlevels, scale
        w, h = cam["size"]
        fx, fy, cx, cy = cam["K"]
        # reference camera at a random pose; points in its frustum (and a margin around it)
        self.R0 = rodrigues(rng.normal(size=3) * 0.2)
        self.t0 = rng.normal(size=3) * 0.5
        z = rng.uniform(1.5, 14.0, n_pts)
        u = rng.uniform(-0.15, 1.15, n_pts) * w
        v = rng.uniform(-0.15, 1.15, n_pts) * h
        if "kb" in cam:   # spread over the fisheye field of view
            ang = rng.uniform(0, 2 * np.pi, n_pts); rad = rng.uniform(0, 1.1, n_pts) ** 0.5 * 1.2
            Pc = np.stack([np.tan(rad) * np.cos(ang) * z, np.tan(rad) * np.sin(ang) * z, z], 1)
        else:
            Pc = np.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], 1)
        self.P = ((Pc - self.t0) @ self.R0).astype(np.float32)
        self.n = n_pts
        self.desc = rng.integers(0, 256, (n_pts, 32), dtype=np.uint8)
        Ow0 = -(self.R0.T @ self.t0)
        d = np.linalg.norm(self.P - Ow0, axis=1)
        self.level0 = rng.integers(0, nlevels, n_pts)
        self.max_dist = (d * scale ** (self.level0 + rng.uniform(-0.45, 0.45, n_pts))).astype(np.float32)
        self.min_dist = (self.max_dist / np.float32(scale ** (nlevels - 1))).astype(np.float32)
        nrm = (self.P - Ow0) / d[:, None] + rng.normal(size=(n_pts, 3)) * 0.5   # mean viewing direction: camera -> point
        self.normal = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
        self.n_obs = rng.integers(0, 6, n_pts).astype(np.int32)
        self.bad = (rng.random(n_pts) < 0.03).astype(np.uint8)
        self.angle = rng.uniform(0, 360, n_pts)
        # duplicated map points (what Fuse exists for): the last n_dup points are near-copies of earlier ones; no camera sees them
        self.dup_of = np.full(n_pts, -1, np.int64)
        if n_dup:
            src = rng.choice(n_pts - n_dup, n_dup, replace=False)
            dst = np.arange(n_pts - n_dup, n_pts)
            self.dup_of[dst] = src
            self.P[dst] = self.P[src] + rng.normal(size=(n_dup, 3)).astype(np.float32) * np.float32(0.003)
            self.desc[dst] = self.desc[src]
            for i in dst:
                for b in rng.integers(0, 256, rng.integers(0, 8)):
                    self.desc[i, b >> 3] ^= np.uint8(1 << (b & 7))
            self.max_dist[dst] = self.max_dist[src]; self.min_dist[dst] = self.min_dist[src]; self.normal[dst] = self.normal[src]
            self.bad[dst] = 0

    def pose(self, dr=0.03, dt=0.25, seed=1):
        """a camera near the reference one: (R, t) world -> camera"""
        rng = np.random.default_rng(seed)
        Rd = rodrigues(rng.normal(size=3) * dr)
        td = rng.normal(size=3) * dt
        return Rd @ self.R0, Rd @ self.t0 + td

    def observe(self, R, t, seed=2, pix_noise=0.8, flips=10, distractors=0.3, mapped=0.6, stereo=0.7, level_jitter=0.15):
        """keypoints a camera at (R, t) extracts: dict(kps, desc, u_right, mp_idx (the map point behind a keypoint or -1), assoc
        (mvpMapPoints as the tracker would hold them: a subset of mp_idx), outlier, Tcw)"""
        rng = np.random.default_rng(seed)
        cam = self.cam
        w, h = cam["size"]
        Pc = self.P.astype(np.float64) @ R.T + t
        ok = (Pc[:, 2] > 0.3) & (self.dup_of < 0)
        u, v = project(cam, np.where(ok[:, None], Pc, [0, 0, 1]))
        ok &= (u > 1) & (u < w - 2) & (v > 1) & (v < h - 2)
        idx = np.nonzero(ok)[0]
        dist = np.linalg.norm(Pc[idx], axis=1)
        lvl = np.ceil(np.log(self.max_dist[idx] / dist) / np.log(self.scale) - 1e-9).astype(int)
        lvl = lvl + (rng.random(len(idx)) < level_jitter) * rng.choice([-1, 1], len(idx))
        keep = (lvl >= 0) & (lvl < self.nlevels)
        idx, lvl = idx[keep], lvl[keep]
        n_true = len(idx)
        n_dis = int(n_true * distractors)
        n = n_true + n_dis
        kps = np.zeros(n, KP_DTYPE)
        kps["x"][:n_true] = u[idx] + rng.normal(size=n_true) * pix_noise
        kps["y"][:n_true] = v[idx] + rng.normal(size=n_true) * pix_noise
        kps["x"][n_true:] = rng.uniform(2, w - 3, n_dis)
        kps["y"][n_true:] = rng.uniform(2, h - 3, n_dis)
        kps["octave"][:n_true] = lvl
        kps["octave"][n_true:] = rng.integers(0, self.nlevels, n_dis)
        kps["angle"][:n_true] = (self.angle[idx] + rng.normal(size=n_true) * 4.0) % 360.0
        kps["angle"][n_true:] = rng.uniform(0, 360, n_dis)
        kps["size"] = 31.0 * np.float32(self.scale) ** kps["octave"]
        kps["response"] = rng.integers(7, 120, n)
        kps["class_id"] = -1
        desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        desc[:n_true] = self.desc[idx]
        nflip = rng.integers(0, flips + 1, n_true)
        for i in range(n_true):
            for b in rng.integers(0, 256, nflip[i]):
                desc[i, b >> 3] ^= np.uint8(1 << (b & 7))
        mp = np.full(n, -1, np.int32)
        mp[:n_true] = idx
        ur = np.full(n, -1.0, np.float32)
        if cam.get("bf", 0) > 0:
            zz = Pc[idx, 2]
            has = rng.random(n_true) < stereo
            ur[:n_true] = np.where(has, kps["x"][:n_true] - cam["bf"] / zz + rng.normal(size=n_true) * 0.3, -1.0)
        perm = rng.permutation(n)
        kps, desc, mp, ur = kps[perm], desc[perm], mp[perm], ur[perm]
        assoc = np.where(rng.random(n) < mapped, mp, -1).astype(np.int32)
        outlier = ((rng.random(n) < 0.05) & (assoc >= 0)).astype(np.uint8)
        return dict(kps=kps, desc=desc, u_right=ur, mp_idx=mp, assoc=assoc, outlier=outlier, Tcw=T_of(R, t), R=R, t=t)

    def observe_two_cameras(self, R, t, seed=3, paired=0.7, baseline=0.11, **kw):
        """A fisheye-STEREO frame (Frame.cc:1080-1262, Nleft != -1): the left camera at (R, t), the right one at Trl * Tcw with a
        small relative rotation; the two keypoint sets stacked as the reference stacks them (mvKeys, mvKeysRight, descriptors
        vconcat), and mvLeftToRightMatch / mvRightToLeftMatch linking keypoints that show the same map point (a `paired` share)."""
        rng = np.random.default_rng(seed + 77)
        Rrl = rodrigues(rng.normal(size=3) * 0.01)
        trl = np.array([-baseline, 0.002, -0.001])
        L = self.observe(R, t, seed=seed, **kw)
        Rg = self.observe(Rrl @ R, Rrl @ t + trl, seed=seed + 1, **kw)
        nl, nr = len(L["kps"]), len(Rg["kps"])
        l2r, r2l = np.full(nl, -1, np.int32), np.full(nr, -1, np.int32)
        where_r = {int(m): j for j, m in enumerate(Rg["mp_idx"]) if m >= 0}
        for i, mpi in enumerate(L["mp_idx"]):
            j = where_r.get(int(mpi), -1)
            if mpi >= 0 and j >= 0 and rng.random() < paired:
                l2r[i] = j; r2l[j] = i
        Tlr = np.zeros((3, 4), np.float32)           # left <- right (Frame.cc:1186-1195 derives Trl from it)
        Tlr[:, :3] = Rrl.T; Tlr[:, 3] = -(Rrl.T @ trl)
        return dict(kps=L["kps"], kps_right=Rg["kps"], desc=np.vstack([L["desc"], Rg["desc"]]), n_left=nl, l2r=l2r, r2l=r2l, Tlr=Tlr,
                    Tcw=L["Tcw"], mp_idx=np.concatenate([L["mp_idx"], Rg["mp_idx"]]), assoc=np.concatenate([L["assoc"], Rg["assoc"]]))

    def view_two_cameras(self, rm, obs, assoc=None, feat_vec=None):
        cam = self.cam
        cam2 = (1 if "kb" in cam else 0,) + tuple(cam["K"]) + tuple(cam.get("kb", (0, 0, 0, 0)))
        return rm.View(obs["kps"], obs["desc"], self.bounds(), K=cam["K"], scale_factor=self.scale, nlevels=self.nlevels, Tcw=obs["Tcw"],
                       map_point_idx=assoc, kb=cam.get("kb"), kps_right=obs["kps_right"], n_left=obs["n_left"], cam2=cam2, Tlr=obs["Tlr"],
                       left_to_right=obs["l2r"], right_to_left=obs["r2l"], feat_vec=feat_vec)

    def bounds(self):
        w, h = self.cam["size"]
        return (0.0, float(w), 0.0, float(h))

    def view(self, rm, obs, assoc=None, outlier=None, with_pose=True, feat_vec=None):
        """oracle.refmatch.View of an observation"""
        cam = self.cam
        return rm.View(obs["kps"], obs["desc"], self.bounds(), K=cam["K"], bf=cam.get("bf", 0.0), b=cam.get("bf", 0.0) / cam["K"][0],
                       u_right=obs["u_right"], scale_factor=self.scale, nlevels=self.nlevels, Tcw=obs["Tcw"] if with_pose else None,
                       map_point_idx=assoc, outlier=outlier, kb=cam.get("kb"), feat_vec=feat_vec)

    def points(self, rm):
        return rm.Points(self.P, self.desc, self.normal, self.min_dist, self.max_dist, self.n_obs, self.bad)
