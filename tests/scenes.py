"""Seeded synthetic inputs shared by tests and bench.py: procedural room scenes, area-uniform surface
sampling with analytic normals, analytic ray-cast LiDAR scans.  Pure numpy; no reference code involved."""
import numpy as np


def rot_axis_angle(axis, angle):
    axis = np.asarray(axis, dtype=np.float64)
    axis = axis / np.linalg.norm(axis)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(angle) * K + (1 - np.cos(angle)) * (K @ K)


def rigid(R, t):
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = t
    return T


class Scene:
    """Closed room [0,L]x[0,W]x[0,H] + axis-aligned boxes + vertical cylinders."""

    def __init__(self, seed, size=(120.0, 120.0, 12.0), n_boxes=60, n_cyl=40, keepout=None, keepout_r=3.0):
        rng = np.random.default_rng(seed)
        self.size = np.array(size, dtype=np.float64)
        L, W, H = size
        e = rng.uniform(1.0, 6.0, size=(n_boxes, 3))
        e[:, 2] = np.minimum(e[:, 2], H - 0.5)
        c = np.stack([rng.uniform(3, L - 3, n_boxes), rng.uniform(3, W - 3, n_boxes), e[:, 2] / 2], axis=1)
        self.box_lo = c - e / 2
        self.box_hi = c + e / 2
        self.cyl_r = rng.uniform(0.2, 1.0, n_cyl)
        self.cyl_h = np.minimum(rng.uniform(2.0, 8.0, n_cyl), H - 0.5)
        self.cyl_c = np.stack([rng.uniform(3, L - 3, n_cyl), rng.uniform(3, W - 3, n_cyl)], axis=1)
        if keepout is not None:   # drop primitives whose footprint comes within keepout_r of the sensor path
            ko = np.atleast_2d(np.asarray(keepout, dtype=np.float64))[:, :2]
            d = np.maximum(np.maximum(self.box_lo[:, None, :2] - ko[None], ko[None] - self.box_hi[:, None, :2]), 0)
            keep = np.linalg.norm(d, axis=2).min(axis=1) > keepout_r
            self.box_lo, self.box_hi = self.box_lo[keep], self.box_hi[keep]
            dc = np.linalg.norm(self.cyl_c[:, None] - ko[None], axis=2).min(axis=1) - self.cyl_r
            keep = dc > keepout_r
            self.cyl_c, self.cyl_r, self.cyl_h = self.cyl_c[keep], self.cyl_r[keep], self.cyl_h[keep]

    # ---- area-uniform surface sampling with analytic outward normals
    def _faces(self):
        L, W, H = self.size
        faces = []  # (origin, u, v, normal)
        # room (normals point inward = toward the free space)
        faces += [((0, 0, 0), (L, 0, 0), (0, W, 0), (0, 0, 1)), ((0, 0, H), (L, 0, 0), (0, W, 0), (0, 0, -1)),
                  ((0, 0, 0), (L, 0, 0), (0, 0, H), (0, 1, 0)), ((0, W, 0), (L, 0, 0), (0, 0, H), (0, -1, 0)),
                  ((0, 0, 0), (0, W, 0), (0, 0, H), (1, 0, 0)), ((L, 0, 0), (0, W, 0), (0, 0, H), (-1, 0, 0))]
        for lo, hi in zip(self.box_lo, self.box_hi):
            d = hi - lo
            faces += [((lo[0], lo[1], hi[2]), (d[0], 0, 0), (0, d[1], 0), (0, 0, 1)),
                      ((lo[0], lo[1], lo[2]), (d[0], 0, 0), (0, 0, d[2]), (0, -1, 0)),
                      ((lo[0], hi[1], lo[2]), (d[0], 0, 0), (0, 0, d[2]), (0, 1, 0)),
                      ((lo[0], lo[1], lo[2]), (0, d[1], 0), (0, 0, d[2]), (-1, 0, 0)),
                      ((hi[0], lo[1], lo[2]), (0, d[1], 0), (0, 0, d[2]), (1, 0, 0))]
        return faces

    def sample_surface(self, n, seed):
        rng = np.random.default_rng(seed)
        faces = self._faces()
        f_o = np.array([f[0] for f in faces], dtype=np.float64)
        f_u = np.array([f[1] for f in faces], dtype=np.float64)
        f_v = np.array([f[2] for f in faces], dtype=np.float64)
        f_n = np.array([f[3] for f in faces], dtype=np.float64)
        f_area = np.linalg.norm(np.cross(f_u, f_v), axis=1)
        c_area = 2 * np.pi * self.cyl_r * self.cyl_h
        areas = np.concatenate([f_area, c_area])
        counts = rng.multinomial(n, areas / areas.sum())
        nf = len(faces)
        fi = np.repeat(np.arange(nf), counts[:nf])
        a, b = rng.random(fi.size), rng.random(fi.size)
        pts = [f_o[fi] + a[:, None] * f_u[fi] + b[:, None] * f_v[fi]]
        nrm = [f_n[fi]]
        ci = np.repeat(np.arange(len(self.cyl_r)), counts[nf:])
        th, hh = rng.uniform(0, 2 * np.pi, ci.size), rng.random(ci.size) * self.cyl_h[ci]
        cn = np.stack([np.cos(th), np.sin(th), np.zeros_like(th)], axis=1)
        pts.append(np.concatenate([self.cyl_c[ci] + self.cyl_r[ci, None] * cn[:, :2], hh[:, None]], axis=1))
        nrm.append(cn)
        pts, nrm = np.concatenate(pts), np.concatenate(nrm)
        perm = rng.permutation(pts.shape[0])
        return pts[perm], nrm[perm]

    def make_map(self, m, seed, dedupe_voxel=0.05, oversample=1.6):
        pts, nrm = self.sample_surface(int(m * oversample), seed)
        if dedupe_voxel > 0:
            key = np.floor(pts / dedupe_voxel).astype(np.int64)
            h = (key[:, 0] * 73856093) ^ (key[:, 1] * 19349663) ^ (key[:, 2] * 83492791)
            _, first = np.unique(h, return_index=True)
            first.sort()
            pts, nrm = pts[first], nrm[first]
        if pts.shape[0] < m:
            raise RuntimeError(f"only {pts.shape[0]} points after dedupe, need {m}")
        return pts[:m], nrm[:m]

    # ---- analytic ray casting (every ray hits: the room is closed)
    def raycast(self, origin, dirs):
        o = np.asarray(origin, dtype=np.float64)
        d = np.asarray(dirs, dtype=np.float64)
        n = d.shape[0]
        t = np.full(n, np.inf)
        with np.errstate(divide="ignore", invalid="ignore"):
            inv = 1.0 / d
            # room walls: exit distance of the box [0,size]
            t1 = (0 - o) * inv
            t2 = (self.size - o) * inv
            t = np.minimum(t, np.min(np.maximum(t1, t2), axis=1))
            for lo, hi in zip(self.box_lo, self.box_hi):
                ta, tb = (lo - o) * inv, (hi - o) * inv
                tn, tf = np.max(np.minimum(ta, tb), axis=1), np.min(np.maximum(ta, tb), axis=1)
                hit = (tn <= tf) & (tn > 1e-6)
                t = np.where(hit & (tn < t), tn, t)
            for c, r, h in zip(self.cyl_c, self.cyl_r, self.cyl_h):
                oc = o[:2] - c
                A = d[:, 0] ** 2 + d[:, 1] ** 2
                B = 2 * (oc[0] * d[:, 0] + oc[1] * d[:, 1])
                C = oc @ oc - r * r
                disc = B * B - 4 * A * C
                tc = (-B - np.sqrt(np.maximum(disc, 0))) / (2 * A)
                z = o[2] + tc * d[:, 2]
                hit = (disc > 0) & (tc > 1e-6) & (z >= 0) & (z <= h) & (A > 1e-12)
                t = np.where(hit & (tc < t), tc, t)
        return t


def lidar_dirs(n_beams=128, n_az=1024, fov_deg=(-22.5, 22.5)):
    el = np.deg2rad(np.linspace(fov_deg[0], fov_deg[1], n_beams))
    az = np.linspace(0, 2 * np.pi, n_az, endpoint=False)
    E, A = np.meshgrid(el, az, indexing="ij")
    return np.stack([np.cos(E) * np.cos(A), np.cos(E) * np.sin(A), np.sin(E)], axis=-1).reshape(-1, 3)


def lidar_scan(scene, pose, n_beams=128, n_az=1024, noise_sigma=0.01, seed=0):
    """Scan in the SENSOR frame (N x 3, float64); pose = 4x4 sensor-to-world."""
    dirs_s = lidar_dirs(n_beams, n_az)
    dirs_w = dirs_s @ pose[:3, :3].T
    t = scene.raycast(pose[:3, 3], dirs_w)
    rng = np.random.default_rng(seed)
    if noise_sigma > 0:
        t = t + rng.normal(0, noise_sigma, t.shape)
    return dirs_s * t[:, None]


def config2(m=5_000_000, n_beams=128, n_az=1024, scene_seed=1001, size=(120.0, 120.0, 12.0), n_boxes=60, n_cyl=40):
    """BASELINE config #2: returns map (xyz, normals), scan (sensor frame), T_true, T_init."""
    c = np.array([size[0] / 2 + 3.0, size[1] / 2 - 2.0, 1.8])
    sc = Scene(scene_seed, size=size, n_boxes=n_boxes, n_cyl=n_cyl, keepout=c)
    map_xyz, map_nrm = sc.make_map(m, scene_seed + 10)
    T_true = rigid(rot_axis_angle([0, 0, 1], np.deg2rad(17.0)), c)
    scan = lidar_scan(sc, T_true, n_beams, n_az, 0.01, scene_seed + 1)
    pert = rigid(rot_axis_angle([1, 2, 3], np.deg2rad(1.5)), [0.15, -0.10, 0.05])
    T_init = T_true @ pert
    return dict(scene=sc, map_xyz=map_xyz, map_nrm=map_nrm, scan=scan, T_true=T_true, T_init=T_init)
