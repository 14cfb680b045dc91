"""Second, independent restatement of the reference math in numpy float64 (TEST INFRASTRUCTURE).

Used by the CPU tests to cross-check oracle/ndt_oracle.cpp: the oracle follows the reference's
float32 per-term arithmetic; this file follows the same formulas (reference
xchu_mapping/include/pclomp/ndt_omp_impl.hpp = NDT_I, voxel_grid_covariance_omp_impl.hpp = VGC_I)
in plain float64 and also differentiates the score numerically, so a transcription error in the
oracle's Jacobian / Hessian tables would show up as a mismatch of order 1, not 1e-7.
"""
import numpy as np


def gauss_constants(resolution, outlier_ratio):
    """NDT_I:55-62 / 80-87."""
    c1 = 10.0 * (1.0 - outlier_ratio)
    c2 = outlier_ratio / resolution ** 3
    d3 = -np.log(c2)
    d1 = -np.log(c1 + c2) - d3
    d2 = -2.0 * np.log((-np.log(c1 * np.exp(-0.5) + c2) - d3) / d1)
    return d1, d2, d3


def rot(p):
    """R = Rx(p3) Ry(p4) Rz(p5), NDT_H:213-216."""
    cx, sx, cy, sy, cz, sz = np.cos(p[3]), np.sin(p[3]), np.cos(p[4]), np.sin(p[4]), np.cos(p[5]), np.sin(p[5])
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rx @ Ry @ Rz


def pose_matrix(p):
    T = np.eye(4)
    T[:3, :3] = rot(p)
    T[:3, 3] = p[:3]
    return T


def voxel_stats(pts, leaf, min_points=6, eig_mult=0.01):
    """VGC_I:209-369 in float64 except the cell index (float32 floor(x * inv_leaf), VGC_I:218-220).
    Returns dict cell(i,j,k) -> (n, mean, cov, icov or None)."""
    pts = np.asarray(pts, dtype=np.float32)[:, :3]
    inv = np.float32(1.0) / np.float32(leaf)
    ijk = np.floor(pts * inv).astype(np.int64)
    out = {}
    order = np.lexsort((ijk[:, 0], ijk[:, 1], ijk[:, 2]))
    s = ijk[order]
    cuts = np.flatnonzero(np.any(s[1:] != s[:-1], axis=1)) + 1
    for grp in np.split(order, cuts):
        x = pts[np.sort(grp)].astype(np.float64)
        n = x.shape[0]
        key = tuple(int(v) for v in ijk[grp[0]])
        if n < min_points:
            out[key] = (n, x.mean(axis=0), None, None)
            continue
        sx = x.sum(axis=0)
        sxx = np.eye(3) + x.T @ x                      # Leaf::cov_ starts at Identity (VGC_H:101)
        mean = sx / n
        # VGC_I:329-330: cov = (cov_ - 2 pt_sum mean^T) / n + mean mean^T, then * (n-1)/n
        cov = (sxx - 2.0 * np.outer(sx, mean)) / n + np.outer(mean, mean)
        cov = cov * ((n - 1.0) / n)
        w, V = np.linalg.eigh(cov)
        if w[0] < 0 or w[1] < 0 or w[2] <= 0:
            out[key] = (-1, mean, cov, None)
            continue
        lo = eig_mult * w[2]
        if w[0] < lo:                                  # VGC_I:345-356
            w[0] = lo
            if w[1] < lo:
                w[1] = lo
            cov = V @ np.diag(w) @ np.linalg.inv(V)
        icov = np.linalg.inv(cov)
        out[key] = (n, mean, cov, icov)
    return out


OFF7 = [(0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]   # VGC_I:423-430


def neighbors_direct(vox, xt, leaf, offsets, min_points=6):
    """VGC_I:373-442: query cell floor(x / leaf) in float32, then the offsets in reference order."""
    c = np.floor(np.asarray(xt, dtype=np.float32) / np.float32(leaf)).astype(np.int64)
    res = []
    for o in offsets:
        k = (int(c[0] + o[0]), int(c[1] + o[1]), int(c[2] + o[2]))
        v = vox.get(k)
        if v is not None and v[0] >= min_points:
            res.append(k)
    return res


def score_only(p, src, vox, leaf, d1, d2, offsets=OFF7):
    """sum over points and neighbour voxels of -d1 exp(-d2/2 q)  (NDT_I:505-509), float64."""
    T = pose_matrix(p)
    # the voxel lookup uses the float32 transformed point like the reference; the math is float64
    xt32 = (src[:, :3].astype(np.float32) @ T[:3, :3].astype(np.float32).T + T[:3, 3].astype(np.float32)).astype(np.float32)
    xt = src[:, :3].astype(np.float64) @ T[:3, :3].T + T[:3, 3]
    s = 0.0
    for i in range(src.shape[0]):
        for k in neighbors_direct(vox, xt32[i], leaf, offsets):
            n, mean, cov, icov = vox[k]
            d = xt[i] - mean
            s += -d1 * np.exp(-d2 * (d @ icov @ d) / 2.0)
    return s


def derivatives(p, src, vox, leaf, d1, d2, offsets=OFF7, nbr_sets=None):
    """Analytic score / gradient / Hessian (NDT_I:175-283, 286-445, 491-548) in float64.
    nbr_sets: optional list of neighbour-cell lists to hold the voxel association fixed."""
    cx, sx, cy, sy, cz, sz = np.cos(p[3]), np.sin(p[3]), np.cos(p[4]), np.sin(p[4]), np.cos(p[5]), np.sin(p[5])
    jang = np.array([                                   # NDT_I:319-326
        [-sx * sz + cx * sy * cz, -sx * cz - cx * sy * sz, -cx * cy],
        [cx * sz + sx * sy * cz, cx * cz - sx * sy * sz, -sx * cy],
        [-sy * cz, sy * sz, cy],
        [sx * cy * cz, -sx * cy * sz, sx * sy],
        [-cx * cy * cz, cx * cy * sz, -cx * sy],
        [-cy * sz, -cy * cz, 0],
        [cx * cz - sx * sy * sz, -cx * sz - sx * sy * cz, 0],
        [sx * cz + cx * sy * sz, cx * sy * cz - sx * sz, 0]])
    hang = np.array([                                   # NDT_I:340-359
        [-cx * sz - sx * sy * cz, -cx * cz + sx * sy * sz, sx * cy],      # a2
        [-sx * sz + cx * sy * cz, -cx * sy * sz - sx * cz, -cx * cy],     # a3
        [cx * cy * cz, -cx * cy * sz, cx * sy],                           # b2
        [sx * cy * cz, -sx * cy * sz, sx * sy],                           # b3
        [-sx * cz - cx * sy * sz, sx * sz - cx * sy * cz, 0],             # c2
        [cx * cz - sx * sy * sz, -sx * sy * cz - cx * sz, 0],             # c3
        [-cy * cz, cy * sz, sy],                                          # d1
        [-sx * sy * cz, sx * sy * sz, sx * cy],                           # d2
        [cx * sy * cz, -cx * sy * sz, -cx * cy],                          # d3
        [sy * sz, sy * cz, 0],                                            # e1
        [-sx * cy * sz, -sx * cy * cz, 0],                                # e2
        [cx * cy * sz, cx * cy * cz, 0],                                  # e3
        [-cy * cz, cy * sz, 0],                                           # f1
        [-cx * sz - sx * sy * cz, -cx * cz + sx * sy * sz, 0],            # f2
        [-sx * sz + cx * sy * cz, -cx * sy * sz - sx * cz, 0]])           # f3
    T = pose_matrix(p)
    xt32 = (src[:, :3].astype(np.float32) @ T[:3, :3].astype(np.float32).T + T[:3, 3].astype(np.float32)).astype(np.float32)
    xt = src[:, :3].astype(np.float64) @ T[:3, :3].T + T[:3, 3]
    score, g, H = 0.0, np.zeros(6), np.zeros((6, 6))
    for i in range(src.shape[0]):
        x = src[i, :3].astype(np.float64)
        nb = nbr_sets[i] if nbr_sets is not None else neighbors_direct(vox, xt32[i], leaf, offsets)
        if not nb:
            continue
        ja, ha = jang @ x, hang @ x
        J = np.zeros((3, 6))
        J[:, :3] = np.eye(3)
        J[1, 3], J[2, 3], J[0, 4], J[1, 4], J[2, 4], J[0, 5], J[1, 5], J[2, 5] = ja     # NDT_I:419-426
        a = np.array([0, ha[0], ha[1]]); b = np.array([0, ha[2], ha[3]]); c = np.array([0, ha[4], ha[5]])
        d = ha[6:9]; e = ha[9:12]; f = ha[12:15]
        Hp = {(3, 3): a, (3, 4): b, (3, 5): c, (4, 3): b, (4, 4): d, (4, 5): e, (5, 3): c, (5, 4): e, (5, 5): f}  # NDT_I:435-443
        for k in nb:
            n, mean, cov, C = vox[k]
            xr = xt[i] - mean
            q = xr @ C @ xr
            ee = np.exp(-d2 * q / 2.0)
            score += -d1 * ee
            ee2 = d2 * ee
            if ee2 > 1 or ee2 < 0 or np.isnan(ee2):
                continue
            ee2 *= d1
            xCJ = xr @ C @ J
            g += ee2 * xCJ
            for ii in range(6):
                for jj in range(6):
                    v = -d2 * xCJ[ii] * xCJ[jj] + J[:, jj] @ C @ J[:, ii]
                    if (ii, jj) in Hp:
                        v += xr @ C @ Hp[(ii, jj)]
                    H[ii, jj] += ee2 * v
    return score, g, H
