"""Synthetic inputs for the matching paths, shared by CPU and GPU tests."""
import numpy as np

from oracle import oracle_py as op
from orbslam3_cpp_fork_b200 import synth


def flip_bits(desc, rng, kmax=40):
    d = desc.copy()
    for row in d:
        k = int(rng.integers(0, kmax + 1))
        pos = rng.choice(256, size=k, replace=False)
        for p in pos:
            row[p >> 3] ^= np.uint8(1 << (p & 7))
    return d


def extracted_frame(w=752, h=480, nf=1000, seed=31):
    e = op.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, k, d = e.extract(synth.synth_image(w, h, seed))
    return e, k, d


def mappoint_case(k, d, scale_factors, M, seed, w=752, h=480, with_uright=False, dup_frac=0.1):
    """cfg4: M MapPoints vs the frame (k, d). ~20% true matches (bit-flipped descriptors, jittered
    projections), some duplicated targets (order-dependent conflicts), the rest distractors."""
    rng = np.random.default_rng(seed)
    N = len(k)
    n_true = int(0.2 * M)
    tgt = rng.integers(0, N, n_true)
    n_dup = int(dup_frac * n_true)
    tgt[:n_dup] = tgt[n_dup:2 * n_dup]                     # several MapPoints chase the same feature
    desc = rng.integers(0, 256, (M, 32), dtype=np.uint8)
    px = rng.uniform(0, w, M).astype(np.float32); py = rng.uniform(0, h, M).astype(np.float32)
    level = rng.integers(0, 8, M).astype(np.int32)
    sel = rng.permutation(M)[:n_true]
    desc[sel] = flip_bits(d[tgt], rng)
    oct_t = k["octave"][tgt]
    level[sel] = np.minimum(oct_t + rng.integers(0, 2, n_true), 7)
    rad = 2.5 * np.asarray(scale_factors)[level[sel]]
    px[sel] = k["x"][tgt] + rng.uniform(-0.5, 0.5, n_true).astype(np.float32) * rad
    py[sel] = k["y"][tgt] + rng.uniform(-0.5, 0.5, n_true).astype(np.float32) * rad
    view_cos = np.where(rng.random(M) < 0.5, 0.9985, 0.9995).astype(np.float32)
    depth = rng.uniform(1, 80, M).astype(np.float32)
    in_view = rng.random(M) < 0.95
    bad = rng.random(M) < 0.02
    has_obs = rng.random(M) < 0.9
    flags = (in_view * 1 + bad * 2 + has_obs * 4).astype(np.uint8)
    u_right = None
    pxr = (px - np.float32(40.0) / depth).astype(np.float32)
    if with_uright:
        u_right = np.where(rng.random(N) < 0.6, k["x"] - rng.uniform(1, 60, N), -1).astype(np.float32)
        ok = u_right[tgt] > 0
        pxr[sel] = np.where(ok, u_right[tgt] + rng.uniform(-3, 3, n_true), pxr[sel]).astype(np.float32)
    bound = (rng.random(N) < 0.1).astype(np.uint8)         # features already bound on entry
    return dict(desc=desc, px=px, py=py, pxr=pxr, level=level, view_cos=view_cos, depth=depth, flags=flags,
                u_right=u_right, bound=bound)


def lastframe_case(k, d, seed, w=752, h=480):
    """SearchByProjection(Cur, Last): last-frame points projected near current features."""
    rng = np.random.default_rng(seed)
    N = len(k)
    n = N + 200
    tgt = rng.integers(0, N, n)
    desc = flip_bits(d[tgt], rng, 60)
    rnd = rng.random(n) < 0.25
    desc[rnd] = rng.integers(0, 256, (int(rnd.sum()), 32), dtype=np.uint8)
    u = (k["x"][tgt] + rng.uniform(-6, 6, n)).astype(np.float32)
    v = (k["y"][tgt] + rng.uniform(-6, 6, n)).astype(np.float32)
    u[:5] = -3.0; v[5:10] = h + 2.0                         # out of image bounds
    inv_z = rng.uniform(0.02, 1.0, n).astype(np.float32); inv_z[10:15] = -0.1
    octave = np.clip(k["octave"][tgt] + rng.integers(-1, 2, n), 0, 7).astype(np.int32)
    angle = np.mod(k["angle"][tgt] + rng.normal(0, 8, n), 360).astype(np.float32)
    angle[rng.random(n) < 0.1] = rng.uniform(0, 360)
    valid = rng.random(n) < 0.8
    has_obs = rng.random(n) < 0.8
    flags = (valid * 1 + has_obs * 4).astype(np.uint8)
    u_right = np.where(rng.random(N) < 0.5, k["x"] - rng.uniform(1, 60, N), -1).astype(np.float32)
    return dict(desc=desc, u=u, v=v, inv_z=inv_z, octave=octave, angle=angle, flags=flags, u_right=u_right)


def frustum_case(n, seed, w=752, h=480):
    """Frame::isInFrustum inputs: a camera pose, n world points scattered in and around the frustum (behind the
    camera, outside the image, too near / too far, grazing normals included), with the Frame members it reads."""
    rng = np.random.default_rng(seed)
    ang = rng.uniform(-0.4, 0.4, 3)
    cx_, sx_ = np.cos(ang[0]), np.sin(ang[0]); cy_, sy_ = np.cos(ang[1]), np.sin(ang[1]); cz_, sz_ = np.cos(ang[2]), np.sin(ang[2])
    Rx = np.array([[1, 0, 0], [0, cx_, -sx_], [0, sx_, cx_]]); Ry = np.array([[cy_, 0, sy_], [0, 1, 0], [-sy_, 0, cy_]])
    Rz = np.array([[cz_, -sz_, 0], [sz_, cz_, 0], [0, 0, 1]])
    R = (Rz @ Ry @ Rx).astype(np.float32)
    t = rng.uniform(-2, 2, 3).astype(np.float32)
    K4 = (458.654, 457.296, 367.215, 248.375)
    cam = op.make_frustum_camera(R, t, K4, 47.906, w, h)
    # points: camera-frame samples pushed back to the world (float64 here: they are just inputs)
    z = rng.uniform(-2, 40, n); z[rng.random(n) < 0.05] = rng.uniform(0.0, 0.05, int((rng.random(n) < 0.05).sum()) or 1)[0]
    u = rng.uniform(-120, w + 120, n); v = rng.uniform(-120, h + 120, n)
    pc = np.stack([(u - K4[2]) / K4[0] * z, (v - K4[3]) / K4[1] * z, z], 1)
    P = ((pc - t.astype(np.float64)) @ R.astype(np.float64)).astype(np.float32)          # R^T (pc - t)
    ow = np.array(cam.ow[:], np.float64)
    po = P.astype(np.float64) - ow
    dist = np.linalg.norm(po, axis=1) + 1e-9
    nrm = po / dist[:, None] + rng.normal(0, 0.6, (n, 3))
    nrm /= np.linalg.norm(nrm, axis=1)[:, None] + 1e-12
    max_d = (dist * rng.uniform(0.6, 3.0, n)).astype(np.float32)
    min_d = (max_d / np.float32(1.2 ** 7) * rng.uniform(0.5, 1.5, n)).astype(np.float32)
    # exact bin edges of PredictScale: ratio = 1.2^k
    k = rng.integers(0, 8, n)
    edge = rng.random(n) < 0.1
    max_d[edge] = (dist[edge] * 1.2 ** k[edge]).astype(np.float32)
    return cam, P, nrm.astype(np.float32), min_d, max_d


def lastframe_pose_case(k, d, seed, motion, w=752, h=480):
    """Tracking::TrackWithMotionModel inputs with the geometry live: current / last camera poses, world points of the
    last frame's MapPoints (back-projected near current features, some behind the camera / out of the image), a mix of
    0- and >0-observation MapPoints and duplicated targets, so that a feature IS bound twice and the later binding
    overwrites the earlier one (src/ORBmatcher.cc:1763-1767, 1786).  motion: 'forward' | 'backward' | 'none'."""
    rng = np.random.default_rng(seed)
    N = len(k)
    K4 = np.array([458.654, 457.296, 367.215, 248.375], np.float32)
    a = rng.uniform(-0.05, 0.05, 3)
    Rx = np.array([[1, 0, 0], [0, np.cos(a[0]), -np.sin(a[0])], [0, np.sin(a[0]), np.cos(a[0])]])
    Ry = np.array([[np.cos(a[1]), 0, np.sin(a[1])], [0, 1, 0], [-np.sin(a[1]), 0, np.cos(a[1])]])
    Rz = np.array([[np.cos(a[2]), -np.sin(a[2]), 0], [np.sin(a[2]), np.cos(a[2]), 0], [0, 0, 1]])
    R = (Rz @ Ry @ Rx).astype(np.float32)
    t = rng.uniform(-0.5, 0.5, 3).astype(np.float32)
    pose_cur = np.concatenate([R.ravel(), t]).astype(np.float32)
    dz = {"forward": -0.6, "backward": 0.6, "none": 0.01}[motion]      # tlc(2) = Tlw * twc: last sees the current centre at +-z
    t_last = (t.astype(np.float64) + np.array([0.0, 0.0, -dz])).astype(np.float32)
    pose_last = np.concatenate([R.ravel(), t_last]).astype(np.float32)
    n = N + 300
    tgt = rng.integers(0, N, n)
    tgt[:150] = tgt[150:300]                                           # several last-frame points chase one feature
    u = k["x"][tgt] + rng.uniform(-5, 5, n); v = k["y"][tgt] + rng.uniform(-5, 5, n)
    u[300:306] = -4.0; v[306:312] = h + 3.0                            # projected out of the image
    z = rng.uniform(1.0, 30.0, n); z[312:320] = -2.0                   # behind the camera
    pc = np.stack([(u - K4[2]) / K4[0] * z, (v - K4[3]) / K4[1] * z, z], 1)
    world = ((pc - t.astype(np.float64)) @ R.astype(np.float64)).astype(np.float32)
    desc = flip_bits(d[tgt], rng, 50)
    rnd_ = rng.random(n) < 0.2
    desc[rnd_] = rng.integers(0, 256, (int(rnd_.sum()), 32), dtype=np.uint8)
    octave = np.clip(k["octave"][tgt] + rng.integers(-1, 2, n), 0, 7).astype(np.int32)
    angle = np.mod(k["angle"][tgt] + rng.normal(0, 8, n), 360).astype(np.float32)
    angle[rng.random(n) < 0.1] = rng.uniform(0, 360)
    has_mp = (rng.random(n) < 0.85).astype(np.uint8)
    outlier = (rng.random(n) < 0.05).astype(np.uint8)
    n_obs = np.where(rng.random(n) < 0.5, 0, rng.integers(1, 6, n)).astype(np.int32)
    u_right = np.where(rng.random(N) < 0.5, k["x"] - rng.uniform(1, 60, N), -1).astype(np.float32)
    bound = (rng.random(N) < 0.05).astype(np.uint8)
    return dict(K4=K4, pose_cur=pose_cur, pose_last=pose_last, world=world, desc=desc, octave=octave, angle=angle,
                has_mp=has_mp, outlier=outlier, n_obs=n_obs, u_right=u_right, bound=bound)


def localmap_case(k, d, scale_factors, M, seed, w=752, h=480):
    """Tracking::SearchLocalPoints inputs with the geometry live: a camera pose and M world points; ~25 % are
    back-projections of frame features (bit-flipped descriptors, distance range chosen so that PredictScale lands on
    or next to the feature's octave, normals near the viewing ray), several points per feature, the rest are
    distractors all around the frustum; some points are bad, some have no observations, some are already matched."""
    rng = np.random.default_rng(seed)
    N = len(k)
    a = rng.uniform(-0.3, 0.3, 3)
    Rx = np.array([[1, 0, 0], [0, np.cos(a[0]), -np.sin(a[0])], [0, np.sin(a[0]), np.cos(a[0])]])
    Ry = np.array([[np.cos(a[1]), 0, np.sin(a[1])], [0, 1, 0], [-np.sin(a[1]), 0, np.cos(a[1])]])
    Rz = np.array([[np.cos(a[2]), -np.sin(a[2]), 0], [np.sin(a[2]), np.cos(a[2]), 0], [0, 0, 1]])
    R = (Rz @ Ry @ Rx).astype(np.float32)
    t = rng.uniform(-2, 2, 3).astype(np.float32)
    K4 = (458.654, 457.296, 367.215, 248.375)
    mbf = 47.906
    cam = op.make_frustum_camera(R, t, K4, mbf, w, h)
    n_true = int(0.25 * M)
    tgt = rng.integers(0, N, n_true)
    tgt[:n_true // 10] = tgt[n_true // 10:2 * (n_true // 10)]          # several MapPoints chase the same feature
    sel = rng.permutation(M)[:n_true]
    u = rng.uniform(-100, w + 100, M); v = rng.uniform(-100, h + 100, M)
    z = rng.uniform(-2, 40, M)
    u[sel] = k["x"][tgt] + rng.uniform(-1.5, 1.5, n_true); v[sel] = k["y"][tgt] + rng.uniform(-1.5, 1.5, n_true)
    z[sel] = rng.uniform(1.0, 30.0, n_true)
    pc = np.stack([(u - K4[2]) / K4[0] * z, (v - K4[3]) / K4[1] * z, z], 1)
    P = ((pc - t.astype(np.float64)) @ R.astype(np.float64)).astype(np.float32)
    ow = np.array(cam.ow[:], np.float64)
    po = P.astype(np.float64) - ow
    dist = np.linalg.norm(po, axis=1) + 1e-9
    nrm = po / dist[:, None] + rng.normal(0, 0.5, (M, 3))
    nrm[sel] = po[sel] / dist[sel, None] + rng.normal(0, 0.15, (n_true, 3))
    nrm /= np.linalg.norm(nrm, axis=1)[:, None] + 1e-12
    max_d = (dist * rng.uniform(0.6, 3.0, M)).astype(np.float32)
    lvl = np.minimum(k["octave"][tgt] + rng.integers(0, 2, n_true), 7)
    max_d[sel] = (dist[sel] * 1.2 ** (lvl - 0.5)).astype(np.float32)    # PredictScale -> lvl
    min_d = (max_d / np.float32(1.2 ** 7) * rng.uniform(0.3, 1.0, M)).astype(np.float32)
    desc = rng.integers(0, 256, (M, 32), dtype=np.uint8)
    desc[sel] = flip_bits(d[tgt], rng)
    bad = (rng.random(M) < 0.03).astype(np.uint8)
    has_obs = (rng.random(M) < 0.92).astype(np.uint8)
    skip_idx = rng.permutation(M)[:M // 40].astype(np.int32)            # mnLastFrameSeen == this frame
    u_right = np.where(rng.random(N) < 0.6, k["x"] - rng.uniform(1, 60, N), -1).astype(np.float32)
    ok = u_right[tgt] > 0
    # depth consistent with the feature's right coordinate for most stereo targets: z = mbf / disparity
    zz = np.where(ok & (rng.random(n_true) < 0.8), mbf / np.maximum(k["x"][tgt] - u_right[tgt], 1e-3), z[sel])
    pc[sel] = np.stack([(u[sel] - K4[2]) / K4[0] * zz, (v[sel] - K4[3]) / K4[1] * zz, zz], 1)
    P[sel] = ((pc[sel] - t.astype(np.float64)) @ R.astype(np.float64)).astype(np.float32)
    po = P.astype(np.float64) - ow
    dist2 = np.linalg.norm(po, axis=1) + 1e-9
    max_d[sel] = (dist2[sel] * 1.2 ** (lvl - 0.5)).astype(np.float32)
    min_d[sel] = (max_d[sel] / np.float32(1.2 ** 7) * 0.5).astype(np.float32)
    nrm[sel] = po[sel] / dist2[sel, None] + rng.normal(0, 0.15, (n_true, 3))
    nrm /= np.linalg.norm(nrm, axis=1)[:, None] + 1e-12
    bound = (rng.random(N) < 0.1).astype(np.uint8)
    return dict(cam=cam, P=P, N=nrm.astype(np.float32), min_d=min_d, max_d=max_d, desc=desc, bad=bad, has_obs=has_obs,
                skip_idx=skip_idx, u_right=u_right, bound=bound, mbf=mbf)


def localmap_expected(c, k, d, sf, frustum_fn, search_fn, th, far, th_far, ratio, limit=0.5, w=752, h=480):
    """The same computation from its two parts: isInFrustum for the points that are neither skipped nor bad, then
    SearchByProjection over all points with the resulting tracking fields.  frustum_fn / search_fn: the oracle's or the
    reference's (oracle_py / ref_py)."""
    M = len(c["P"])
    f = frustum_fn(c["cam"], c["P"], c["N"], c["min_d"], c["max_d"], limit)
    skip = np.zeros(M, bool); skip[c["skip_idx"]] = True
    in_view = (f["in_view"] > 0) & ~skip & (c["bad"] == 0)
    flags = (in_view * 1 + c["bad"] * 2 + c["has_obs"] * 4).astype(np.uint8)
    fv, keep = op.make_frame_view(k, d, c["u_right"], c["bound"], w, h, sf, mb=0.1, mbf=c["mbf"])
    res = search_fn(fv, c["desc"], f["proj_x"], f["proj_y"], f["proj_xr"], np.maximum(f["level"], 0), f["view_cos"], f["depth"],
                    flags, th, far, th_far, ratio)
    return int(in_view.sum()), res[0], res[1]
