"""Synthetic inputs for the two-camera (Frame::Nleft != -1) branches of the tracking-time searches, shared by the CPU
and GPU tests.  The two cameras' features are real extractions of a synthetic stereo pair; mvLeftToRightMatch /
mvRightToLeftMatch form a random partial bijection; MapPoints aim at a left feature and at a right feature (often
that left feature's stereo partner, so the cross-bound entries collide with direct matches), with duplicates,
MapPoints without observations, bad ones, and mnTrackScaleLevelR == -1."""
import numpy as np

from oracle import oracle_py as op
from orbslam3_cpp_fork_b200 import synth
from tests.matching_cases import flip_bits

W, H = 480, 320
MP_IN_VIEW, MP_BAD, MP_HAS_OBS, MP_IN_VIEW_R = 1, 2, 4, 8


def rig_views(seed, nf=700):
    L, R = synth.synth_stereo_pair(W, H, seed)
    e1, e2 = op.OracleExtractor(nf, 1.2, 8, 20, 7), op.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, k1, d1 = e1.extract(L)
    _, k2, d2 = e2.extract(R)
    sf = np.asarray(e1.tables()[0], np.float32)
    rng = np.random.default_rng(seed)
    n1, n2 = len(k1), len(k2)
    npair = int(0.45 * min(n1, n2))
    a = rng.permutation(n1)[:npair]; b = rng.permutation(n2)[:npair]
    l2r = np.full(n1, -1, np.int32); r2l = np.full(n2, -1, np.int32)
    l2r[a] = b; r2l[b] = a
    bound_l = (rng.random(n1) < 0.08).astype(np.uint8); bound_r = (rng.random(n2) < 0.08).astype(np.uint8)
    return dict(k1=k1, d1=d1, k2=k2, d2=d2, sf=sf, l2r=l2r, r2l=r2l, bound_l=bound_l, bound_r=bound_r)


def frame_views(rv):
    fvl, ka = op.make_frame_view(rv["k1"], rv["d1"], None, rv["bound_l"], W, H, rv["sf"], mb=0.1, mbf=40.0)
    fvr, kb = op.make_frame_view(rv["k2"], rv["d2"], None, rv["bound_r"], W, H, rv["sf"], mb=0.1, mbf=40.0)
    return fvl, fvr, (ka, kb)


def rig_mappoints(rv, M, seed, no_obs_frac=0.1):
    rng = np.random.default_rng(seed)
    k1, d1, k2, d2, sf = rv["k1"], rv["d1"], rv["k2"], rv["d2"], rv["sf"]
    n1, n2 = len(k1), len(k2)
    tl = rng.integers(0, n1, M)                               # with replacement: several MapPoints per feature
    tr = rng.integers(0, n2, M)
    partner = rv["l2r"][tl]
    use_partner = (partner >= 0) & (rng.random(M) < 0.6)
    tr = np.where(use_partner, partner, tr)
    desc = flip_bits(d1[tl], rng, 40)
    from_r = rng.random(M) < 0.4                              # descriptor closer to the right target
    desc[from_r] = flip_bits(d2[tr[from_r]], rng, 40)
    rnd = rng.random(M) < 0.15
    desc[rnd] = rng.integers(0, 256, (int(rnd.sum()), 32), dtype=np.uint8)
    level = np.clip(k1["octave"][tl] + rng.integers(0, 2, M), 0, 7).astype(np.int32)
    level_r = np.clip(k2["octave"][tr] + rng.integers(0, 2, M), 0, 7).astype(np.int32)
    level_r[rng.random(M) < 0.05] = -1                        # :154
    px = (k1["x"][tl] + rng.uniform(-2, 2, M)).astype(np.float32); py = (k1["y"][tl] + rng.uniform(-2, 2, M)).astype(np.float32)
    pxr = (k2["x"][tr] + rng.uniform(-2, 2, M)).astype(np.float32); pyr = (k2["y"][tr] + rng.uniform(-2, 2, M)).astype(np.float32)
    view_cos = np.where(rng.random(M) < 0.5, 0.9985, 0.9995).astype(np.float32)
    view_cos_r = np.where(rng.random(M) < 0.5, 0.9985, 0.9995).astype(np.float32)
    depth = rng.uniform(1, 80, M).astype(np.float32)
    flags = ((rng.random(M) < 0.8) * MP_IN_VIEW + (rng.random(M) < 0.04) * MP_BAD + (rng.random(M) >= no_obs_frac) * MP_HAS_OBS +
             (rng.random(M) < 0.7) * MP_IN_VIEW_R).astype(np.uint8)
    return dict(desc=desc, px=px, py=py, level=level, view_cos=view_cos, pxr=pxr, pyr=pyr, level_r=level_r,
                view_cos_r=view_cos_r, depth=depth, flags=flags)


def rig_lastframe(rv, seed, trl=(-37.5, 1.25)):
    """Last-frame points: (u, v) lands near a left feature; the right projection is (u + trl_x, v + trl_y)."""
    rng = np.random.default_rng(seed)
    k1, d1, k2, d2 = rv["k1"], rv["d1"], rv["k2"], rv["d2"]
    n1, n2 = len(k1), len(k2)
    n = n1 + 150
    tl = rng.integers(0, n1, n)
    desc = flip_bits(d1[tl], rng, 60)
    # half of the points sit where the right projection lands on a right feature whose descriptor they carry
    tr = rng.integers(0, n2, n)
    on_r = rng.random(n) < 0.5
    u = (k1["x"][tl] + rng.uniform(-5, 5, n)).astype(np.float32); v = (k1["y"][tl] + rng.uniform(-5, 5, n)).astype(np.float32)
    u[on_r] = (k2["x"][tr[on_r]] - np.float32(trl[0]) + rng.uniform(-4, 4, int(on_r.sum()))).astype(np.float32)
    v[on_r] = (k2["y"][tr[on_r]] - np.float32(trl[1]) + rng.uniform(-4, 4, int(on_r.sum()))).astype(np.float32)
    desc[on_r] = flip_bits(d2[tr[on_r]], rng, 60)
    u[:5] = -3.0; v[5:10] = H + 2.0
    z = rng.uniform(0.5, 30.0, n).astype(np.float32); z[10:15] = -2.0
    inv_z = (1.0 / z.astype(np.float64)).astype(np.float32)                        # const float invzc = 1.0 / x3Dc(2)
    ur = (u + np.float32(trl[0])).astype(np.float32); vr = (v + np.float32(trl[1])).astype(np.float32)
    octave = np.clip(np.where(on_r, k2["octave"][tr], k1["octave"][tl]) + rng.integers(-1, 2, n), 0, 7).astype(np.int32)
    angle = np.mod(np.where(on_r, k2["angle"][tr], k1["angle"][tl]) + rng.normal(0, 8, n), 360).astype(np.float32)
    angle[rng.random(n) < 0.1] = rng.uniform(0, 359)
    angle = np.clip(angle, 0, np.float32(359.99)).astype(np.float32)
    flags = ((rng.random(n) < 0.85) * MP_IN_VIEW + (rng.random(n) < 0.8) * MP_HAS_OBS).astype(np.uint8)
    return dict(desc=desc, u=u, v=v, z=z, inv_z=inv_z, ur=ur, vr=vr, octave=octave, angle=angle, flags=flags, trl=trl)


def rig_replay(best_l, best_r, rv, kept_l=None, kept_r=None, cross=True):
    """Per-query results -> what the reference leaves in F.mvpMapPoints (indices [0, Nleft) left, then right):
    query index, -1 none, -2 the entry that was there before.  cross: also bind the stereo partner
    (src/ORBmatcher.cc:138-142, :212-216).  Returns (binding, nmatches, per_query[M, 4] = the features pointing at
    query i right after it was processed, ascending, -1 padded)."""
    n1, n2 = len(rv["k1"]), len(rv["k2"])
    out = np.concatenate([np.where(rv["bound_l"] != 0, -2, -1), np.where(rv["bound_r"] != 0, -2, -1)]).astype(np.int32)
    per = np.full((len(best_l), 4), -1, np.int32)
    n = 0
    for i in range(len(best_l)):
        mine = []
        a = int(best_l[i])
        if a >= 0 and (kept_l is None or kept_l[i]):
            out[a] = i; n += 1; mine.append(a)
            if cross and rv["l2r"][a] != -1:
                out[n1 + rv["l2r"][a]] = i; n += 1; mine.append(n1 + int(rv["l2r"][a]))
        b = int(best_r[i])
        if b >= 0 and (kept_r is None or kept_r[i]):
            if cross and rv["r2l"][b] != -1:
                out[rv["r2l"][b]] = i; n += 1; mine.append(int(rv["r2l"][b]))
            out[n1 + b] = i; n += 1; mine.append(n1 + b)
        mine = sorted(set(mine))
        per[i, :len(mine)] = mine
    return out, n, per


def rig_replay_lastframe(best_l, best_r, rot_l, rot_r, keep, rv, check_ori):
    """SearchByProjection(Cur, Last) on a rig: all writes in query order, then every (query, feature) entry whose
    rotation bin was not kept is NULLed (src/ORBmatcher.cc:1889-1905) whoever holds the feature by then."""
    n1 = len(rv["k1"])
    out = np.concatenate([np.where(rv["bound_l"] != 0, -2, -1), np.where(rv["bound_r"] != 0, -2, -1)]).astype(np.int32)
    n = 0
    for i in range(len(best_l)):
        if best_l[i] >= 0:
            out[best_l[i]] = i; n += 1
        if best_r[i] >= 0:
            out[n1 + best_r[i]] = i; n += 1
    if check_ori:
        keep = set(int(b) for b in keep)
        for i in range(len(best_l)):
            if best_l[i] >= 0 and int(rot_l[i]) not in keep:
                out[best_l[i]] = -1; n -= 1
            if best_r[i] >= 0 and int(rot_r[i]) not in keep:
                out[n1 + best_r[i]] = -1; n -= 1
    return out, n
