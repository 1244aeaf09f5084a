"""How the flat inputs of the N3 searches (tests/search_cases.py, tests/bow_cases.py) relate to what the reference's
own functions are given through oracle/ref_compat/ref_capi_search.cpp.  Shared by the CPU (oracle vs reference) and
GPU (libcuorb vs reference) tests."""
import numpy as np

from oracle import oracle_py as op
from tests.search_cases import H, W

f32 = np.float32


def in_image(u, v):
    """KeyFrame::IsInImage (src/KeyFrame.cc:836-838) with the KeyFrame's int bounds [0, W) x [0, H)."""
    u = np.asarray(u, np.float32); v = np.asarray(v, np.float32)
    return ((u >= 0) & (u < W) & (v >= 0) & (v < H)).astype(np.uint8)


def binding_from_best(best, kept, bound):
    """Per-query result (best feature, -1 none) -> per-feature result as the reference leaves it in
    mvpMapPoints / vpMatched: query index, -1 none, -2 the entry that was there before the call."""
    out = np.where(np.asarray(bound) != 0, -2, -1).astype(np.int32)
    kept = np.broadcast_to(kept, best.shape)
    for i in np.nonzero((best >= 0) & kept)[0]:
        assert out[best[i]] == -1, "a feature was bound twice"
        out[best[i]] = i
    return out


def has_mp_codes(valid, seed):
    """valid (0/1) -> 0 no MapPoint, 1 a good MapPoint, 2 a bad MapPoint (isBad()): a third of the 0s become 2s."""
    rng = np.random.default_rng(1000 + seed)
    v = np.asarray(valid, np.uint8).copy()
    bad = (v == 0) & (rng.random(len(v)) < 0.33)
    v[bad] = 2
    return v


def bow_views(c, sf=None):
    """Frame views for a tests/bow_cases.py pair: arbitrary in-image positions, octave 0, the case's angles."""
    sf = np.asarray([1.2 ** i for i in range(8)], np.float32) if sf is None else sf
    out = []
    keep = []
    for desc, angle, seed in ((c["desc1"], c["angle1"], 1), (c["desc2"], c["angle2"], 2)):
        rng = np.random.default_rng(seed)
        k = np.zeros(len(desc), op.KP_DTYPE)
        k["x"] = rng.uniform(1, W - 1, len(desc)).astype(np.float32); k["y"] = rng.uniform(1, H - 1, len(desc)).astype(np.float32)
        k["size"] = 31; k["angle"] = angle; k["octave"] = 0; k["class_id"] = -1
        fv, ka = op.make_frame_view(k, desc, None, None, W, H, sf)
        out.append(fv); keep.append(ka)
    return out[0], out[1], keep


def fuse_depths(q, mbf, seed):
    """Camera-frame depth per query and the right-image coordinate the reference derives from it
    (src/ORBmatcher.cc:1198, :1208: invz = 1 / z; ur = uv(0) - bf * invz, all in float)."""
    rng = np.random.default_rng(2000 + seed)
    z = rng.uniform(0.5, 6.0, len(q["u"])).astype(np.float32)
    # aim most queries' ur at the target feature's uRight so the 3-dof chi-square gate lets some through
    want = q["u"] - q["ur"]
    ok = want > 1
    z = np.where(ok, (f32(mbf) / np.where(ok, want, 1)).astype(np.float32), z).astype(np.float32)
    invz = (f32(1.0) / z).astype(np.float32)
    ur = (q["u"] - (f32(mbf) * invz).astype(np.float32)).astype(np.float32)
    return z, ur


def triangulation_poses(kind):
    """T1w (R row-major, t) for KF1; KF2 sits at the identity, K = I."""
    if kind == "sideways":
        # near-rectified pair: t12 ~ (1, 0, 0) so epipolar lines are image rows, plus a small rotation and
        # off-axis translation so every entry of F12 takes part (K = I: the entries multiply pixel coordinates)
        ax, ay, az = 2e-5, -1.5e-5, 2.5e-5
        R = np.array([[1, -az, ay], [az, 1, -ax], [-ay, ax, 1]], np.float64)
        t = np.array([1.0, 1e-4, 2e-4])
    else:
        # KF1's centre projects to (W/2 + 30, H/2 - 12) in KF2: the epipole-distance gate (:1029-1040) bites
        R = np.eye(3)
        cw = np.array([W * 0.5 + 30.0, H * 0.5 - 12.0, 1.0])
        t = -R @ cw
    return np.concatenate([R.reshape(9), t]).astype(np.float32)


def fuse_bookkeeping(best, bound, nobs_kf, nobs_q, sim3):
    """What ORBmatcher::Fuse does once a query has chosen KeyFrame feature best[i] (src/ORBmatcher.cc:1308-1322 and
    :1437-1447), replayed in query order.  Returns per query: 1 = the query was replaced by the point the KeyFrame
    holds (Sim3 overload: recorded in vpReplacePoint), 2 = the KeyFrame's original point was replaced by the query,
    3 = the query was added to the KeyFrame, 0 = nothing (no feature, or the KeyFrame's point was already bad)."""
    holder = {int(i): ("kf", int(i)) for i in np.nonzero(np.asarray(bound))[0]}     # feature -> point held
    nobs = {("kf", int(i)): int(nobs_kf[i]) for i in range(len(bound))}
    nobs.update({("q", i): int(nobs_q[i]) for i in range(len(best))})
    bad, replaced_by, in_kf, rep = set(), {}, set(), {}
    for i in range(len(best)):
        idx = int(best[i])
        if idx < 0:
            continue
        me = ("q", i)
        p = holder.get(idx)
        if p is not None:
            if p not in bad:
                if sim3:
                    rep[i] = p
                elif nobs[p] > nobs[me]:
                    replaced_by[me] = p; bad.add(me)
                else:
                    replaced_by[p] = me; bad.add(p)
        else:
            in_kf.add(me); nobs[me] += 1; holder[idx] = me
    out = np.zeros(len(best), np.int32)
    for i in range(len(best)):
        me = ("q", i)
        if (i in rep) if sim3 else (me in replaced_by):
            out[i] = 1
        elif me in in_kf:
            out[i] = 3
        elif best[i] >= 0 and replaced_by.get(("kf", int(best[i]))) == me:
            out[i] = 2
    return out


def bow_rig_views(c, nleft1, nleft2, sf=None):
    """Four frame views for a tests/bow_cases.py pair read as a rig KeyFrame (features [0, nleft1) left camera, the
    rest right camera) and a rig Frame (split at nleft2): arbitrary in-image positions, octave 0, the case's angles."""
    sf = np.asarray([1.2 ** i for i in range(8)], np.float32) if sf is None else sf
    out, keep = [], []
    for desc, angle, seed, cut in ((c["desc1"], c["angle1"], 1, nleft1), (c["desc2"], c["angle2"], 2, nleft2)):
        rng = np.random.default_rng(seed)
        k = np.zeros(len(desc), op.KP_DTYPE)
        k["x"] = rng.uniform(1, W - 1, len(desc)).astype(np.float32); k["y"] = rng.uniform(1, H - 1, len(desc)).astype(np.float32)
        k["size"] = 31; k["angle"] = angle; k["octave"] = 0; k["class_id"] = -1
        for lo, hi in ((0, cut), (cut, len(desc))):
            fv, ka = op.make_frame_view(k[lo:hi], desc[lo:hi], None, None, W, H, sf)
            out.append(fv); keep.append(ka)
    return out, keep
