"""CPU: the oracle's restatement of the remaining ORBmatcher searches (oracle/search_oracle.cpp) against
an independent brute-force restatement in numpy/Python (no grid: candidates are found by scanning all
features and ordered by (cell x, cell y, index), the order the reference's grid walk visits them), plus
invariants the reference loops guarantee."""
import numpy as np
import pytest

from oracle import oracle_py as op
from tests.search_cases import H, W, point_queries, sim3_case, triangulation_case, two_views
from orbslam3_cpp_fork_b200.matcher import join_feature_vectors

f32 = np.float32
GW, GH = f32(64.0) / f32(W), f32(48.0) / f32(H)


def ham(a, b):
    return int(np.unpackbits(a ^ b).sum())


def cells(k):
    cx = np.rint((k["x"] - f32(0)) * GW).astype(int); cy = np.rint((k["y"] - f32(0)) * GH).astype(int)
    # std::round is half-away-from-zero; coordinates are positive and never land on .5 exactly here
    ok = (cx >= 0) & (cx < 64) & (cy >= 0) & (cy < 48)
    return cx, cy, ok


def area(k, cx, cy, ok, x, y, r):
    """GetFeaturesInArea without level filter, in visiting order."""
    x, y, r = f32(x), f32(y), f32(r)
    x0 = max(0, int(np.floor((x - f32(0) - r) * GW))); x1 = min(63, int(np.ceil((x - f32(0) + r) * GW)))
    y0 = max(0, int(np.floor((y - f32(0) - r) * GH))); y1 = min(47, int(np.ceil((y - f32(0) + r) * GH)))
    if x0 >= 64 or x1 < 0 or y0 >= 48 or y1 < 0:
        return []
    m = ok & (cx >= x0) & (cx <= x1) & (cy >= y0) & (cy <= y1) & (np.abs(k["x"] - x) < r) & (np.abs(k["y"] - y) < r)
    idx = np.nonzero(m)[0]
    return sorted(idx.tolist(), key=lambda i: (cx[i], cy[i], i))


def py_window_search(k, d, sf, q, th, lo, hi, max_dist, bound=None, bounds_check=False, chi2=None, u_right=None):
    cx, cy, ok = cells(k)
    taken = None if bound is None else bound.astype(bool).copy()
    out = np.full(len(q["u"]), -1, np.int32)
    for i in range(len(q["u"])):
        if not q["active"][i]:
            continue
        x, y, lvl = q["u"][i], q["v"][i], int(q["level"][i])
        if bounds_check and (x < 0 or x > W or y < 0 or y > H):
            continue
        best, bi = 256, -1
        for j in area(k, cx, cy, ok, x, y, f32(th) * sf[lvl]):
            if taken is not None and taken[j]:
                continue
            o = int(k["octave"][j])
            if o < lvl + lo or o > lvl + hi:
                continue
            if chi2 is not None:
                ex, ey = f32(x) - k["x"][j], f32(y) - k["y"][j]
                e2 = f32(f32(ex * ex) + f32(ey * ey))
                if u_right is not None and u_right[j] >= 0:
                    er = f32(q["ur"][i]) - u_right[j]
                    e2 = f32(e2 + f32(er * er))
                    if float(f32(e2 * chi2[o])) > 7.8:
                        continue
                elif float(f32(e2 * chi2[o])) > 5.99:
                    continue
            dist = ham(q["desc"][i], d[j])
            if dist < best:
                best, bi = dist, j
        if bi >= 0 and f32(best) <= f32(max_dist):
            out[i] = bi
            if taken is not None:
                taken[bi] = True
    return out


@pytest.fixture(scope="module")
def tv():
    return two_views(5, nf=800)


def test_reloc_sim3_fuse_vs_bruteforce(tv):
    k, d, sf = tv["k1"], tv["d1"], tv["sf"]
    q = point_queries(k, d, sf, 500, 3)
    # relocalisation: Frame window with levels [l-1, l+1], any bound feature skipped, bounds test
    fv, keep = op.make_frame_view(k, d, None, q["bound"], W, H, sf)
    n, best, rot, keepb = op.search_by_projection_reloc(fv, q["desc"], q["u"], q["v"], q["level"], q["angle"], q["active"],
                                                        10.0, 80, True)
    ref = py_window_search(k, d, sf, q, 10.0, -1, 1, 80, bound=q["bound"], bounds_check=True)
    assert (best == ref).all() and (best >= 0).sum() > 100
    hit = best[best >= 0]
    assert len(set(hit.tolist())) == len(hit) and (q["bound"][hit] == 0).all()      # one MapPoint per feature
    assert n == sum(1 for i in range(len(best)) if best[i] >= 0 and rot[i] in keepb)
    assert (best[:8] == -1).all()
    # Sim3 projection: KeyFrame window, levels [l-1, l], threshold TH_LOW * ratioHamming
    for th, ratio in ((10, 1.0), (6, 1.5)):
        n, best = op.search_by_projection_sim3(fv, q["desc"], q["u"], q["v"], q["level"], q["active"], th, ratio)
        ref = py_window_search(k, d, sf, q, th, -1, 0, f32(50) * f32(ratio), bound=q["bound"])
        assert (best == ref).all() and n == (best >= 0).sum() and n > 50
    # Fuse: no binding, chi-square gates
    fvs, keep2 = op.make_frame_view(k, d, q["u_right"], None, W, H, sf, mb=0.1, mbf=40.0)
    n1, b1 = op.fuse(fvs, tv["inv_sigma2"], q["desc"], q["u"], q["v"], q["ur"], q["level"], q["active"], 3.0, True)
    ref = py_window_search(k, d, sf, q, 3.0, -1, 0, 50, chi2=tv["inv_sigma2"], u_right=q["u_right"])
    assert (b1 == ref).all() and n1 == (b1 >= 0).sum()
    n0, b0 = op.fuse(fvs, None, q["desc"], q["u"], q["v"], None, q["level"], q["active"], 3.0, False)
    ref0 = py_window_search(k, d, sf, q, 3.0, -1, 0, 50)
    assert (b0 == ref0).all() and n0 >= n1 > 30 and (b0 != b1).any()               # the gate rejects something


def test_search_by_sim3_agreement(tv):
    a, b = sim3_case(tv, 11)
    fv1, k1 = op.make_frame_view(tv["k1"], tv["d1"], None, None, W, H, tv["sf"])
    fv2, k2 = op.make_frame_view(tv["k2"], tv["d2"], None, None, W, H, tv["sf"])
    n, m12 = op.search_by_sim3(fv1, fv2, *a, *b, 7.5)
    qa = dict(desc=a[0], u=a[1], v=a[2], level=a[3], active=a[4]); qb = dict(desc=b[0], u=b[1], v=b[2], level=b[3], active=b[4])
    r1 = py_window_search(tv["k2"], tv["d2"], tv["sf"], qa, 7.5, -1, 0, 100)
    r2 = py_window_search(tv["k1"], tv["d1"], tv["sf"], qb, 7.5, -1, 0, 100)
    ref = np.array([r1[i] if r1[i] >= 0 and r2[r1[i]] == i else -1 for i in range(len(r1))], np.int32)
    assert (m12 == ref).all() and n == (m12 >= 0).sum() and n > 40


def py_init(k1, d1, k2, d2, prev, window, ratio):
    cx, cy, ok = cells(k2)
    m12 = np.full(len(k1), -1, np.int32); m21 = np.full(len(k2), -1, np.int32)
    md = np.full(len(k2), 2 ** 31 - 1, np.int64)
    events = []
    for i1 in range(len(k1)):
        if k1["octave"][i1] > 0:
            continue
        best = best2 = 2 ** 31 - 1; bi = -1
        for j in area(k2, cx, cy, ok, prev[i1, 0], prev[i1, 1], f32(window)):
            if k2["octave"][j] != 0:
                continue
            dist = ham(d1[i1], d2[j])
            if md[j] <= dist:
                continue
            if dist < best:
                best2, best, bi = best, dist, j
            elif dist < best2:
                best2 = dist
        if best <= 50 and f32(best) < f32(f32(best2) * f32(ratio)):
            if m21[bi] >= 0:
                m12[m21[bi]] = -1
            m12[i1] = bi; m21[bi] = i1; md[bi] = best
            events.append((i1, bi))
    return m12, events


@pytest.mark.parametrize("tie_heavy", [False, True])
def test_search_for_initialization_vs_bruteforce(tie_heavy):
    t = two_views(6, nf=400, tie_heavy=tie_heavy)
    k1, d1, k2, d2, sf = t["k1"], t["d1"], t["k2"], t["d2"], t["sf"]
    fv1, a = op.make_frame_view(k1, d1, None, None, W, H, sf); fv2, b = op.make_frame_view(k2, d2, None, None, W, H, sf)
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    n, m12, pm = op.search_for_initialization(fv1, fv2, prev, 60, 0.9, False)
    ref, events = py_init(k1, d1, k2, d2, prev, 60, 0.9)
    assert (m12 == ref).all() and n == (m12 >= 0).sum() and n > (3 if tie_heavy else 10)
    if tie_heavy:
        assert len(events) > n                                   # some matches were stolen by later features
    hit = m12 >= 0
    assert (pm[hit, 0] == k2["x"][m12[hit]]).all() and (pm[~hit] == prev[~hit]).all()
    # orientation check only removes matches, never adds or moves them
    n2, m12o, _ = op.search_for_initialization(fv1, fv2, prev, 60, 0.9, True)
    assert n2 <= n and ((m12o == m12) | (m12o == -1)).all()


def py_triangulate(t, c, only_stereo, coarse):
    k1, d1, k2, d2, sf = t["k1"], t["d1"], t["k2"], t["d2"], t["sf"]
    off1, idx1, off2, idx2 = join_feature_vectors(c["fv1"], c["fv2"])
    F = c["F12"].reshape(9)
    out = np.full(len(k1), -1, np.int32)
    for nd in range(len(off1) - 1):
        for i1 in idx1[off1[nd]:off1[nd + 1]]:
            if c["has1"][i1]:
                continue
            s1 = c["ur1"][i1] >= 0
            if only_stereo and not s1:
                continue
            x1, y1 = k1["x"][i1], k1["y"][i1]
            a = f32(f32(f32(x1 * F[0]) + f32(y1 * F[3])) + F[6]); b = f32(f32(f32(x1 * F[1]) + f32(y1 * F[4])) + F[7])
            cc = f32(f32(f32(x1 * F[2]) + f32(y1 * F[5])) + F[8])
            best, bi = 50, -1
            for i2 in idx2[off2[nd]:off2[nd + 1]]:
                if c["has2"][i2]:
                    continue
                s2 = c["ur2"][i2] >= 0
                if only_stereo and not s2:
                    continue
                dist = ham(d1[i1], d2[i2])
                if dist > 50 or dist > best:
                    continue
                x2, y2 = k2["x"][i2], k2["y"][i2]
                if not s1 and not s2:
                    ex, ey = c["ep"][0] - x2, c["ep"][1] - y2
                    if f32(f32(ex * ex) + f32(ey * ey)) < f32(f32(100) * sf[k2["octave"][i2]]):
                        continue
                if not coarse:
                    num = f32(f32(f32(a * x2) + f32(b * y2)) + cc); den = f32(f32(a * a) + f32(b * b))
                    if den == 0 or not (float(f32(f32(num * num) / den)) < 3.84 * float(t["sigma2"][k2["octave"][i2]])):
                        continue
                best, bi = dist, i2
            out[i1] = bi
    return out


@pytest.mark.parametrize("only_stereo,coarse", [(False, False), (True, False), (False, True)])
def test_search_for_triangulation_vs_bruteforce(tv, only_stereo, coarse):
    c = triangulation_case(tv, 4)
    fv1, a = op.make_frame_view(tv["k1"], tv["d1"], c["ur1"], c["has1"], W, H, tv["sf"])
    fv2, b = op.make_frame_view(tv["k2"], tv["d2"], c["ur2"], c["has2"], W, H, tv["sf"])
    off1, idx1, off2, idx2 = join_feature_vectors(c["fv1"], c["fv2"])
    n, m12 = op.search_for_triangulation(fv1, fv2, off1, idx1, off2, idx2, c["F12"], c["ep"], tv["sigma2"], only_stereo, coarse, False)
    ref = py_triangulate(tv, c, only_stereo, coarse)
    assert (m12 == ref).all() and n == (m12 >= 0).sum() and n > (0 if only_stereo else 8)
    hit = m12 >= 0
    assert (c["has1"][hit] == 0).all() and (c["has2"][m12[hit]] == 0).all()
    if only_stereo:
        assert (c["ur1"][hit] >= 0).all() and (c["ur2"][m12[hit]] >= 0).all()
    n2, m12o = op.search_for_triangulation(fv1, fv2, off1, idx1, off2, idx2, c["F12"], c["ep"], tv["sigma2"], only_stereo, coarse, True)
    assert n2 <= n and ((m12o == m12) | (m12o == -1)).all()
