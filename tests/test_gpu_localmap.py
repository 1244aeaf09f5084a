"""GPU (-m gpu): the resident local map (cuorb_localmap_*, include/cuorb.h) - Tracking::SearchLocalPoints' isInFrustum
loop and ORBmatcher::SearchByProjection in one call on device-resident MapPoints - against the same computation made of
its two parts by the oracle and by the reference's own functions (oracle/_ref: Frame::isInFrustum lifted verbatim, then
the unmodified ORBmatcher::SearchByProjection).  Bit-exact: number of points in view, matches per point, pair list."""
import numpy as np
import pytest

from oracle import oracle_py as op
from oracle import ref_py as rp
from tests.matching_cases import extracted_frame, localmap_case, localmap_expected

pytestmark = pytest.mark.gpu
W, H = 752, 480


@pytest.fixture(scope="module")
def fr():
    e, k, d = extracted_frame(seed=33)
    return k, d, np.asarray(e.tables()[0], np.float32)


def _cam(c):
    from orbslam3_cpp_fork_b200 import _lib
    return _lib.FrustumCamera.from_buffer_copy(bytes(c["cam"]))


def _frame(k, d, sf, c):
    from orbslam3_cpp_fork_b200.matcher import Frame
    F = Frame(k, d, sf, W, H, mvuRight=c["u_right"], mb=0.1, mbf=c["mbf"])
    F.mvpMapPoints[c["bound"] > 0] = 10 ** 6; F.mvpHasObs[c["bound"] > 0] = 1
    return F


def _fill(m, c, lo=0, hi=None):
    hi = len(c["P"]) if hi is None else hi
    m.update(lo, c["P"][lo:hi], c["N"][lo:hi], c["min_d"][lo:hi], c["max_d"][lo:hi], c["desc"][lo:hi], c["bad"][lo:hi],
             c["has_obs"][lo:hi])


@pytest.mark.parametrize("M,seed,th,far", [(50000, 1, 1.0, False), (8000, 2, 3.0, True), (20000, 3, 1.0, True)])
def test_localmap_search_equals_its_two_parts(fr, M, seed, th, far):
    from orbslam3_cpp_fork_b200 import LocalMap
    k, d, sf = fr
    c = localmap_case(k, d, sf, M, seed)
    m = LocalMap(M + 100)
    _fill(m, c, 0, M // 3); _fill(m, c, M // 3, M)                 # written in two pieces
    assert m.count == M
    F = _frame(k, d, sf, c)
    n, n_in, mp, ft, best = m.search(_cam(c), F, c["skip_idx"], 0.5, th, far, 30.0, 0.8)
    o_in, o_n, o_best = localmap_expected(c, k, d, sf, op.is_in_frustum, op.search_by_projection, th, far, 30.0, 0.8)
    assert n_in == o_in and n == o_n and (best == o_best).all() and n > 500
    hit = np.nonzero(best >= 0)[0]
    assert (mp == hit).all() and (ft == best[hit]).all()           # pair list: point order
    assert not np.isin(c["skip_idx"], mp).any() and (c["bad"][mp] == 0).all()
    if rp.available():
        r_in, r_n, r_best = localmap_expected(c, k, d, sf, rp.is_in_frustum, rp.search_by_projection, th, far, 30.0, 0.8)
        assert n_in == r_in and n == r_n and (best == r_best).all()


def test_localmap_incremental_update_and_next_frame(fr):
    """The map persists between frames: a few points change (turn bad, get a new descriptor, move), the pose changes."""
    from orbslam3_cpp_fork_b200 import LocalMap
    k, d, sf = fr
    M = 12000
    c = localmap_case(k, d, sf, M, 5)
    m = LocalMap(M)
    _fill(m, c)
    F = _frame(k, d, sf, c)
    n0, in0, mp0, ft0, best0 = m.search(_cam(c), F, c["skip_idx"], 0.5, 1.0, False, 30.0, 0.8)
    # frame 2: points 1000..1499 are replaced, another pose's case supplies the camera
    c2 = localmap_case(k, d, sf, M, 6)
    for key in ("P", "N", "min_d", "max_d", "desc", "bad", "has_obs"):
        c[key][1000:1500] = c2[key][1000:1500]
    _fill(m, c, 1000, 1500)
    n1, in1, mp1, ft1, best1 = m.search(_cam(c), _frame(k, d, sf, c), [], 0.5, 1.0, False, 30.0, 0.8, want_best=True)
    c_noskip = dict(c); c_noskip["skip_idx"] = np.zeros(0, np.int32)
    o_in, o_n, o_best = localmap_expected(c_noskip, k, d, sf, op.is_in_frustum, op.search_by_projection, 1.0, False, 30.0, 0.8)
    assert in1 == o_in and n1 == o_n and (best1 == o_best).all()
    # same map, the other case's camera: nothing is uploaded but the pose and the frame
    c_pose = dict(c_noskip); c_pose["cam"] = c2["cam"]
    n2, in2, mp2, ft2, best2 = m.search(_cam(c_pose), _frame(k, d, sf, c), [], 0.5, 1.0, False, 30.0, 0.8)
    o_in, o_n, o_best = localmap_expected(c_pose, k, d, sf, op.is_in_frustum, op.search_by_projection, 1.0, False, 30.0, 0.8)
    assert in2 == o_in and n2 == o_n and (best2 == o_best).all()
    assert n0 > 300 and n1 > 300


def test_localmap_limits(fr):
    from orbslam3_cpp_fork_b200 import LocalMap, _lib
    k, d, sf = fr
    c = localmap_case(k, d, sf, 4000, 7)
    m = LocalMap(4000)
    F = _frame(k, d, sf, c)
    n, n_in, mp, ft, best = m.search(_cam(c), F)                   # empty map
    assert n == 0 and n_in == 0 and len(mp) == 0
    _fill(m, c)
    with pytest.raises(_lib.CuorbError):                           # more matches than the caller's pair arrays hold
        m.search(_cam(c), _frame(k, d, sf, c), [], 0.5, 1.0, False, 30.0, 0.8, match_capacity=5)
    assert "match_capacity" in _lib.last_error()
    with pytest.raises(_lib.CuorbError):
        m.update(3990, c["P"][:20], c["N"][:20], c["min_d"][:20], c["max_d"][:20], c["desc"][:20], c["bad"][:20], c["has_obs"][:20])
    n, n_in, mp, ft, best = m.search(_cam(c), _frame(k, d, sf, c), np.arange(4000, dtype=np.int32))   # everything skipped
    assert n == 0 and n_in == 0
    n, n_in, mp, ft, best = m.search(_cam(c), _frame(k, d, sf, c), [], 0.5, 1.0, False, 30.0, 0.8, want_best=False)
    assert best is None and n > 100 and len(mp) == n


def test_localmap_processing_order_and_in_view(fr):
    """Storage stays put while Tracking rebuilds mvpLocalMapPoints in another order every frame: `order` lists the
    storage slots in this frame's processing order (which decides contested features); all results are positions."""
    from orbslam3_cpp_fork_b200 import LocalMap
    k, d, sf = fr
    M = 15000
    c = localmap_case(k, d, sf, M, 8)
    m = LocalMap(M + 7)
    _fill(m, c)
    rng = np.random.default_rng(3)
    order = rng.permutation(M)[:M - 500].astype(np.int32)          # a permuted subset: 500 points left the local map
    skip_pos = rng.permutation(len(order))[:300].astype(np.int32)
    n, n_in, mp, ft, best, inv = m.search(_cam(c), _frame(k, d, sf, c), skip_pos, 0.5, 1.0, False, 30.0, 0.8, order=order,
                                          want_in_view=True)
    cp = {key: (c[key][order] if key in ("P", "N", "min_d", "max_d", "desc", "bad", "has_obs") else c[key]) for key in c}
    cp["skip_idx"] = skip_pos
    o_in, o_n, o_best = localmap_expected(cp, k, d, sf, op.is_in_frustum, op.search_by_projection, 1.0, False, 30.0, 0.8)
    assert n_in == o_in and n == o_n and (best == o_best).all() and n > 300
    assert inv.sum() == n_in and (inv[best >= 0] == 1).all() and (inv[skip_pos] == 0).all()
    # the order matters: the identity order over the same subset gives another assignment of contested features
    srt = np.sort(order)
    n2, _, _, _, best2 = m.search(_cam(c), _frame(k, d, sf, c), [], 0.5, 1.0, False, 30.0, 0.8, order=srt)
    n3, _, _, _, best3 = m.search(_cam(c), _frame(k, d, sf, c), [], 0.5, 1.0, False, 30.0, 0.8, order=order)
    by_slot = np.full(M, -1, np.int32); by_slot[order] = best3
    by_slot2 = np.full(M, -1, np.int32); by_slot2[srt] = best2
    assert (by_slot != by_slot2).any()
    if rp.available():
        r_in, r_n, r_best = localmap_expected(cp, k, d, sf, rp.is_in_frustum, rp.search_by_projection, 1.0, False, 30.0, 0.8)
        assert n_in == r_in and n == r_n and (best == r_best).all()


def test_localmap_empty_frame_and_inactive_slots(fr):
    from orbslam3_cpp_fork_b200 import LocalMap
    from orbslam3_cpp_fork_b200.matcher import Frame
    k, d, sf = fr
    c = localmap_case(k, d, sf, 3000, 9)
    m = LocalMap(5000)
    _fill(m, c)
    empty = Frame(k[:0], d[:0], sf, W, H, mb=0.1, mbf=c["mbf"])
    n, n_in, mp, ft, best, inv = m.search(_cam(c), empty, [], 0.5, 1.0, False, 30.0, 0.8, want_in_view=True)
    f = op.is_in_frustum(c["cam"], c["P"], c["N"], c["min_d"], c["max_d"], 0.5)
    want = (f["in_view"] > 0) & (c["bad"] == 0)
    assert n == 0 and n_in == want.sum() and (inv == want).all() and (best == -1).all()     # isInFrustum still runs
    # slots beyond the written range and out-of-range slots in `order` count as bad points
    order = np.array([0, 1, 4999, 2, -7, 3, 6000], np.int32)
    n, n_in, mp, ft, best, inv = m.search(_cam(c), _frame(k, d, sf, c), [], 0.5, 1.0, False, 30.0, 0.8, order=order, want_in_view=True)
    assert (inv[[2, 4, 6]] == 0).all() and (best[[2, 4, 6]] == -1).all()
    assert (inv[[0, 1, 3, 5]] == want[[0, 1, 2, 3]]).all()


@pytest.mark.parametrize("env_over", [{"CUORB_ROUND_BATCH": "1"}, {"CUORB_ROUND_BATCH": "6"}, {"CUORB_PDL": "0"}],
                         ids=lambda d: ",".join(f"{k[6:]}={v}" for k, v in d.items()))
def test_round_scheduling_switches_subprocess(env_over):
    """How the binding rounds are enqueued (rounds per host synchronisation, dependent launches or plain stream order)
    must not show in the result: one round per sync walks the batch-to-batch carry-over of the dirty cells every
    round.  The switches are read once per process, hence the subprocess, which re-runs tests of this file and the
    crafted tie-heavy frame of test_gpu_localmap_resident.py."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, **env_over)
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-m", "gpu", "-x",
                        "tests/test_gpu_localmap.py::test_localmap_search_equals_its_two_parts",
                        "tests/test_gpu_localmap.py::test_localmap_processing_order_and_in_view",
                        "tests/test_gpu_localmap_resident.py::test_resident_cell_order_decides_ties"],
                       env=env, cwd=root, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and " passed" in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]
