"""CPU: the oracle's restatement of the two-camera (Frame::Nleft != -1) branches of the tracking-time searches
(oracle/rig_oracle.cpp) against the reference's OWN functions (unmodified src/ORBmatcher.cc in oracle/_ref, driven by
oracle/ref_compat/ref_capi_rig.cpp).  Bit for bit: final F.mvpMapPoints over both cameras' features, the match count
(which counts cross-bound stereo partners), and per MapPoint the features it was written to."""
import numpy as np
import pytest

from oracle import oracle_py as op
from oracle import ref_py as rp
from tests.rig_cases import frame_views, rig_lastframe, rig_mappoints, rig_replay, rig_replay_lastframe, rig_views

pytestmark = pytest.mark.skipif(not rp.available(), reason="oracle/_ref not built and no reference checkout")


@pytest.fixture(scope="module")
def rv():
    return rig_views(7)


@pytest.mark.parametrize("seed,M,th,far,no_obs,ratio", [(1, 3000, 1.0, False, 0.1, 0.8), (2, 3000, 3.0, True, 0.0, 0.8),
                                                      (3, 1500, 3.0, False, 0.5, 0.9), (4, 4000, 1.0, False, 0.1, 0.6)])
def test_mappoints_rig_equals_reference(rv, seed, M, th, far, no_obs, ratio):
    q = rig_mappoints(rv, M, seed, no_obs_frac=no_obs)
    args = [q[k] for k in ("desc", "px", "py", "level", "view_cos", "pxr", "pyr", "level_r", "view_cos_r", "depth", "flags")]
    fvl, fvr, keep = frame_views(rv)
    n, bl, br = op.search_by_projection_rig(fvl, fvr, rv["l2r"], rv["r2l"], *args, th, far, 40.0, ratio)
    fvl, fvr, keep = frame_views(rv)
    rn, rbind, rper = rp.search_by_projection_rig(fvl, fvr, rv["l2r"], rv["r2l"], *args, th, far, 40.0, ratio)
    bind, cnt, per = rig_replay(bl, br, rv)
    assert rn == n == cnt and n > 300
    assert (rbind == bind).all()
    assert (rper == per).all()
    assert (bl >= 0).sum() > 100 and (br >= 0).sum() > 100 and n - (bl >= 0).sum() - (br >= 0).sum() > 100   # cross-binds


@pytest.mark.parametrize("seed,fwd,bwd,th,ori", [(1, False, False, 7.0, True), (2, True, False, 15.0, True),
                                                (3, False, True, 7.0, True), (4, False, False, 15.0, False)])
def test_lastframe_rig_equals_reference(rv, seed, fwd, bwd, th, ori):
    q = rig_lastframe(rv, seed)
    fvl, fvr, keep = frame_views(rv)
    n, bl, br, rl, rr, kb = op.search_by_projection_frame_rig(fvl, fvr, q["desc"], q["u"], q["v"], q["inv_z"], q["ur"], q["vr"],
                                                              q["octave"], q["angle"], q["flags"], th, fwd, bwd, ori)
    fvl, fvr, keep = frame_views(rv)
    rn, rbind = rp.search_by_projection_frame_rig(fvl, fvr, q["desc"], q["u"], q["v"], q["z"], q["trl"][0], q["trl"][1],
                                                  q["octave"], q["angle"], q["flags"], th, fwd, bwd, ori)
    bind, cnt = rig_replay_lastframe(bl, br, rl, rr, kb, rv, ori)
    assert rn == n == cnt and n > 100
    assert (rbind == bind).all()
    assert (bl >= 0).sum() > 60 and (br >= 0).sum() > 60
