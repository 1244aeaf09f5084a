"""CPU: sanity and invariants of the oracle's matching half (stereo association, projection
searches) on the synthetic cases the GPU parity tests use."""
import numpy as np

from oracle import oracle_py as op
from orbslam3_cpp_fork_b200 import synth
from tests.matching_cases import extracted_frame, lastframe_case, mappoint_case


def test_stereo_oracle_recovers_disparity():
    w, h = 480, 320
    L, R = synth.synth_stereo_pair(w, h, 3)
    eL, eR = op.OracleExtractor(600), op.OracleExtractor(600)
    _, kl, dl = eL.extract(L); _, kr, dr = eR.extract(R)
    mbf = 200.0; mb = mbf / 400.0
    s = op.stereo(eL, eR, kl, dl, kr, dr, mb, mbf)
    m = s["u_right"] >= 0
    assert m.sum() > 100
    d = synth.disparity_field(w, h, 3)
    xi = np.clip(np.rint(s["u_right"][m]).astype(int), 0, w - 1); yi = np.clip(np.rint(kl["y"][m]).astype(int), 0, h - 1)
    err = np.abs((kl["x"][m] - s["u_right"][m]) - d[yi, xi])
    assert np.median(err) < 1.5                                  # sub-pixel refinement lands on the true disparity
    assert np.allclose(s["depth"][m], mbf / (kl["x"][m] - s["u_right"][m]), rtol=1e-5)
    assert ((s["depth"] < 0) == (s["u_right"] < 0)).all()


def test_projection_oracle_invariants():
    e, k, d = extracted_frame(480, 320, 600, 9)
    sf = e.tables()[0]
    c = mappoint_case(k, d, sf, 4000, 1, 480, 320, with_uright=True)
    fv, keep = op.make_frame_view(k, d, c["u_right"], c["bound"], 480, 320, sf, mb=0.1, mbf=40.0)
    n, best = op.search_by_projection(fv, c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"],
                                      c["depth"], c["flags"], 3.0, True, 60.0, 0.8)
    assert n == (best >= 0).sum() and n > 200
    hit = best[best >= 0]
    assert (c["bound"][hit] == 0).all()                          # never binds an already-bound feature
    obs = (c["flags"][best >= 0] & 4) != 0
    assert len(np.unique(hit[obs])) == obs.sum()                 # a feature taken by an observed MapPoint is taken once
    assert ((c["flags"][best >= 0] & 1) != 0).all() and ((c["flags"][best >= 0] & 2) == 0).all()


def test_lastframe_oracle_invariants():
    e, k, d = extracted_frame(480, 320, 600, 10)
    sf = e.tables()[0]
    c = lastframe_case(k, d, 2, 480, 320)
    fv, keep = op.make_frame_view(k, d, c["u_right"], None, 480, 320, sf, mb=0.1, mbf=40.0)
    n, best, rot, keepb = op.search_by_projection_frame(fv, c["desc"], c["u"], c["v"], c["inv_z"], c["octave"],
                                                        c["angle"], c["flags"], 7.0, False, False, True)
    assert (best[:15] == -1).all()                               # out-of-bounds / behind-camera points
    assert keepb[0] >= 0 and n == sum(1 for i in range(len(best)) if best[i] >= 0 and rot[i] in keepb)
    assert n > 100
