"""GPU: ORBmatcher::SearchByBoW (both overloads) through the C ABI against the CPU oracle — identical
matches, tie cases included."""
import numpy as np
import pytest

from tests.bow_cases import make_pair

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,tie_heavy,ratio,ori", [(1, False, 0.7, True), (2, True, 0.9, True), (3, False, 0.6, False), (4, True, 0.75, True)])
def test_search_by_bow_vs_oracle(seed, tie_heavy, ratio, ori):
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import ORBmatcher
    from orbslam3_cpp_fork_b200.matcher import join_feature_vectors
    c = make_pair(seed, tie_heavy=tie_heavy)
    m = ORBmatcher(ratio, ori)
    off1, idx1, off2, idx2 = join_feature_vectors(c["fv1"], c["fv2"])
    # KeyFrame - Frame
    n, match = m.SearchByBoW_KF_F(c["desc1"], c["valid1"], c["angle1"], c["fv1"], c["desc2"], c["angle2"], c["fv2"])
    rn, rmatch = op.search_by_bow(c["desc1"], c["valid1"], c["angle1"], c["desc2"], np.ones(len(c["desc2"]), np.uint8), c["angle2"],
                                  off1, idx1, off2, idx2, 0, ratio, ori)
    assert n == rn and (match == rmatch).all() and n == (match >= 0).sum() and n > 50
    assert len(set(match[match >= 0].tolist())) == n                    # one-to-one
    # KeyFrame - KeyFrame
    n, match = m.SearchByBoW_KF_KF(c["desc1"], c["valid1"], c["angle1"], c["fv1"], c["desc2"], c["elig2"], c["angle2"], c["fv2"])
    rn, rmatch = op.search_by_bow(c["desc1"], c["valid1"], c["angle1"], c["desc2"], c["elig2"], c["angle2"], off1, idx1, off2, idx2,
                                  1, ratio, ori)
    assert n == rn and (match == rmatch).all()
    assert (c["elig2"][match[match >= 0]] == 1).all() and (c["valid1"][match >= 0] == 1).all()


def test_search_by_bow_degenerate():
    from orbslam3_cpp_fork_b200 import ORBmatcher
    c = make_pair(9, n1=40, n2=30, n_nodes=5)
    m = ORBmatcher(0.7, True)
    n, match = m.SearchByBoW_KF_F(c["desc1"], c["valid1"], c["angle1"], {}, c["desc2"], c["angle2"], c["fv2"])   # no common node
    assert n == 0 and (match == -1).all()
    n, match = m.SearchByBoW_KF_F(c["desc1"], np.zeros(40, np.uint8), c["angle1"], c["fv1"], c["desc2"], c["angle2"], c["fv2"])
    assert n == 0 and (match == -1).all()                                   # no MapPoints on the KeyFrame
