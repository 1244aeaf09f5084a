"""CPU: the BoW-search oracle against an independent pure-Python transcription of the reference loop."""
import numpy as np
import pytest

from oracle import oracle_py as op
from orbslam3_cpp_fork_b200.matcher import join_feature_vectors
from tests.bow_cases import make_pair


def bow_python(c, elig2, mode, ratio, ori):
    d1, d2 = c["desc1"], c["desc2"]
    match = np.full(len(d1), -1, np.int64); matched2 = np.zeros(len(d2), bool)
    hist = [[] for _ in range(30)]
    n = 0
    for node in sorted(set(c["fv1"]) & set(c["fv2"])):
        for i1 in c["fv1"][node]:
            if not c["valid1"][i1]:
                continue
            b1, b2, bi = 256, 256, -1
            for i2 in c["fv2"][node]:
                if matched2[i2] or not elig2[i2]:
                    continue
                d = int(np.unpackbits(d1[i1] ^ d2[i2]).sum())
                if d < b1:
                    b2, b1, bi = b1, d, i2
                elif d < b2:
                    b2 = d
            if (b1 <= 50 if mode == 0 else b1 < 50) and np.float32(b1) < np.float32(ratio) * np.float32(b2):
                match[i1] = bi; matched2[bi] = True; n += 1
                if ori:
                    rot = np.float32(c["angle1"][i1]) - np.float32(c["angle2"][bi])
                    if rot < 0:
                        rot = np.float32(rot + np.float32(360.0))
                    x = float(np.float32(rot * np.float32(1.0 / 30)))
                    b = int(np.floor(x + 0.5)) if x >= 0 else int(np.ceil(x - 0.5))    # std::round: half away from zero
                    hist[0 if b == 30 else b].append(i1)
    if ori:
        sizes = [len(h) for h in hist]
        m1 = m2 = m3 = 0; i1_ = i2_ = i3_ = -1
        for i, s in enumerate(sizes):
            if s > m1:
                m3, m2, m1, i3_, i2_, i1_ = m2, m1, s, i2_, i1_, i
            elif s > m2:
                m3, m2, i3_, i2_ = m2, s, i2_, i
            elif s > m3:
                m3, i3_ = s, i
        if m2 < np.float32(0.1) * np.float32(m1):
            i2_ = i3_ = -1
        elif m3 < np.float32(0.1) * np.float32(m1):
            i3_ = -1
        for i, h in enumerate(hist):
            if i not in (i1_, i2_, i3_):
                for k in h:
                    match[k] = -1; n -= 1
    return n, match


@pytest.mark.parametrize("seed,tie_heavy,mode", [(1, False, 0), (2, True, 1), (5, True, 0)])
def test_bow_oracle_equals_python_transcription(seed, tie_heavy, mode):
    c = make_pair(seed, n1=260, n2=240, n_nodes=30, tie_heavy=tie_heavy)
    elig2 = np.ones(len(c["desc2"]), np.uint8) if mode == 0 else c["elig2"]
    off1, idx1, off2, idx2 = join_feature_vectors(c["fv1"], c["fv2"])
    n, match = op.search_by_bow(c["desc1"], c["valid1"], c["angle1"], c["desc2"], elig2, c["angle2"], off1, idx1, off2, idx2,
                                mode, 0.75, True)
    pn, pmatch = bow_python(c, elig2, mode, 0.75, True)
    assert n == pn and (match == pmatch).all() and n > 10
