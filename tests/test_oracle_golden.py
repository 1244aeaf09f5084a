"""CPU: the C++ oracle against the committed golden vectors (made with the real cv2 primitives by
tests/golden/make_golden.py) and, when cv2 is importable, against cv2 directly."""
import glob
import hashlib
import os

import numpy as np
import pytest

from oracle import oracle_py as op
from orbslam3_cpp_fork_b200 import synth

GOLDEN = sorted(p for p in glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz"))
                if not os.path.basename(p).startswith(("preproc_", "frameglue_", "bfmatch_")))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def golden_image(g):
    if "image" in g:
        return g["image"]
    w, h, seed = (int(v) for v in g["image_spec"])
    img = synth.synth_image(w, h, seed)
    assert sha(img) == str(g["image_sha"]), "synthetic generator drifted from the golden input"
    return img


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_matches_golden(path):
    g = np.load(path)
    img = golden_image(g)
    e = op.OracleExtractor(int(g["nfeatures"]), 1.2, 8, 20, 7)
    mono, k, d = e.extract(img, tuple(int(v) for v in g["lap"]))
    assert mono == int(g["mono"])
    assert len(k) == len(g["kps"])
    assert (k == g["kps"]).all()          # bit-exact incl. float angle
    assert (d == g["desc"]).all()
    for lv in range(8):
        assert sha(e.plane(lv, 0)) == str(g["pyr_sha"][lv])
        b = e.plane(lv, 2)
        assert (sha(b) if b is not None else "") == str(g["blur_sha"][lv])
        xs, ys, rs = e.candidates(lv)
        assert len(xs) == int(g["cand_count"][lv])
        assert sha(np.stack([xs, ys, rs], 1).astype(np.float32).reshape(-1, 3)) == str(g["cand_sha"][lv])


def test_goldens_exist():
    assert len(GOLDEN) >= 5


def test_tables_cfg1():
    e = op.OracleExtractor(1000, 1.2, 8, 20, 7)
    sf, inv, s2, is2, per = e.tables()
    assert per.tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert e.umax().tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert [e2 for e2 in op.OracleExtractor(2000, 1.2, 8, 20, 7).tables()[4]] == [434, 362, 302, 251, 209, 175, 145, 122]


def test_empty_image_returns_minus_one():
    e = op.OracleExtractor()
    mono, k, d = e.extract(np.zeros((0, 0), np.uint8))
    assert mono == -1 and len(k) == 0


cv2 = pytest.importorskip("cv2") if os.environ.get("CUORB_SKIP_CV2") is None else None


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
@pytest.mark.parametrize("w,h", [(97, 131), (752, 480), (1241, 376), (64, 50)])
def test_primitives_vs_cv2(w, h):
    rng = np.random.default_rng(w * 1000 + h)
    img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    assert (op.blur(img) == cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)).all()
    dw, dh = int(np.rint(np.float32(w) / np.float32(1.2))), int(np.rint(np.float32(h) / np.float32(1.2)))
    assert (op.resize(img, dw, dh) == cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR)).all()
    assert (op.border(img) == cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)).all()
    for t in (20, 7):
        k = op.fast(img, t)
        det = cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        ref = np.array([(p.pt[0], p.pt[1], p.response) for p in det.detect(img)], np.float32).reshape(-1, 3)
        assert ref.shape[0] == len(k)
        assert (np.stack([k["x"], k["y"], k["response"]], 1) == ref).all()   # same corners, scores AND order


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_fast_atan2_vs_cv2():
    rng = np.random.default_rng(3)
    y = rng.integers(-300000, 300000, 20000).astype(np.float32)
    x = rng.integers(-300000, 300000, 20000).astype(np.float32)
    y[:10] = 0; x[10:20] = 0
    ref = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    assert (op.fast_atan2(y, x) == ref).all()


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_full_pipeline_vs_cv2_reference():
    from oracle import cv2_reference as cr
    img = synth.synth_image(400, 300, 21)
    e = op.OracleExtractor(600, 1.2, 8, 20, 7)
    mono, k, d = e.extract(img, (50, 120))
    mono2, k2, d2 = cr.extract(img, 600, 1.2, 8, 20, 7, (50, 120))
    assert mono == mono2 and (k == k2).all() and (d == d2).all()
