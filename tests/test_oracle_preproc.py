"""CPU: the oracle's restatement of cv::remap / cv::resize / cv::cvtColor (oracle/preproc_oracle.cpp)
against golden vectors made with the real cv2 (tests/golden/make_preproc_golden.py) and, when cv2 is
importable, against cv2 directly at the reference's frame sizes."""
import numpy as np
import pytest

from oracle import oracle_py as op
from tests.preproc_cases import CASES, FMT_CH, GOLDEN, color_frames, rectify_maps, sha


@pytest.mark.parametrize("fmt", sorted(FMT_CH))
@pytest.mark.parametrize("case", sorted(CASES))
def test_oracle_preproc_matches_cv2_golden(case, fmt):
    g = np.load(GOLDEN)
    mode, arg = CASES[case]
    src = g[f"src{fmt}"]
    if mode == 1:
        out = op.preprocess(src, fmt, op.PRE_REMAP, g[arg[0]], g[arg[1]])
    elif mode == 2:
        out = op.preprocess(src, fmt, op.PRE_RESIZE, dsize=arg)
    else:
        out = op.preprocess(src, fmt)
    assert list(out.shape) == g[f"{case}{fmt}_shape"].tolist()
    if f"{case}{fmt}" in g:
        assert (out == g[f"{case}{fmt}"]).all()
    assert sha(out) == str(g[f"{case}{fmt}_sha"])


def test_oracle_preproc_matches_cv2_live():
    cv2 = pytest.importorskip("cv2")
    w, h = 752, 480
    mx, my = rectify_maps(w, h)
    for fmt, code in ((1, cv2.COLOR_RGB2GRAY), (2, cv2.COLOR_BGR2GRAY), (3, cv2.COLOR_RGBA2GRAY), (4, cv2.COLOR_BGRA2GRAY)):
        src = color_frames(1, h, w, FMT_CH[fmt], 40 + fmt)[0]
        assert (op.preprocess(src, fmt) == cv2.cvtColor(src, code)).all()
        assert (op.preprocess(src, fmt, op.PRE_REMAP, mx, my) ==
                cv2.cvtColor(cv2.remap(src, mx, my, cv2.INTER_LINEAR), code)).all()
    big = color_frames(1, 960, 1280, 3, 50)[0]
    assert (op.preprocess(big, 2, op.PRE_RESIZE, dsize=(752, 480)) ==
            cv2.cvtColor(cv2.resize(big, (752, 480)), cv2.COLOR_BGR2GRAY)).all()
    assert (op.preprocess(big, 2, op.PRE_RESIZE, dsize=(640, 480)) ==
            cv2.cvtColor(cv2.resize(big, (640, 480)), cv2.COLOR_BGR2GRAY)).all()
    g = color_frames(1, h, w, 1, 51)[0]
    assert (op.preprocess(g, 0, op.PRE_REMAP, mx, my) == cv2.remap(g, mx, my, cv2.INTER_LINEAR)).all()
