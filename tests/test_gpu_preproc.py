"""GPU: libcuorb's pre-extraction pass (k_preproc, through the C ABI) against the cv2 goldens and
the CPU oracle — bit-exact — and the fused preprocess+extract call against extraction of the
oracle-prepared frame."""
import numpy as np
import pytest

from tests.preproc_cases import CASES, FMT_CH, GOLDEN, color_frames, rectify_maps, sha

pytestmark = pytest.mark.gpu


def _pre(src_w, src_h, mode, maps=None, dsize=None, max_batch=1):
    from orbslam3_cpp_fork_b200 import FramePreprocessor
    return FramePreprocessor((src_w, src_h), mode, maps=maps, dsize=dsize, max_batch=max_batch)


@pytest.mark.parametrize("fmt", sorted(FMT_CH))
@pytest.mark.parametrize("case", sorted(CASES))
def test_preproc_matches_cv2_golden(case, fmt):
    g = np.load(GOLDEN)
    mode, arg = CASES[case]
    src = g[f"src{fmt}"]
    h, w = src.shape[:2]
    pre = _pre(w, h, mode, maps=(g[arg[0]], g[arg[1]]) if mode == 1 else None, dsize=arg if mode == 2 else None)
    out = pre(src[None], fmt)[0]
    assert list(out.shape) == g[f"{case}{fmt}_shape"].tolist()
    if f"{case}{fmt}" in g:
        assert (out == g[f"{case}{fmt}"]).all()
    assert sha(out) == str(g[f"{case}{fmt}_sha"])


@pytest.mark.parametrize("fmt", [0, 1, 4])
def test_preproc_full_size_batch_vs_oracle(fmt):
    from oracle import oracle_py as op
    w, h, n = 752, 480, 3
    mx, my = rectify_maps(w, h)
    frames = color_frames(n, h, w, FMT_CH[fmt], 60 + fmt)
    pre = _pre(w, h, 1, maps=(mx, my), max_batch=n)
    out = pre(frames, fmt)
    for i in range(n):
        assert (out[i] == op.preprocess(frames[i], fmt, op.PRE_REMAP, mx, my)).all()
    # device-resident source and a host source with row padding
    import torch
    out_d = pre(torch.from_numpy(frames).cuda(), fmt)
    assert (out_d == out).all()
    padded = np.zeros((n, h + 3, w + 5) + frames.shape[3:], np.uint8)
    padded[:, :h, :w] = frames
    assert (pre(padded[:, :h, :w], fmt) == out).all()


def test_resize_then_extract_fused_equals_oracle_chain():
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import ORBextractor
    n = 2
    frames = color_frames(n, 960, 1280, 3, 70)
    pre = _pre(1280, 960, 2, dsize=(752, 480), max_batch=n)
    ext = ORBextractor(1000, 1.2, 8, 20, 7, max_width=752, max_height=480, max_batch=n)
    pre.extract_batch_async(ext, frames, 2, (0, 0))
    counts, mono, kps, desc = ext.fetch()
    ref = op.OracleExtractor(1000, 1.2, 8, 20, 7)
    for i in range(n):
        gray = op.preprocess(frames[i], 2, op.PRE_RESIZE, dsize=(752, 480))
        m, k, d = ref.extract(gray, (0, 0))
        assert counts[i] == len(k) and mono[i] == m
        got = kps[i, :len(k)]
        for f in ("x", "y", "size", "response", "octave"):
            assert (got[f] == k[f]).all()
        assert np.abs(np.deg2rad(got["angle"] - k["angle"])).max() <= 1e-4   # north_star tolerance
        assert (desc[i, :len(k)] == d).all()


def test_remap_then_extract_fused_gray_and_errors():
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import ORBextractor, _lib
    w, h = 752, 480
    mx, my = rectify_maps(w, h, 1)
    frame = color_frames(1, h, w, 1, 80)
    pre = _pre(w, h, 1, maps=(mx, my))
    ext = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h)
    pre.extract_batch_async(ext, frame, 0, (0, 0))
    counts, mono, kps, desc = ext.fetch()
    m, k, d = op.OracleExtractor(1000, 1.2, 8, 20, 7).extract(op.preprocess(frame[0], 0, op.PRE_REMAP, mx, my), (0, 0))
    assert counts[0] == len(k) and (desc[0, :len(k)] == d).all()
    assert (kps[0, :len(k)]["x"] == k["x"]).all() and (kps[0, :len(k)]["y"] == k["y"]).all()
    with pytest.raises(_lib.CuorbError):            # batch larger than the handle's max_batch
        pre(np.zeros((2, h, w), np.uint8), 0)
    with pytest.raises(ValueError):                 # wrong channel count for the format
        pre(np.zeros((1, h, w, 3), np.uint8), 4)
