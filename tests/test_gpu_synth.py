"""GPU (-m gpu): the device-side frame generator of cfg5 (csrc/synth.cu) against its numpy definition
(orbslam3_cpp_fork_b200/synth_cb.py), byte for byte."""
import numpy as np
import pytest

from orbslam3_cpp_fork_b200 import synth_cb

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("w,h,first,n", [(752, 480, 0, 24), (752, 480, 99_990, 8), (1241, 376, 5, 4), (333, 251, 7, 3)])
def test_device_generator_equals_host(w, h, first, n):
    import torch
    dl, dr = synth_cb.stereo_batch_device(first, n, w, h)
    torch.cuda.synchronize()
    dl, dr = dl.cpu().numpy(), dr.cpu().numpy()
    for i in range(n):
        l, r = synth_cb.stereo_pair(first + i, w, h)
        assert (dl[i] == l).all(), f"left eye of frame {first + i}"
        assert (dr[i] == r).all(), f"right eye of frame {first + i}"


def _host_digest(frame_id):
    import hashlib
    l, r = synth_cb.stereo_pair(int(frame_id))
    return hashlib.sha256(l.tobytes() + r.tobytes()).hexdigest()


def test_certified_subset_1000_pairs():
    """a 1,000-pair subset of the 100k sweep (every 100th frame id) generated on the host
    (numpy, one process per core) and compared with what the device produced for the same frame ids,
    through SHA-256 of the 2 x 752 x 480 bytes of each pair."""
    import hashlib
    import os
    from concurrent.futures import ProcessPoolExecutor
    import torch
    ids = list(range(0, 100_000, 100))
    import multiprocessing as mp
    with ProcessPoolExecutor(max_workers=min(32, os.cpu_count() or 1), mp_context=mp.get_context("spawn")) as ex:
        fut = ex.map(_host_digest, ids, chunksize=8)
        dev = {}
        for f in ids:
            dl, dr = synth_cb.stereo_batch_device(f, 1)
            dev[f] = hashlib.sha256(dl[0].cpu().numpy().tobytes() + dr[0].cpu().numpy().tobytes()).hexdigest()
        host = dict(zip(ids, fut))
    bad = [f for f in ids if host[f] != dev[f]]
    assert not bad, bad
    assert len(set(dev.values())) == len(ids)        # 1,000 distinct pairs


def test_generated_frames_feed_the_extractor():
    """The generated frames are as feature-rich as the numpy-PCG ones: quotas saturate on every level."""
    import orbslam3_cpp_fork_b200 as cu
    dl, dr = synth_cb.stereo_batch_device(1234, 4)
    e = cu.ORBextractor(1000, max_batch=4)
    e.extract_batch_async(dl)
    counts, mono, k, d = e.fetch()
    assert (counts >= 1000).all() and (counts <= 1040).all()
