"""Synthetic (KeyFrame, Frame) pairs for the vocabulary-node searches: side-2 descriptors are noisy
copies of side-1 descriptors (plus outliers), features are binned into ~100 vocabulary nodes by a hash of
the clean descriptor (so true matches mostly share a node, as with a real vocabulary), a fraction of the
side-1 features has no MapPoint, and keypoint angles follow a common rotation plus outliers."""
import numpy as np


def make_pair(seed, n1=1100, n2=1000, n_nodes=100, tie_heavy=False):
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
    if tie_heavy:
        base[:, 4:] = 0                                     # few distinct distances -> many ties
    node_of_base = (base[:, 0].astype(np.int64) * 7 + base[:, 1]) % n_nodes
    src = rng.integers(0, n1, n2)
    noise = rng.integers(0, 256, (n2, 32), dtype=np.uint8) & rng.integers(0, 256, (n2, 32), dtype=np.uint8) & \
        rng.integers(0, 256, (n2, 32), dtype=np.uint8) & rng.integers(0, 256, (n2, 32), dtype=np.uint8)
    if tie_heavy:
        noise[:, 4:] = 0
    desc2 = base[src] ^ noise
    outl = rng.random(n2) < 0.15
    desc2[outl] = rng.integers(0, 256, (int(outl.sum()), 32), dtype=np.uint8)
    node2 = node_of_base[src].copy()
    node2[outl] = rng.integers(0, n_nodes + 5, int(outl.sum()))   # some nodes only exist on one side
    valid1 = (rng.random(n1) < 0.7).astype(np.uint8)
    elig2 = (rng.random(n2) < 0.8).astype(np.uint8)
    angle1 = rng.uniform(0, 360, n1).astype(np.float32)
    angle2 = ((angle1[src] - np.float32(37.0) + rng.normal(0, 3, n2)) % 360).astype(np.float32)
    wild = rng.random(n2) < 0.2
    angle2[wild] = rng.uniform(0, 360, int(wild.sum())).astype(np.float32)
    angle2 = np.clip(angle2, 0, np.float32(359.99))
    fv1 = {}
    for i in rng.permutation(n1):                           # FeatureVector lists are in insertion order, not sorted
        fv1.setdefault(int(node_of_base[i]), []).append(int(i))
    fv2 = {}
    for j in rng.permutation(n2):
        fv2.setdefault(int(node2[j]), []).append(int(j))
    return dict(desc1=base, valid1=valid1, angle1=angle1, fv1=fv1, desc2=desc2, elig2=elig2, angle2=angle2, fv2=fv2)
