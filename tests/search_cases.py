"""Synthetic inputs for the ORBmatcher searches outside the tracking path, shared
by the CPU and GPU tests.  Frames / KeyFrames are real extractions of the synthetic stereo pair (so the
second view's features are shifted copies of the first's); queries are MapPoints whose descriptors are
bit-flipped copies of target features and whose projections land near them, with duplicates (several
queries chasing one feature: the order-dependent paths), inactive entries and pre-bound features."""
import numpy as np

from oracle import oracle_py as op
from orbslam3_cpp_fork_b200 import synth
from tests.matching_cases import flip_bits

W, H = 480, 320


def two_views(seed, nf=700, tie_heavy=False):
    L, R = synth.synth_stereo_pair(W, H, seed)
    e1, e2 = op.OracleExtractor(nf, 1.2, 8, 20, 7), op.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, k1, d1 = e1.extract(L)
    _, k2, d2 = e2.extract(R)
    if tie_heavy:                                            # few distinct distances: ties, steals
        d1 = d1.copy(); d2 = d2.copy(); d1[:, 1:] = 0; d2[:, 1:] = 0
    sf = np.asarray(e1.tables()[0], np.float32)
    return dict(k1=k1, d1=d1, k2=k2, d2=d2, sf=sf, sigma2=(sf * sf).astype(np.float32),
                inv_sigma2=(np.float32(1.0) / (sf * sf)).astype(np.float32))


def point_queries(k, d, sf, n, seed, jitter=4.0, kmax=45):
    """n MapPoints aimed at random features of (k, d) (with replacement: collisions are wanted)."""
    rng = np.random.default_rng(seed)
    N = len(k)
    tgt = rng.integers(0, N, n)
    desc = flip_bits(d[tgt], rng, kmax)
    rnd = rng.random(n) < 0.2
    desc[rnd] = rng.integers(0, 256, (int(rnd.sum()), 32), dtype=np.uint8)
    u = (k["x"][tgt] + rng.uniform(-jitter, jitter, n)).astype(np.float32)
    v = (k["y"][tgt] + rng.uniform(-jitter, jitter, n)).astype(np.float32)
    u[:4] = -5.0; v[4:8] = H + 3.0                           # outside the image
    u[8:12] = W - 0.5                                        # window hangs over the right edge
    level = np.clip(k["octave"][tgt] + rng.integers(0, 2, n), 0, len(sf) - 1).astype(np.int32)
    angle = np.mod(k["angle"][tgt] + rng.normal(0, 6, n), 360).astype(np.float32)
    angle[rng.random(n) < 0.15] = rng.uniform(0, 359, 1).astype(np.float32)[0]
    angle = np.clip(angle, 0, np.float32(359.99)).astype(np.float32)
    active = (rng.random(n) < 0.85).astype(np.uint8)
    ur = (u - rng.uniform(1, 40, n)).astype(np.float32)
    bound = (rng.random(N) < 0.2).astype(np.uint8)
    u_right = np.where(rng.random(N) < 0.5, k["x"] - rng.uniform(1, 40, N), -1).astype(np.float32)
    ok = u_right[tgt] >= 0
    ur = np.where(ok & (rng.random(n) < 0.8), u_right[tgt] + rng.uniform(-2, 2, n), ur).astype(np.float32)
    return dict(tgt=tgt, desc=desc, u=u, v=v, level=level, angle=angle, active=active, ur=ur, bound=bound,
                u_right=u_right)


def sim3_case(tv, seed):
    """Mutual projections between two KeyFrames along a random partial bijection."""
    rng = np.random.default_rng(seed)
    k1, d1, k2, d2, sf = tv["k1"], tv["d1"], tv["k2"], tv["d2"], tv["sf"]
    N1, N2 = len(k1), len(k2)
    npair = int(0.6 * min(N1, N2))
    p1 = rng.permutation(N1)[:npair]; p2 = rng.permutation(N2)[:npair]

    def side(ka, Na, kb, db, Nb, pa, pb):
        desc = rng.integers(0, 256, (Na, 32), dtype=np.uint8)
        u = rng.uniform(0, W, Na).astype(np.float32); v = rng.uniform(0, H, Na).astype(np.float32)
        level = rng.integers(0, len(sf), Na).astype(np.int32)
        desc[pa] = flip_bits(db[pb], rng, 50)
        u[pa] = kb["x"][pb] + rng.uniform(-3, 3, npair).astype(np.float32)
        v[pa] = kb["y"][pb] + rng.uniform(-3, 3, npair).astype(np.float32)
        level[pa] = np.clip(kb["octave"][pb] + rng.integers(0, 2, npair), 0, len(sf) - 1)
        active = (rng.random(Na) < 0.8).astype(np.uint8)
        return desc, u, v, level, active

    a = side(k1, N1, k2, d2, N2, p1, p2)
    b = side(k2, N2, k1, d1, N1, p2, p1)
    return a, b


def row_nodes(k, band=10):
    """Stand-in vocabulary nodes: features of the same image band share a node (insertion order shuffled)."""
    fv = {}
    order = np.random.default_rng(len(k)).permutation(len(k))
    for i in order:
        fv.setdefault(int(k["y"][i]) // band, []).append(int(i))
    return fv


def triangulation_case(tv, seed):
    rng = np.random.default_rng(seed)
    k1, k2 = tv["k1"], tv["k2"]
    N1, N2 = len(k1), len(k2)
    # rectified pair with a small perturbation so every entry of F12 takes part in the arithmetic
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32) + rng.normal(0, 2e-4, (3, 3)).astype(np.float32)
    ep = np.array([W * 0.5 + 30.0, H * 0.5 - 12.0], np.float32)      # inside the image: the epipole gate bites
    has1 = (rng.random(N1) < 0.3).astype(np.uint8); has2 = (rng.random(N2) < 0.3).astype(np.uint8)
    ur1 = np.where(rng.random(N1) < 0.4, k1["x"] - rng.uniform(1, 30, N1), -1).astype(np.float32)
    ur2 = np.where(rng.random(N2) < 0.4, k2["x"] - rng.uniform(1, 30, N2), -1).astype(np.float32)
    return dict(F12=F12, ep=ep, has1=has1, has2=has2, ur1=ur1, ur2=ur2, fv1=row_nodes(k1), fv2=row_nodes(k2))
