"""Stage-by-stage parity report (GPU box): libcuorb vs the CPU oracle. Prints, never asserts."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from orbslam3_cpp_fork_b200 import ORBextractor, ORBmatcher, synth
from oracle import oracle_py as op

def report(w, h, nf, seed, lap=(0, 0), kind=None):
    img = synth.degenerate_image(w, h, kind, seed) if kind else synth.synth_image(w, h, seed)
    e = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=2)
    o = op.OracleExtractor(nf, 1.2, 8, 20, 7)
    t = time.time(); mono, k, d = e(img, None, lap); tg = time.time() - t
    mo, ko, do = o.extract(img, lap)
    print(f"== {w}x{h} nf={nf} seed={seed} lap={lap} kind={kind}: gpu n={len(k)} mono={mono} | oracle n={len(ko)} mono={mo} ({tg*1e3:.1f} ms)")
    for lv in range(8):
        okp = (e.download_plane(lv, 0) == o.plane(lv, 0))
        ob = o.plane(lv, 2)
        okb = None if ob is None else (e.download_plane(lv, 2) != ob).sum()
        xs, ys, rs = e.download_candidates(lv); xo, yo, ro = o.candidates(lv)
        same = len(xs) == len(xo) and (xs == xo).all() and (ys == yo).all() and (rs == ro).all()
        print(f"  L{lv}: pyr mism {(~okp).sum()}  blur mism {okb}  cand gpu {len(xs)} oracle {len(xo)} same {same}")
    if len(k) == len(ko):
        pos = (k['x'] == ko['x']) & (k['y'] == ko['y']) & (k['octave'] == ko['octave']) & (k['response'] == ko['response']) & (k['size'] == ko['size'])
        ang = k['angle'] == ko['angle']
        print(f"  kps pos/oct/resp equal {pos.sum()}/{len(k)}  angle bit-equal {ang.sum()}  max|dangle| {np.abs(k['angle']-ko['angle']).max() if len(k) else 0}")
        dd = (d != do)
        print(f"  descriptor rows differing {dd.any(axis=1).sum()}  bits differing {np.unpackbits(d ^ do).sum()}")
    else:
        ks = set(zip(k['x'].tolist(), k['y'].tolist(), k['octave'].tolist())); kos = set(zip(ko['x'].tolist(), ko['y'].tolist(), ko['octave'].tolist()))
        print(f"  COUNT MISMATCH: common {len(ks & kos)}")
    return e, o, k, d, ko, do

report(752, 480, 1000, 1)
report(752, 480, 1000, 2, (0, 1000))
report(640, 480, 1000, 3)
report(1241, 376, 2000, 4)
report(320, 240, 500, 5, (100, 200))
for kind in ("black", "flat", "noise"):
    report(752, 480, 1000, 6, kind=kind)

# stereo
w, h = 752, 480
L, R = synth.synth_stereo_pair(w, h, 11)
eL = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=2)
eR = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=2)
eL.extract_batch_async(np.stack([L, L])); eR.extract_batch_async(np.stack([R, R]))
cl, ml, kl, dl = eL.fetch(); cr, mr, kr, dr = eR.fetch()
mbf = 435.2; mb = mbf / 458.654
u, dep, dbg = ORBmatcher().SearchStereo(eL, eR, mb, mbf, debug=True)
oL, oR = op.OracleExtractor(), op.OracleExtractor()
_, k0, d0 = oL.extract(L); _, k1, d1 = oR.extract(R)
so = op.stereo(oL, oR, k0, d0, k1, d1, mb, mbf)
n = cl[0]
print("stereo: n", n, len(k0), "batch images equal", (kl[0] == kl[1]).all(), (u[0] == u[1]).all())
if n == len(k0):
    for name, a, b in (("u_right", u[0, :n], so["u_right"]), ("depth", dep[0, :n], so["depth"]),
                       ("best_idx_r", dbg["best_idx_r"][0, :n], so["best_idx_r"]), ("best_dist", dbg["best_dist"][0, :n], so["best_dist"]),
                       ("sad", dbg["sad"][0, :n], so["sad"])):
        print(f"  {name}: mismatches {(a != b).sum()} / {n}   (matched: {(b >= 0).sum()})")
