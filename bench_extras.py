"""Measurements of the BASELINE.json configs bench.py's headline line does not cover:
  cfg1  one 752x480 mono frame through cuorb_extract (the call Tracking makes): latency, next to the CPU
  cfg2  640x480 / 1000 features, batch 64 / 256 / 1024: frames/s and per-stage GB/s
  cfg3  1241x376 / 2000 features stereo pairs: pairs/s, stage ms, candidate descriptor pairs/s of the matcher
  cfg4  SearchByProjection: 1000 frame descriptors vs 50k MapPoints through the C ABI
  sustained  >= 5 s of the headline loop with the clock trace
They run on rank 0 at N=1 after the headline measurement and land in the bench line under "extra".
The CPU sides use oracle/_ref (the reference's own code) when present, else the oracle port — only as the
thing measured beside, never inside a GPU figure.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
import time

import numpy as np


def _pool(n, w, h, base=8, seed0=2000):
    from orbslam3_cpp_fork_b200 import synth
    Ls, Rs = [], []
    for i in range(min(base, n)):
        l, r = synth.synth_stereo_pair(w, h, seed0 + i)
        Ls.append(l); Rs.append(r)
    i = 0
    while len(Ls) < n:
        sh = 37 * (len(Ls) // base) + 11
        Ls.append(np.roll(Ls[i % base], sh, axis=1)); Rs.append(np.roll(Rs[i % base], sh, axis=1))
        i += 1
    return np.stack(Ls), np.stack(Rs)


def _stats(ts):
    a = np.sort(np.asarray(ts)) * 1e3
    return {"median_ms": float(a[len(a) // 2]), "p99_ms": float(a[min(len(a) - 1, int(0.99 * len(a)))]),
            "min_ms": float(a[0]), "calls": int(len(a))}


def _cpu_module():
    from oracle import ref_py
    if ref_py.available():
        try:
            ref_py.lib()
            return ref_py, "reference (oracle/_ref: the reference's own ORBextractor.cc / ComputeStereoMatches over scalar restated OpenCV primitives)"
        except Exception:
            pass
    from oracle import oracle_py
    return oracle_py, "port (oracle restatement)"


def cfg1_latency(local, with_cpu=True, calls=300):
    """One frame at a time through the synchronous drop-in calls, host buffers in and out."""
    from orbslam3_cpp_fork_b200 import ORBextractor, _lib, synth
    L_ = _lib.load()
    W, H = 752, 480
    imgL, imgR = synth.synth_stereo_pair(W, H, 4242)
    out = {"workload": "cfg1: one 752x480 frame, 1000 features, 8 levels, 1.2, FAST 20/7, lapping {0,1000}; "
                       "wall clock around the synchronous C-ABI call, pageable host image in, keypoints + descriptors out"}
    e = ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=1, device=local)
    eR = ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=1, device=local)
    cap = e.capacity
    bufs = [(np.zeros(cap, _lib.KP_DTYPE), np.zeros((cap, 32), np.uint8), C.c_int(0)) for _ in range(2)]

    def extract(ext, img, b, lap1):
        return L_.cuorb_extract(ext.handle, _lib.ptr(img), W, H, W, 0, lap1, _lib.ptr(b[0]), _lib.ptr(b[1]), cap, C.byref(b[2]))

    for _ in range(20):
        _lib.check(extract(e, imgL, bufs[0], 1000))
    ts = []
    for _ in range(calls):
        t0 = time.perf_counter(); extract(e, imgL, bufs[0], 1000); ts.append(time.perf_counter() - t0)
    out["cuorb_extract_mono"] = _stats(ts)
    out["keypoints"] = int(bufs[0][2].value)
    # the same call with the frame in pinned host memory (cuorb_host_alloc), as a capture pipeline that owns its
    # buffers would hand it over: the 361 KB upload is one DMA instead of a staged pageable copy
    nbytes = imgL.nbytes
    pin = L_.cuorb_host_alloc(nbytes)
    if pin:
        pimg = np.frombuffer((C.c_char * nbytes).from_address(pin), dtype=np.uint8).reshape(imgL.shape)
        pimg[...] = imgL
        for _ in range(10):
            _lib.check(extract(e, pimg, bufs[0], 1000))
        ts = []
        for _ in range(calls):
            t0 = time.perf_counter(); extract(e, pimg, bufs[0], 1000); ts.append(time.perf_counter() - t0)
        out["cuorb_extract_mono_pinned_frame"] = _stats(ts)
        del pimg
        L_.cuorb_host_free(pin)
    # where the time of one call goes on the device: per-stage CUDA events (stages serialised while profiling)
    e.set_profiling(True)
    acc = {}
    for _ in range(20):
        extract(e, imgL, bufs[0], 1000)
        for k_, v_ in e.stage_ms().items():
            acc.setdefault(k_, []).append(v_)
    e.set_profiling(False)
    out["device_stage_ms_batch1"] = {k_: float(np.median(v_)) for k_, v_ in acc.items() if k_ != "stereo"}
    out["device_stage_ms_batch1"]["sum"] = float(sum(out["device_stage_ms_batch1"].values()))

    # the stereo Frame constructor's sequence (src/Frame.cc:202-231): two ExtractORB threads on two handles,
    # then ComputeStereoMatches on the results
    u = np.zeros(cap, np.float32); d = np.zeros(cap, np.float32)
    MBF, FX = 47.90639384423901, 458.654

    def stereo_frame(threads):
        if threads:
            t = threading.Thread(target=extract, args=(eR, imgR, bufs[1], 0))
            t.start(); extract(e, imgL, bufs[0], 0); t.join()
        else:
            extract(e, imgL, bufs[0], 0); extract(eR, imgR, bufs[1], 0)
        nl, nr = bufs[0][2].value, bufs[1][2].value
        return L_.cuorb_stereo_match(e.handle, eR.handle, nl, _lib.ptr(bufs[0][0]), _lib.ptr(bufs[0][1]), nr,
                                     _lib.ptr(bufs[1][0]), _lib.ptr(bufs[1][1]), MBF / FX, MBF, _lib.ptr(u), _lib.ptr(d))

    for thr, key in ((False, "stereo_frame_sequential"), (True, "stereo_frame_two_threads")):
        for _ in range(10):
            _lib.check(stereo_frame(thr))
        ts = []
        for _ in range(calls // 2):
            t0 = time.perf_counter(); stereo_frame(thr); ts.append(time.perf_counter() - t0)
        out[key] = _stats(ts)
    if hasattr(L_, "cuorb_extract_stereo"):
        def fused():
            return L_.cuorb_extract_stereo(e.handle, eR.handle, _lib.ptr(imgL), _lib.ptr(imgR), W, H, W, MBF / FX, MBF,
                                           _lib.ptr(bufs[0][0]), _lib.ptr(bufs[0][1]), C.byref(bufs[0][2]),
                                           _lib.ptr(bufs[1][0]), _lib.ptr(bufs[1][1]), C.byref(bufs[1][2]), cap,
                                           _lib.ptr(u), _lib.ptr(d))
        for _ in range(10):
            _lib.check(fused())
        ts = []
        for _ in range(calls // 2):
            t0 = time.perf_counter(); fused(); ts.append(time.perf_counter() - t0)
        out["stereo_frame_fused_call"] = _stats(ts)
    out["stereo_matches"] = int((u[:bufs[0][2].value] >= 0).sum())
    e.close(); eR.close()

    if with_cpu:
        mod, kind = _cpu_module()
        Ext = getattr(mod, "RefExtractor", None) or mod.OracleExtractor
        oL, oR = Ext(1000), Ext(1000)
        oL.extract(imgL, (0, 1000))
        ts = []
        for _ in range(8):
            t0 = time.perf_counter(); oL.extract(imgL, (0, 1000)); ts.append(time.perf_counter() - t0)
        out["cpu_extract_mono_1core"] = dict(_stats(ts), kind=kind)
        sec, _, _ = mod.bench_stereo(np.stack([imgL] * 4), np.stack([imgR] * 4), MBF / FX, MBF, 1)
        out["cpu_stereo_frame_1thread_ms"] = 1e3 * sec / 4
        # two threads as the Frame ctor uses them: approximated by two frames in flight on two threads
        sec2, _, _ = mod.bench_stereo(np.stack([imgL] * 4), np.stack([imgR] * 4), MBF / FX, MBF, 2)
        out["cpu_stereo_frame_2threads_throughput_ms"] = 1e3 * sec2 / 4
        out["cpu_simd_primitives_1core"] = simd_primitives(imgL, reps=10)
    return out


def simd_primitives(img, reps=10, threads=1):
    """The OpenCV calls of one extraction (resize + copyMakeBorder + per-cell FAST with retry + GaussianBlur on
    8 levels) through the real SIMD cv2 of this image: what the reference's pixel stages cost with the library
    it links.  Quadtree / orientation / descriptors are not included (they are the reference's own scalar code)."""
    try:
        import cv2
    except Exception as ex:  # pragma: no cover
        return {"unavailable": f"cv2 not importable: {ex}"}
    cv2.setNumThreads(1)
    fast_ini = cv2.FastFeatureDetector_create(20, True)
    fast_min = cv2.FastFeatureDetector_create(7, True)

    def one(im):
        lv = im
        s = np.float32(1.0)
        n = 0
        for l in range(8):
            if l:
                s = np.float32(float(s) * float(np.float32(1.2)))
                inv = np.float32(1.0) / s
                sz = (int(np.rint(np.float32(im.shape[1]) * inv)), int(np.rint(np.float32(im.shape[0]) * inv)))
                lv = cv2.resize(lv, sz, interpolation=cv2.INTER_LINEAR)
            cv2.copyMakeBorder(lv, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
            h, w = lv.shape
            minB, maxBX, maxBY = 16, w - 16, h - 16
            width, height = maxBX - minB, maxBY - minB
            nC, nR = int(width / 35), int(height / 35)
            wC, hC = int(np.ceil(width / nC)), int(np.ceil(height / nR))
            for i in range(nR):
                iy = minB + i * hC
                my = min(iy + hC + 6, maxBY)
                if iy >= maxBY - 3:
                    continue
                for j in range(nC):
                    ix = minB + j * wC
                    mx = min(ix + wC + 6, maxBX)
                    if ix >= maxBX - 6:
                        continue
                    k = fast_ini.detect(lv[iy:my, ix:mx])
                    if not k:
                        k = fast_min.detect(lv[iy:my, ix:mx])
                    n += len(k)
            cv2.GaussianBlur(lv, (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101)
        return n

    one(img)
    if threads <= 1:
        t0 = time.perf_counter()
        for _ in range(reps):
            one(img)
        dt = (time.perf_counter() - t0) / reps
        return {"ms_per_image": dt * 1e3, "images_per_s": 1.0 / dt, "threads": 1, "cv2": cv2.__version__,
                "note": "resize + border + per-cell cv::FAST (ini/min retry) + GaussianBlur, 8 levels; python loop overhead included"}
    from concurrent.futures import ThreadPoolExecutor
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda _: one(img), range(reps * threads)))
    dt = time.perf_counter() - t0
    return {"images_per_s": reps * threads / dt, "threads": threads, "cv2": cv2.__version__,
            "note": "as above on a thread pool (cv2 releases the GIL inside its calls; the python cell loop does not)"}


def _stage_gbs(stage_ms, w, h, B):
    from bench import algorithmic_bytes
    ab = algorithmic_bytes(w, h)
    return {k: ab[k] * B / (stage_ms[k] * 1e-3) / 1e9 for k in ("pyramid", "blur", "fast") if stage_ms.get(k)}


def _resident_loop(pairs_or_handles, enqueue, sync, steps):
    import torch
    for i in range(3):
        enqueue(i)
    sync()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(steps):
        enqueue(i)
    sync()
    torch.cuda.synchronize()
    return time.perf_counter() - t0


def cfg2_batched(local, peak, batches=(64, 256, 1024)):
    """TUM RGB-D shape: 640x480, 1000 features, mono extraction only, frames resident in HBM."""
    import torch
    from orbslam3_cpp_fork_b200 import ORBextractor
    W, H = 640, 480
    Bmax = max(batches)
    Lh, _ = _pool(Bmax, W, H)
    Ld = torch.from_numpy(Lh).cuda(local)
    hs = [ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=Bmax, device=local) for _ in range(2)]
    res = {"workload": "cfg2: 640x480 frames, 1000 features, ORBextractor::operator() only, frames resident in HBM, "
                       "two handles alternating; stage times with one handle alone (CUDA events inside libcuorb)"}
    for B in batches:
        x = Ld[:B]
        steps = max(8, 4096 // B)
        dt = _resident_loop(hs, lambda i: hs[i & 1].extract_batch_async(x), lambda: [h.sync() for h in hs], steps)
        hs[0].set_profiling(True)
        acc = {}
        for _ in range(3):
            hs[0].extract_batch_async(x); hs[0].sync()
            for k, v in hs[0].stage_ms().items():
                acc.setdefault(k, []).append(v)
        hs[0].set_profiling(False)
        st = {k: float(np.median(v)) for k, v in acc.items() if k not in ("h2d", "stereo")}
        gbs = _stage_gbs(st, W, H, B)
        res[f"batch_{B}"] = {"frames_per_s": B * steps / dt, "ms_per_batch": 1e3 * dt / steps, "stage_ms": st,
                             "stage_gbs": gbs, "stage_frac_of_hbm_peak": {k: v / peak for k, v in gbs.items()}}
    for h in hs:
        h.close()
    return res


def cfg3_kitti(local, peak, B=96):
    """KITTI shape: 1241x376 stereo pairs, 2000 features: left + right extraction + ComputeStereoMatches."""
    import torch
    from orbslam3_cpp_fork_b200 import ORBextractor, _lib
    L_ = _lib.load()
    W, H, NF = 1241, 376, 2000
    FX, MBF = 718.856, 386.1448
    MB = MBF / FX
    Lh, Rh = _pool(B, W, H)
    Ld, Rd = torch.from_numpy(Lh).cuda(local), torch.from_numpy(Rh).cuda(local)
    pairs = [(ORBextractor(NF, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B, device=local),
              ORBextractor(NF, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B, device=local)) for _ in range(2)]

    def enq(i):
        eL, eR = pairs[i & 1]
        eL.extract_batch_async(Ld); eR.extract_batch_async(Rd)
        _lib.check(L_.cuorb_stereo_match_batch_async(eL.handle, eR.handle, MB, MBF))

    steps = 24
    dt = _resident_loop(pairs, enq, lambda: [h.sync() for p in pairs for h in p], steps)
    eL, eR = pairs[0]
    eL.set_profiling(True)
    acc = {}
    for _ in range(3):
        eR.extract_batch_async(Rd); eR.sync()
        eL.extract_batch_async(Ld); eL.sync()
        _lib.check(L_.cuorb_stereo_match_batch_async(eL.handle, eR.handle, MB, MBF)); eL.sync()
        for k, v in eL.stage_ms().items():
            acc.setdefault(k, []).append(v)
    eL.set_profiling(False)
    st = {k: float(np.median(v)) for k, v in acc.items() if k != "h2d"}
    # candidate descriptor pairs the matcher has to score: right keypoints in the row band of a left
    # keypoint that pass the octave and disparity predicates (src/Frame.cc:961-1017), counted on the host for
    # the first pairs of the batch from the fetched keypoints
    cl, _, kl, _ = eL.fetch(); cr, _, kr, _ = eR.fetch()
    sf = eL.GetScaleFactors()
    maxD = np.float32(MBF) / np.float32(MB)
    tot, nimg = 0, min(B, 4)
    for b in range(nimg):
        a, r_ = kl[b, :cl[b]], kr[b, :cr[b]]
        rr = np.float32(2.0) * sf[r_["octave"]]
        lo, hi = np.floor(r_["y"] - rr), np.ceil(r_["y"] + rr)
        vL = np.floor(a["y"])[:, None]
        band = (vL >= lo[None, :]) & (vL <= hi[None, :])
        octv = np.abs(a["octave"][:, None] - r_["octave"][None, :]) <= 1
        disp = (r_["x"][None, :] >= (a["x"] - maxD)[:, None]) & (r_["x"][None, :] <= a["x"][:, None])
        tot += int((band & octv & disp).sum())
    pairs_per_img = tot / nimg
    u = np.zeros((B, eL.capacity), np.float32); d = np.zeros_like(u)
    _lib.check(L_.cuorb_stereo_fetch(eL.handle, _lib.ptr(u), _lib.ptr(d), eL.capacity, None, None, None))
    res = {"workload": "cfg3: 1241x376 stereo pairs, 2000 features; 2x operator() + ComputeStereoMatches, frames resident in HBM",
           "pairs_per_step": B, "stereo_pairs_per_s": B * steps / dt, "frames_per_s": B * steps / dt,
           "ms_per_step": 1e3 * dt / steps, "stage_ms_left_handle": st, "stage_gbs": _stage_gbs(st, W, H, B),
           "keypoints_per_image": float(cl.mean()), "stereo_matches_per_pair": float((u >= 0).sum() / B),
           "stereo_candidate_desc_pairs_per_image": pairs_per_img,
           "stereo_stage_candidate_pairs_per_s": pairs_per_img * B / (st["stereo"] * 1e-3) if st.get("stereo") else None,
           "stereo_stage_popc_per_s": 8 * pairs_per_img * B / (st["stereo"] * 1e-3) if st.get("stereo") else None}
    res["stage_frac_of_hbm_peak"] = {k: v / peak for k, v in res["stage_gbs"].items()}
    for p in pairs:
        for h in p:
            h.close()
    return res


def cfg4_projection(local, with_cpu=True, M=50000):
    """ORBmatcher::SearchByProjection(F, 50k MapPoints) through the C ABI, host SoA buffers in, indices out."""
    import orbslam3_cpp_fork_b200 as cu
    from orbslam3_cpp_fork_b200 import _lib
    from orbslam3_cpp_fork_b200.matcher import _c
    from oracle import oracle_py as op            # input generator (extracts the frame on the CPU) + CPU side figure
    from tests.matching_cases import extracted_frame, mappoint_case
    L_ = _lib.load()
    e, k, d = extracted_frame()
    sf = e.tables()[0]
    c = mappoint_case(k, d, sf, M, 77 + M, with_uright=True)
    F = cu.Frame(k, d, sf, 752, 480, mvuRight=c["u_right"], mb=0.1, mbf=40.0)
    F.mvpMapPoints[c["bound"] > 0] = 10 ** 9
    F.mvpHasObs[c["bound"] > 0] = 1
    mps = cu.MapPointSoA(c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"], c["depth"],
                         (c["flags"] & 1) > 0, (c["flags"] & 2) > 0, (c["flags"] & 4) > 0)
    fvv, keep_alive = F.view()
    flags = mps.flags()
    arrs = [_c(mps.descriptor, np.uint8), _c(mps.mTrackProjX, np.float32), _c(mps.mTrackProjY, np.float32),
            _c(mps.mTrackProjXR, np.float32), _c(mps.mnTrackScaleLevel, np.int32), _c(mps.mTrackViewCos, np.float32),
            _c(mps.mTrackDepth, np.float32), flags]
    best = np.zeros(M, np.int32)

    def call():
        return _lib.check(L_.cuorb_search_by_projection(local, C.byref(fvv), M, *[_lib.ptr(a) for a in arrs], 3.0, 0, 60.0,
                                                        0.8, _lib.ptr(best)))
    n = call()
    for _ in range(5):
        call()
    ts = []
    for _ in range(60):
        t0 = time.perf_counter(); call(); ts.append(time.perf_counter() - t0)
    st = _stats(ts)
    out = {"workload": f"cfg4: SearchByProjection(F, {M} MapPoints), frame 752x480 / {len(k)} keypoints, th 3, nnratio 0.8; "
                       "wall clock around the C-ABI call, host SoA in, best indices out",
           "matches": int(n),
           "ms_per_call": st, "mappoints_per_s": M / (st["median_ms"] * 1e-3)}
    # the same call with the caller's SoA in PINNED host memory (cuorb_host_alloc): the 2.9 MB upload becomes one DMA
    # at link speed instead of a staged pageable copy - what a caller that keeps its local-map arrays pinned gets
    pinned = []
    for a in arrs:
        ptr_ = L_.cuorb_host_alloc(a.nbytes)
        buf = np.frombuffer((C.c_char * a.nbytes).from_address(ptr_), dtype=a.dtype).reshape(a.shape)
        buf[...] = a
        pinned.append((ptr_, buf))
    best_p = np.zeros(M, np.int32)

    def call_pinned():
        return _lib.check(L_.cuorb_search_by_projection(local, C.byref(fvv), M, *[C.c_void_p(p_) for p_, _ in pinned], 3.0, 0,
                                                        60.0, 0.8, _lib.ptr(best_p)))
    for _ in range(5):
        call_pinned()
    ts = []
    for _ in range(60):
        t0 = time.perf_counter(); call_pinned(); ts.append(time.perf_counter() - t0)
    out["ms_per_call_pinned_inputs"] = _stats(ts)
    out["mappoints_per_s_pinned_inputs"] = M / (out["ms_per_call_pinned_inputs"]["median_ms"] * 1e-3)
    out["pinned_equals_pageable"] = bool((best_p == best).all())
    for p_, _ in pinned:
        L_.cuorb_host_free(p_)
    if with_cpu:
        fv, keep = op.make_frame_view(k, d, c["u_right"], c["bound"], 752, 480, sf, mb=0.1, mbf=40.0)
        t0 = time.perf_counter()
        for _ in range(3):
            n_o, best_o = op.search_by_projection(fv, c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"],
                                                  c["depth"], c["flags"], 3.0, False, 60.0, 0.8)
        out["cpu_port_ms_per_call_1core"] = 1e3 * (time.perf_counter() - t0) / 3
        out["equal_to_cpu"] = bool(n_o == n and (best_o == best).all())
    out["resident_local_map"] = localmap_track(local, with_cpu, M, (e, k, d))
    return out


def localmap_track(local, with_cpu=True, M=50000, frame=None):
    """Tracking::SearchLocalPoints' two loops (isInFrustum over the local map, then SearchByProjection) on a local map that
    stays in HBM (cuorb_localmap_*): per call only the pose, the frame's features and the skip list go up."""
    import orbslam3_cpp_fork_b200 as cu
    from orbslam3_cpp_fork_b200 import _lib
    from oracle import oracle_py as op
    from tests.matching_cases import extracted_frame, localmap_case, localmap_expected
    e, k, d = frame if frame is not None else extracted_frame()
    sf = np.asarray(e.tables()[0], np.float32)
    c = localmap_case(k, d, sf, M, 11)
    cam = _lib.FrustumCamera.from_buffer_copy(bytes(c["cam"]))
    m = cu.LocalMap(M, device=local)
    t0 = time.perf_counter()
    m.update(0, c["P"], c["N"], c["min_d"], c["max_d"], c["desc"], c["bad"], c["has_obs"])
    t_upload = time.perf_counter() - t0
    L_ = _lib.load()
    F = cu.Frame(k, d, sf, 752, 480, mvuRight=c["u_right"], mb=0.1, mbf=c["mbf"])
    F.mvpMapPoints[c["bound"] > 0] = 10 ** 9; F.mvpHasObs[c["bound"] > 0] = 1
    fvv, keep_alive = F.view()
    skip = np.ascontiguousarray(c["skip_idx"], np.int32)
    cap = 4096
    mp = np.zeros(cap, np.int32); ft = np.zeros(cap, np.int32); nin = C.c_int32(0)

    def call():
        return _lib.check(L_.cuorb_localmap_search(m._h, C.byref(cam), C.byref(fvv), 0, None, len(skip), _lib.ptr(skip), 0.5, 1.0,
                                                   0, 50.0, 0.8, None, None, cap, _lib.ptr(mp), _lib.ptr(ft), C.byref(nin)))
    n = call()
    for _ in range(5):
        call()
    ts = []
    for _ in range(60):
        t0 = time.perf_counter(); call(); ts.append(time.perf_counter() - t0)
    st = _stats(ts)
    # the same call with the Frame's features left on the device (cuorb_localmap_search_resident): grid and SoA built there
    import torch
    dev = torch.device("cuda", local)
    dk = torch.from_numpy(np.ascontiguousarray(k).view(np.uint8).copy()).to(dev); dd = torch.from_numpy(np.ascontiguousarray(d)).to(dev)
    du = torch.from_numpy(np.ascontiguousarray(c["u_right"], np.float32)).to(dev)
    torch.cuda.synchronize(dev)
    fvr = _lib.FrameView.from_buffer_copy(bytes(fvv))
    fvr.keys_un = dk.data_ptr(); fvr.descriptors = dd.data_ptr(); fvr.u_right = du.data_ptr()
    mp_r = np.zeros(cap, np.int32); ft_r = np.zeros(cap, np.int32); nin_r = C.c_int32(0)

    def call_resident():
        return _lib.check(L_.cuorb_localmap_search_resident(m._h, C.byref(cam), C.byref(fvr), None, None, None, None, 0, None,
                                                            len(skip), _lib.ptr(skip), 0.5, 1.0, 0, 50.0, 0.8, None, None, cap,
                                                            _lib.ptr(mp_r), _lib.ptr(ft_r), C.byref(nin_r)))
    n_r = call_resident()
    for _ in range(5):
        call_resident()
    ts_r = []
    for _ in range(60):
        t0 = time.perf_counter(); call_resident(); ts_r.append(time.perf_counter() - t0)
    out = {"workload": f"isInFrustum + SearchByProjection over a resident local map of {M} MapPoints, frame 752x480 / {len(k)} "
                       "keypoints, th 1, nnratio 0.8; wall clock around cuorb_localmap_search, (point, feature) pairs out",
           "points_in_view": int(nin.value), "matches": int(n), "ms_per_call": st,
           "mappoints_per_s": M / (st["median_ms"] * 1e-3), "one_time_map_upload_ms": 1e3 * t_upload,
           "frame_left_on_device": {"ms_per_call": _stats(ts_r),
                                    "equal_to_host_frame_call": bool(n_r == n and nin_r.value == nin.value and
                                                                     (mp_r[:n] == mp[:n]).all() and (ft_r[:n] == ft[:n]).all()),
                                    "note": "cuorb_localmap_search_resident: keypoints / descriptors / uRight read in place, "
                                            "PosInGrid + cell CSR built by one kernel; only pose, skip list and bound flags go up"}}
    if with_cpu:
        t0 = time.perf_counter()
        o_in, o_n, o_best = localmap_expected(c, k, d, sf, op.is_in_frustum, op.search_by_projection, 1.0, False, 50.0, 0.8)
        out["cpu_port_ms_per_call_1core"] = 1e3 * (time.perf_counter() - t0)
        hit = np.nonzero(o_best >= 0)[0]
        out["equal_to_cpu"] = bool(o_n == n and o_in == nin.value and (mp[:n] == hit).all() and (ft[:n] == o_best[hit]).all())
    return out
