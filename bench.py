#!/usr/bin/env python
"""bench.py — ORB extract + stereo match throughput (BASELINE.json metric) on N B200s.

A "step" = one batch of synthetic 752x480 stereo pairs through the hot path: left + right
ORBextractor::operator() (1000 features, 8 levels, 1.2, FAST 20/7) and Frame::ComputeStereoMatches.
  value : frames/s with the frames already resident in HBM (results stay on the device)
  e2e   : frames/s through the host-facing C ABI: pinned host frames -> H2D -> kernels -> D2H of
          keypoints, descriptors, mvuRight, mvDepth, every step, double-buffered over two handle pairs
  roofline : the dominant kernel stage (CUDA events on the launching stream) against measured HBM
  cpu_baseline : the CPU oracle (port of the reference path) on the host cores, bounded sample
--impl reference times that CPU path alone (rank 0 only).
Frames are sharded across ranks with no collective (each rank: own GPU, own handles, own streams).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, NFEAT = 752, 480, 1000
FX, MBF = 458.654, 47.90639384423901  # EuRoC stereo: Camera.fx, Camera.bf (Examples_old/Stereo/EuRoC.yaml:40)
MB = MBF / FX
METRIC = "ORB extract+match frames/s @752x480,1000kp"
N_FRAMES = 100_000   # cfg5: the sweep runs over frame ids 0 .. 99,999 of the generator, sharded round-robin over ranks


def level_sizes(w, h, nlevels=8, sf=1.2):
    s = np.float32(1.0)
    out = []
    for _ in range(nlevels):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
        s = np.float32(float(s) * float(np.float32(sf)))
    return out


def algorithmic_bytes(w, h):
    """per-image algorithmic bytes of the three pixel stages."""
    lv = level_sizes(w, h)
    P = [a * b for a, b in lv]
    B = [(a + 38) * (b + 38) for a, b in lv]
    return dict(pyramid=sum(P[:-1]) + P[0] + sum(B), blur=2 * sum(P), fast=2 * sum(P))


def _host_pair(frame_id):
    from orbslam3_cpp_fork_b200 import synth_cb
    return synth_cb.stereo_pair(frame_id, W, H)


def make_pool(n_pairs, first_id=0):
    """n_pairs distinct synthetic stereo pairs on the HOST (frames first_id .. of the counter-based generator,
    orbslam3_cpp_fork_b200/synth_cb.py — the same frames the device generator produces), one process per core."""
    from concurrent.futures import ProcessPoolExecutor
    ids = [(first_id + i) % N_FRAMES for i in range(n_pairs)]
    if n_pairs <= 2:
        res = [_host_pair(f) for f in ids]
    else:
        import multiprocessing as mp
        with ProcessPoolExecutor(max_workers=min(n_pairs, os.cpu_count() or 1), mp_context=mp.get_context("spawn")) as ex:
            res = list(ex.map(_host_pair, ids, chunksize=max(1, n_pairs // (4 * (os.cpu_count() or 1)))))
    return np.stack([r[0] for r in res]), np.stack([r[1] for r in res])


class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "20"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=lambda: [self.lines.append(l) for l in self.proc.stdout], daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
                try:
                    pw.append(float(f[3]))
                except ValueError:
                    pass
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "sm_mhz_min": min(sm) if sm else None,
                "sm_mhz_first_last": [sm[0], sm[-1]] if sm else None, "power_w_max": max(pw) if pw else None,
                "power_w_median": float(np.median(pw)) if pw else None}


def cpu_module():
    """(module, kind, label): oracle/_ref (the reference's own ORBextractor.cc + ComputeStereoMatches, compiled
    unmodified over scalar restated OpenCV primitives) when it is there, else the oracle port."""
    from oracle import ref_py
    if ref_py.available():
        try:
            ref_py.lib()
            return ref_py, "reference", ("oracle/_ref: unmodified src/ORBextractor.cc + Frame::ComputeStereoMatches, -O2, "
                                         "OpenCV calls served by scalar restated primitives (no SIMD/IPP)")
        except Exception:
            pass
    from oracle import oracle_py
    return oracle_py, "port", "CPU oracle -O2 (restatement)"


def cpu_baseline(n_threads, frames_per_thread=12):
    """The reference path on the host cores (oracle/_ref, else the oracle port); bounded sample."""
    op, kind, label = cpu_module()
    n = max(4, n_threads * frames_per_thread)
    L, R = make_pool(min(n, 48))
    reps = (n + len(L) - 1) // len(L)
    L = np.concatenate([L] * reps)[:n]; R = np.concatenate([R] * reps)[:n]
    op.bench_stereo(L[:n_threads], R[:n_threads], MB, MBF, n_threads)  # warm-up
    sec, kps, matches = op.bench_stereo(L, R, MB, MBF, n_threads)
    return n / sec, n, kind, label


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    op, kind, label = cpu_module()
    per_step = max(4, 2 * cores)
    L, R = make_pool(min(per_step, 48))
    reps = (per_step + len(L) - 1) // len(L)
    L = np.concatenate([L] * reps)[:per_step]; R = np.concatenate([R] * reps)[:per_step]
    for _ in range(args.warmup):
        op.bench_stereo(L, R, MB, MBF, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        op.bench_stereo(L, R, MB, MBF, cores)
    sec = time.perf_counter() - t0
    fps = per_step * args.steps / sec
    sample = f"{per_step} synthetic 752x480 stereo pairs per step x {args.steps} steps, {cores} std::threads (one frame each); {label}"
    # W2 of the round-1 verdict: the reference links SIMD OpenCV; its pixel stages alone through the real cv2
    simd = None
    try:
        import bench_extras
        simd = bench_extras.simd_primitives(L[0], reps=2, threads=cores)
        simd1 = bench_extras.simd_primitives(L[0], reps=6, threads=1)
        simd["ms_per_image_1core"] = simd1.get("ms_per_image")
    except Exception as ex:
        simd = {"unavailable": str(ex)}
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sec / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "cfg5: 752x480 stereo pairs, 1000 features, 8 levels, 1.2, FAST 20/7; "
                               "2x ORBextractor::operator() + Frame::ComputeStereoMatches per frame",
                   "frames_per_step": per_step},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample,
                         "simd_primitives": simd,
                         "note": "value = the reference's own control code over SCALAR pixel primitives; simd_primitives = the "
                                 "pixel stages of one image (1 of the 2 per stereo frame) through the real SIMD cv2, i.e. an "
                                 "upper bound on what the library the reference links would do for those stages alone"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def preproc_leg(local, nb, peak, with_cpu):
    """SURVEY.md 8f N1 side figure: cv::remap (rectification) + BGR->gray of nb 752x480 frames in one
    k_preproc launch, frames resident in HBM; wall clock around the synchronous C-ABI call."""
    import torch
    from orbslam3_cpp_fork_b200 import FramePreprocessor, _lib, preproc, synth
    L_ = _lib.load()
    mx, my = synth.rectify_maps(W, H)
    base = synth.color_frames(4, H, W, 3, 7)
    frames = np.concatenate([base] * (nb // 4 + 1))[:nb]
    d_src = torch.from_numpy(frames).cuda(local)
    d_dst = torch.empty((nb, H, W), dtype=torch.uint8, device=f"cuda:{local}")
    pre = FramePreprocessor((W, H), preproc.PRE_REMAP, maps=(mx, my), max_batch=nb, device=local)

    def run():
        _lib.check(L_.cuorb_preprocess(pre.handle, nb, ctypes_ptr(d_src.data_ptr()), W * 3, W * H * 3, _lib.MEM_DEVICE,
                                       preproc.PIX_BGR, ctypes_ptr(d_dst.data_ptr()), _lib.MEM_DEVICE))
    for _ in range(3):
        run()
    reps = 20
    t0 = time.perf_counter()
    for _ in range(reps):
        run()
    dt = (time.perf_counter() - t0) / reps
    alg = nb * (W * H * 3 + W * H + 8 * W * H / nb)        # raw frame in, gray out, the two float maps once
    out = {"workload": f"cv::remap(INTER_LINEAR) + BGR2GRAY, {nb} x 752x480x3 resident", "frames_per_s": nb / dt,
           "ms_per_launch": dt * 1e3, "achieved_gbs": alg / dt / 1e9, "frac_of_hbm_peak": alg / dt / 1e9 / peak}
    if with_cpu:
        from oracle import oracle_py as op
        t0 = time.perf_counter()
        for i in range(8):
            op.preprocess(frames[i % len(frames)], op.PIX_BGR, op.PRE_REMAP, mx, my)
        out["cpu_port_frames_per_s_1core"] = 8 / (time.perf_counter() - t0)
    pre.close()
    return out


def ctypes_ptr(addr):
    import ctypes
    return ctypes.c_void_p(int(addr))


def bind_to_gpu_numa(local):
    """Best effort: run this rank (and first-touch its pinned buffers) on the CPUs nvidia-smi reports as
    local to its GPU, so host->device copies do not cross sockets when 8 ranks feed 8 GPUs."""
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        hdr = None
        for ln in out.splitlines():
            cols = [c.strip() for c in ln.split("\t") if c.strip()]
            if not cols:
                continue
            if hdr is None and any("CPU Affinity" in c for c in cols):
                hdr = cols
                continue
            if hdr is not None and cols[0] == f"GPU{local}":
                aff = cols[1 + [i for i, c in enumerate(hdr) if "CPU Affinity" in c][0]]
                cpus = set()
                for part in aff.split(","):
                    a, _, b = part.partition("-")
                    cpus.update(range(int(a), int(b or a) + 1))
                cpus &= os.sched_getaffinity(0)
                if cpus:
                    os.sched_setaffinity(0, cpus)
                return sorted(cpus)[:1] and f"{aff}"
    except Exception:
        pass
    return None


def run_cuorb(args):
    import torch
    import torch.distributed as dist
    from orbslam3_cpp_fork_b200 import ORBextractor, _lib
    L_ = _lib.load()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libcuorb has no CPU fallback (use --impl reference for the CPU arm)")
    # rank -> GPU: when the box shows more GPUs than ranks, spread the ranks over the box (GPU r * ngpu/N): on this
    # pool GPUs 0-3 share host-side PCIe bandwidth (profiles/r02_h2d_sweep.json: 4 ranks on GPUs 0-3 copy 94 GB/s,
    # on GPUs 0,2,4,6 154 GB/s), which is all the e2e leg is bound by.  --gpu-map identity turns it off.
    ngpu = torch.cuda.device_count()
    if args.gpu_map == "spread" and world > 1 and ngpu > world and ngpu % world == 0:
        local = local * (ngpu // world)
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa(local) if world > 1 else None
    # stdout carries exactly one JSON line: libraries that print there from C (NCCL's version banner at
    # communicator creation) are sent to stderr by pointing fd 1 at fd 2 for the whole run; the line
    # itself is written to the saved descriptor at the end.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B = args.batch
    in_bytes = 2 * B * W * H                    # > L2 (126 MB) at the default batch: no L2 flush needed
    # cfg5: every step works on DISTINCT frames, generated on the device before the timed region
    # (cuorb_synth_stereo, csrc/synth.cu).  Step i of rank r takes frames ((i * world + r) * B ...) mod 100k.
    from orbslam3_cpp_fork_b200 import synth_cb
    n_distinct = max(2, min(args.steps + 3, int(args.gen_gb * 1e9 // in_bytes)))
    pool_dev = []
    scratch = torch.empty((B, H, W), dtype=torch.uint8, device=f"cuda:{local}")
    t_gen = time.perf_counter()
    for i in range(n_distinct):
        l_ = torch.empty((B, H, W), dtype=torch.uint8, device=f"cuda:{local}")
        r_ = torch.empty((B, H, W), dtype=torch.uint8, device=f"cuda:{local}")
        synth_cb.stereo_batch_device(((i * world + rank) * B) % N_FRAMES, B, W, H, device=local, out=(l_, r_, scratch))
        pool_dev.append((l_, r_))
    torch.cuda.synchronize()
    t_gen = time.perf_counter() - t_gen
    del scratch
    # spot certification inside the run (the 1,000-pair certification is tests/test_gpu_synth.py)
    cert_ids = [(rank * B) % N_FRAMES, (rank * B + B - 1) % N_FRAMES]
    cert_ok = True
    for j, fid in zip((0, B - 1), cert_ids):
        hl, hr = _host_pair(fid)
        cert_ok &= bool((pool_dev[0][0][j].cpu().numpy() == hl).all() and (pool_dev[0][1][j].cpu().numpy() == hr).all())
    if not cert_ok:
        raise SystemExit("bench.py: device-generated frames differ from the host definition (synth_cb.py)")
    # --- handles: args.pairs pairs in flight on the e2e path; the first two also serve the resident path
    NP = max(2, args.pairs)
    all_pairs = [(ORBextractor(NFEAT, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B, device=local),
                  ORBextractor(NFEAT, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B, device=local))
                 for _ in range(NP)]
    pairs = all_pairs[:2]
    cap = pairs[0][0].capacity
    Ld, Rd = pool_dev[0]

    import ctypes

    def pinned(shape, dtype):
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = L_.cuorb_host_alloc(n)
        if not p:
            raise RuntimeError(_lib.last_error())
        buf = (ctypes.c_char * n).from_address(p)
        return np.frombuffer(buf, dtype=dtype).reshape(shape)

    outs = []
    for _ in range(NP):
        outs.append(dict(
            kl=pinned((B, cap), _lib.KP_DTYPE), dl=pinned((B, cap, 32), np.uint8), nl=pinned((B,), np.int32), ml=pinned((B,), np.int32),
            kr=pinned((B, cap), _lib.KP_DTYPE), dr=pinned((B, cap, 32), np.uint8), nr=pinned((B,), np.int32), mr=pinned((B,), np.int32),
            u=pinned((B, cap), np.float32), d=pinned((B, cap), np.float32)))
    d2h_bytes = sum(v.nbytes for v in outs[0].values())

    def enqueue_step(eL, eR, o, srcL, srcR, d2h):
        """one step on one handle pair: 2 x operator() + ComputeStereoMatches (+ every result back to pinned host)"""
        eL.extract_batch_async(srcL); eR.extract_batch_async(srcR)
        if d2h:
            eR.fetch(o["kr"], o["dr"], o["nr"], o["mr"], sync=False)
        _lib.check(L_.cuorb_stereo_match_batch_async(eL.handle, eR.handle, MB, MBF))
        if d2h:
            eL.fetch(o["kl"], o["dl"], o["nl"], o["ml"], sync=False)
            _lib.check(L_.cuorb_stereo_fetch_async(eL.handle, _lib.ptr(o["u"]), _lib.ptr(o["d"]), cap))

    step_no = [0]

    def step_resident(eL, eR, o=None):
        l_, r_ = pool_dev[step_no[0] % n_distinct]
        step_no[0] += 1
        enqueue_step(eL, eR, o, l_, r_, o is not None)

    # ---------------- value: device-resident, CUDA events on the launching (left) stream ----------
    # Steps alternate between the two handle pairs (as in the e2e leg), so step i+1's extraction runs
    # beside step i's quadtree / describe / stereo kernels; within a pair the library orders the right
    # handle's next extraction after the stereo match that reads its buffers.
    eL, eR = pairs[0]
    sL = torch.cuda.ExternalStream(eL.stream, device=local)
    others = [torch.cuda.ExternalStream(hd.stream, device=local) for p in pairs for hd in p][1:]
    def timed_resident(steps, with_d2h):
        """device-timed loop: frames resident in HBM; with_d2h copies every result (keypoints, descriptors, counts,
        mvuRight, mvDepth of both eyes) back to pinned host memory each step, as SURVEY 8(d) asks"""
        for i in range(max(args.warmup, 2)):
            step_resident(*pairs[i & 1], outs[i & 1] if with_d2h else None)
        for p in pairs:
            p[0].sync(); p[1].sync()
        barrier()
        l0 = L_.cuorb_kernel_launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(sL)                  # every stream is idle here (barrier above)
        for i in range(steps):
            step_resident(*pairs[i & 1], outs[i & 1] if with_d2h else None)
        for st in others:               # the timed region ends when all four handle streams (kernels and copies) are done
            evj = torch.cuda.Event()
            evj.record(st)
            sL.wait_event(evj)
        ev1.record(sL)
        for p in pairs:
            p[0].sync(); p[1].sync()
        barrier()
        return ev0.elapsed_time(ev1), L_.cuorb_kernel_launch_count() - l0

    # clocks / throttle reasons: rank 0's GPU, sampled every 20 ms over BOTH resident legs (one nvidia-smi process per
    # job: eight of them starting at once deliver no sample inside a 165 ms leg)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)                  # nvidia-smi needs a moment before its first line
    ms_nod2h, _ = timed_resident(args.steps, False)
    ms_resident, launches = timed_resident(args.steps, True)
    clocks = sampler.stop() if sampler else None

    # ---------------- per-stage times (one profiled step, CUDA events inside libcuorb) -------------
    eL.set_profiling(True); eR.set_profiling(True)
    stage_acc = {}
    for _ in range(3):
        # stages are timed with the two handles SERIALISED (no overlap between the two streams),
        # so each stage's CUDA-event time is that kernel sequence alone on the GPU
        eR.extract_batch_async(Rd); eR.sync()
        eL.extract_batch_async(Ld); eL.sync()
        _lib.check(L_.cuorb_stereo_match_batch_async(eL.handle, eR.handle, MB, MBF))
        eL.sync()
        for k, v in eL.stage_ms().items():
            stage_acc.setdefault(k, []).append(v)
    eL.set_profiling(False); eR.set_profiling(False)
    stage_ms = {k: float(np.median(v)) for k, v in stage_acc.items()}   # left handle: B images (+ stereo)

    # ---------------- e2e: pinned host frames in, all results out, every step ----------------------
    # host side of the e2e leg: a pinned ring of distinct batches (copied down from the device pool)
    n_host = max(2, min(n_distinct, args.host_batches))
    host_pool = []
    for i in range(n_host):
        hl_, hr_ = pinned((B, H, W), np.uint8), pinned((B, H, W), np.uint8)
        hl_[:] = pool_dev[i][0].cpu().numpy(); hr_[:] = pool_dev[i][1].cpu().numpy()
        host_pool.append((hl_, hr_))

    def step_e2e(i):
        eL, eR = all_pairs[i % NP]
        eL.sync(); eR.sync()                         # buffers of step i-NP are free again
        hl_, hr_ = host_pool[i % n_host]
        enqueue_step(eL, eR, outs[i % NP], hl_, hr_, True)

    for i in range(max(args.warmup, NP)):
        step_e2e(i)
    for p in all_pairs:
        p[0].sync(); p[1].sync()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step_e2e(i)
    for p in all_pairs:
        p[0].sync(); p[1].sync()
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    # a truncated candidate list / node pool would be a silently wrong batch: ask every handle
    for p in all_pairs:
        for hdl in p:
            _lib.check(L_.cuorb_extractor_status(hdl.handle))
    total_kp = int(outs[0]["nl"].sum() + outs[0]["nr"].sum())
    total_matches = int((outs[0]["u"] >= 0).sum())

    # ---------------- max over ranks ----------------
    if world > 1:
        t = torch.tensor([ms_resident, t_e2e, ms_nod2h], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_resident, t_e2e, ms_nod2h = float(t[0]), float(t[1]), float(t[2])
        ln = torch.tensor([launches], dtype=torch.int64, device="cuda")
        dist.all_reduce(ln, op=dist.ReduceOp.SUM)
        launches = int(ln[0])

    if rank == 0:
        frames = B * args.steps * world
        value = frames / (ms_resident * 1e-3)
        e2e = frames / t_e2e
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        ab = algorithmic_bytes(W, H)
        pix = {k: stage_ms[k] for k in ("pyramid", "blur", "fast")}
        dom = max(pix, key=pix.get)
        dom_all = max({k: v for k, v in stage_ms.items() if k != "h2d"}, key=lambda k: stage_ms[k])
        ach = ab[dom] * B / (stage_ms[dom] * 1e-3) / 1e9
        stage_gbs = {k: ab[k] * B / (stage_ms[k] * 1e-3) / 1e9 for k in pix}
        # DRAM traffic of the dominant kernel from the committed `ncu --set full` capture (per image)
        traffic = None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            if dom in tj:
                traffic = tj[dom]["dram_bytes_per_image"] * B
        except Exception:
            pass
        pipe = None
        try:
            pipe = json.load(open(os.path.join(ROOT, "profiles", "pipe_util.json"))).get(dom_all)
        except Exception:
            pass
        # the roofline that binds FAST: instruction issue / the integer ALU pipe, not HBM.  Instruction counts are the
        # committed `ncu --set full` capture's (profiles/pipe_util.json); the kernel time is this run's CUDA-event time.
        alu = None
        try:
            pu = json.load(open(os.path.join(ROOT, "profiles", "pipe_util.json")))["fast"]
            sm_clock = (clocks.get("sm_mhz") or peaks.get("sm_max_mhz") or 1965.0) * 1e6
            n_sm = torch.cuda.get_device_properties(local).multi_processor_count
            winst = pu["warp_inst_per_image"] * B
            t_fast = stage_ms["fast"] * 1e-3
            issue_peak = n_sm * 4 * sm_clock                     # 4 schedulers per SM, 1 warp-instruction per clock each
            alu_peak = n_sm * 2 * sm_clock                       # the ALU pipe takes 2 warp-instructions per clock per SM
            px = sum(a * b for a, b in level_sizes(W, H))
            alu = {"bound": "issue slots / integer ALU pipe", "kernel": pu["kernel"],
                   "warp_inst_per_launch": winst, "thread_inst_per_pixel": pu["warp_inst_per_image"] * pu["thread_inst_per_inst"] / px,
                   "achieved_warp_inst_per_s": winst / t_fast, "issue_peak_warp_inst_per_s": issue_peak,
                   "frac_of_issue_peak": winst / t_fast / issue_peak,
                   "alu_pipe_warp_inst_per_launch": pu["alu_pipe_pct_of_peak"] / 100.0 * alu_peak * pu["time_us"] * 1e-6,
                   "frac_of_alu_pipe_peak_ncu": pu["alu_pipe_pct_of_peak"] / 100.0,
                   "source": pu["source"] + "; time = this run's CUDA-event stage time"}
        except Exception:
            pass
        cores = os.cpu_count() or 1
        with_cpu = not args.no_cpu_baseline and world == 1      # the CPU baseline is an N=1 figure
        cpu_fps, cpu_n, cpu_kind, cpu_label = cpu_baseline(cores) if with_cpu else (None, 0, None, None)
        # the GPU-side extras run after this rank's handles are gone (they allocate their own arenas)
        for p_ in all_pairs:
            for hd in p_:
                hd.close()
        del Ld, Rd
        pool_dev.clear()
        torch.cuda.empty_cache()
        n1 = preproc_leg(local, min(B, 64), peak, with_cpu) if world == 1 else None
        extra = {}
        if world == 1 and not args.no_extras:
            import bench_extras as bx
            for name, fn in (("cfg1_single_frame_latency", lambda: bx.cfg1_latency(local, with_cpu)),
                             ("cfg2_640x480_batched", lambda: bx.cfg2_batched(local, peak)),
                             ("cfg3_kitti_stereo", lambda: bx.cfg3_kitti(local, peak)),
                             ("cfg4_projection_50k", lambda: bx.cfg4_projection(local, with_cpu))):
                try:
                    extra[name] = fn()
                except Exception as ex:             # an extra must never take the headline line down with it
                    extra[name] = {"error": f"{type(ex).__name__}: {ex}"}
                torch.cuda.empty_cache()
            try:
                extra["matcher_pipe_utilisation"] = json.load(open(os.path.join(ROOT, "profiles", "pipe_util.json")))
            except Exception:
                pass
        copy_ceiling = None
        try:
            sw = json.load(open(os.path.join(ROOT, "profiles", "r02_h2d_sweep.json")))
            rows = [r for r in sw["rows"] if r["n"] == world and r["mem"] == "pinned" and r["d2h"]]
            if rows:
                copy_ceiling = max(r["stereo_frames_per_s_ceiling"] for r in rows)
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_resident / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "cfg5: 752x480 stereo pairs, 1000 features, 8 levels, 1.2, FAST 20/7; "
                                   "2x ORBextractor::operator() + Frame::ComputeStereoMatches per frame",
                       "frames_per_step_per_gpu": B, "parallelism": f"frame-sharded x{world}, no collective",
                       "gpu_of_rank0": local, "gpu_map": args.gpu_map,
                       "cpu_affinity_rank0": numa,
                       "l2": f"inputs {in_bytes / 1e6:.0f} MB + pyramids per step exceed the 126 MB L2 (no flush needed)",
                       "value_leg": "frames resident in HBM; keypoints, descriptors, counts, mvuRight, mvDepth of every frame "
                                    "copied back to pinned host memory inside the timed region (SURVEY 8d)",
                       "e2e_pairs_in_flight": NP,
                       "frames": {"generator": "counter-based synth_cb (numpy definition) == cuorb_synth_stereo (device), frame ids "
                                               "0..99,999 sharded round-robin over ranks",
                                  "distinct_pairs_per_gpu_in_this_run": n_distinct * B, "steps_cycle_over_batches": n_distinct,
                                  "generated_on_device_in_s": t_gen, "host_pinned_batches_e2e": n_host,
                                  "host_certified_in_this_run": cert_ids,
                                  "certification": "tests/test_gpu_synth.py::test_certified_subset_1000_pairs (every 100th id)"},
                       "keypoints_per_frame": total_kp / B, "stereo_matches_per_frame": total_matches / B},
            "clocks": clocks,
            "value_results_left_in_hbm": frames / (ms_nod2h * 1e-3),
            "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": in_bytes * world,
                    "d2h_bytes_per_step": d2h_bytes * world, "ms_per_step": 1e3 * t_e2e / args.steps,
                    "h2d_gbs": in_bytes * world / (t_e2e / args.steps) / 1e9,
                    "copy_ceiling_frames_per_s": copy_ceiling,
                    "frac_of_copy_ceiling": (e2e / copy_ceiling) if copy_ceiling else None,
                    "copy_ceiling_source": "profiles/r02_h2d_sweep.json: bare pinned->device + device->pinned copies of one "
                                           "step's bytes per GPU, no kernels, same N (scripts/h2d_sweep.py)"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                         "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                         "note": "stage times: CUDA events on the launching stream, handles serialised; FAST is bound by the "
                                 "integer ALU pipe (min/max network), not HBM - see DESIGN.md section 6",
                         "stage_ms_per_step": stage_ms, "stage_gbs": stage_gbs,
                         "dominant_stage_overall": dom_all, "dominant_stage_pipe_utilisation": pipe, "alu": alu,
                         "algorithmic_bytes_per_image": ab},
            "cpu_baseline": None if cpu_fps is None else {
                "value": cpu_fps, "unit": "frames/s", "cores": cores, "kind": cpu_kind,
                "sample": f"{cpu_n} synthetic 752x480 stereo pairs, {cores} std::threads (one frame each); {cpu_label}"},
            "n1_frame_preparation": n1,
            "extra": extra,
        }
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()
    sys.stdout.flush()
    os.dup2(json_fd, 1)
    os.close(json_fd)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuorb", choices=["cuorb", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="stereo pairs per step per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the cfg1-cfg4 side measurements (N=1 only)")
    ap.add_argument("--pairs", type=int, default=3, help="handle pairs in flight on the e2e leg")
    ap.add_argument("--gpu-map", default="spread", choices=["spread", "identity"],
                    help="rank -> GPU when the box has more GPUs than ranks")
    ap.add_argument("--gen-gb", type=float, default=40.0, help="HBM budget for device-generated distinct input batches")
    ap.add_argument("--host-batches", type=int, default=8, help="distinct batches in the pinned host ring of the e2e leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "cuorb" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuorb(args)


if __name__ == "__main__":
    main()
