#!/usr/bin/env python
"""Host<->device copy ceiling of the box, with no kernels at all: N processes, each streaming one bench step's
worth of frames pinned->device (184.8 MB) and one step's results device->pinned (34.9 MB) per iteration on two
streams, for N = 1, 2, 4, 8, pinned vs write-combined input pool, ranks mapped to GPUs {0..N-1} or spread over
the box.  This is the denominator for bench.py's e2e leg (`e2e.frac_of_copy_ceiling`).

usage (one box):  python scripts/h2d_sweep.py --out profiles/r02_h2d_sweep.json     (spawns the ranks itself)
"""
import argparse
import ctypes as C
import glob
import json
import os
import subprocess
import sys
import time

H2D_BYTES = 2 * 256 * 752 * 480          # left + right, 256 pairs
D2H_BYTES = 34_900_000


def cudart():
    import torch  # noqa: F401  (loads libcudart into the process)
    cands = glob.glob(os.path.join(os.path.dirname(os.__file__), "site-packages", "nvidia", "cuda_runtime", "lib", "libcudart.so*"))
    import site
    for sp in site.getsitepackages():
        cands += glob.glob(os.path.join(sp, "nvidia", "cuda_runtime", "lib", "libcudart.so*"))
    cands += ["/usr/local/cuda/lib64/libcudart.so"]
    for c in cands:
        try:
            return C.CDLL(c)
        except OSError:
            continue
    raise RuntimeError("libcudart not found")


def worker(args):
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    ngpu = torch.cuda.device_count()
    dev = rank if args.map == "identity" else (rank * (ngpu // world)) % ngpu
    torch.cuda.set_device(dev)
    rt = cudart()
    rt.cudaHostAlloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t, C.c_uint]
    rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    flags = 0x04 if args.mem == "wc" else 0x00            # cudaHostAllocWriteCombined
    hin, hout = C.c_void_p(), C.c_void_p()
    assert rt.cudaHostAlloc(C.byref(hin), H2D_BYTES, flags) == 0
    assert rt.cudaHostAlloc(C.byref(hout), D2H_BYTES, 0) == 0
    C.memset(hin, 7, H2D_BYTES)
    din = torch.empty(H2D_BYTES, dtype=torch.uint8, device="cuda")
    dout = torch.zeros(D2H_BYTES, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def it():
        assert rt.cudaMemcpyAsync(din.data_ptr(), hin, H2D_BYTES, 1, C.c_void_p(s1.cuda_stream)) == 0
        if args.d2h:
            assert rt.cudaMemcpyAsync(hout, dout.data_ptr(), D2H_BYTES, 2, C.c_void_p(s2.cuda_stream)) == 0

    for _ in range(5):
        it()
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.iters):
        it()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        dt = float(t[0])
        print(json.dumps({"n": world, "map": args.map, "mem": args.mem, "d2h": bool(args.d2h), "iters": args.iters,
                          "ms_per_iter": 1e3 * dt / args.iters,
                          "h2d_gbs_total": world * H2D_BYTES * args.iters / dt / 1e9,
                          "h2d_gbs_per_gpu": H2D_BYTES * args.iters / dt / 1e9,
                          "d2h_gbs_total": (world * D2H_BYTES * args.iters / dt / 1e9) if args.d2h else 0.0,
                          "stereo_frames_per_s_ceiling": world * 256 * args.iters / dt}), flush=True)
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worker", action="store_true")
    ap.add_argument("--map", default="identity")
    ap.add_argument("--mem", default="pinned")
    ap.add_argument("--d2h", type=int, default=1)
    ap.add_argument("--iters", type=int, default=60)
    ap.add_argument("--out", default=None)
    ap.add_argument("--ns", default="1,2,4,8")
    args = ap.parse_args()
    if args.worker:
        return worker(args)
    import torch
    ngpu = torch.cuda.device_count()
    rows = []
    topo = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout
    port = 29650
    for n in [int(x) for x in args.ns.split(",") if int(x) <= ngpu]:
        for mp in (["identity"] if n in (1, ngpu) else ["identity", "spread"]):
            for mem in ("pinned", "wc"):
                for d2h in (1, 0):
                    port += 1
                    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__), "--worker",
                           "--map", mp, "--mem", mem, "--d2h", str(d2h), "--iters", str(args.iters)]
                    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
                    for ln in r.stdout.splitlines():
                        if ln.startswith("{"):
                            rows.append(json.loads(ln)); print(ln, flush=True)
                    if r.returncode != 0:
                        print(r.stderr[-2000:], file=sys.stderr)
    out = {"what": "bare pinned<->device copies, no kernels; one bench step of bytes per iteration per GPU",
           "h2d_bytes_per_iter_per_gpu": H2D_BYTES, "d2h_bytes_per_iter_per_gpu": D2H_BYTES, "gpus_on_box": ngpu,
           "cpu_count": os.cpu_count(), "rows": rows, "nvidia_smi_topo": topo}
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        json.dump(out, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
