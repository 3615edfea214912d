#!/usr/bin/env python
"""Benchmark of the hot path: brute-force (JIT semantics) encode of a 512x512 grayscale image, 8x8 ranges,
16x16 domains at stride 1, 8 isometries -- the configuration BASELINE.json's metric and the north star's
">= 1000x the reference CPU brute-force encode on 512x512" are quoted on.

    python bench.py --gpus N --steps K --warmup W            # our arm (one process per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...   # the reference arm: CPU restatement, host cores

A step = one full encode of one image per rank (N > 1: image-parallel batch of N images, weak scaling, plus the
small NCCL all-gather of the winner records).  One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "fractal-image-compression_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

H = W = 512
RANGE, DOMAIN = 8, 16
N_RANGES = (H // RANGE) * (W // RANGE)
N_POS = (H - DOMAIN) * (W - DOMAIN)           # exclusive lattice bound, jit_bfencoder.py:41-42
MATCHES_PER_IMAGE = N_RANGES * N_POS * 8      # 8.06e9
FLOP_PER_MATCH = 2 * RANGE * RANGE            # 2n flop per candidate match
METRIC = "range-domain candidate matches/sec (512x512 brute-force encode)"
UNIT = "matches/s"


def load_image(rank: int) -> np.ndarray:
    img = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))["lena512"]
    return np.ascontiguousarray(np.roll(img, 8 * rank, axis=0))


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "tflops": p["bf16_tflops"], "tflops_sustained": p["bf16_tflops_sustained"],
                "source": "MEASURED_PEAKS.json (measured)"}
    return {"hbm_gbs": 6650.0, "tflops": 1590.0, "tflops_sustained": 1400.0, "source": "B200_PROFILING.md fallback"}


class ClockSampler:
    """SM clock and throttle reasons while the timed region runs.

    The timed region is a few tens of milliseconds, far shorter than one nvidia-smi poll, so the samples come from
    NVML in a thread of this process (the encode calls release the GIL); nvidia-smi -lms is the fallback.
    """
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []          # nvidia-smi rows
        self.sm, self.reasons = [], set()
        self.mx = None
        self.proc = None
        self.nvml = None
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else index
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.dev, n.NVML_CLOCK_SM)))
        try:
            r = n.nvmlDeviceGetCurrentClocksEventReasons(self.dev)
        except Exception:
            r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev)
        for name, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20),
                          ("hw_thermal_slowdown", 0x40)):
            if r & bit:
                self.reasons.add(name)

    def _poll(self):
        while not self._stop.is_set():
            try:
                self._sample_nvml()
            except Exception:
                break
            time.sleep(0.001)

    def start(self):
        if self.nvml:
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.nvml:
            self._stop.set()
            self.t.join(timeout=2)
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx,
                    "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


def host_cores() -> int:
    """All host threads this process may use (torchrun exports OMP_NUM_THREADS=1, which would hide them)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_sample(img, n_ranges_sample, threads=0):
    """Oracle (CPU restatement of jit_bfencoder.py) on a seeded spread of ranges; returns matches/s."""
    from oracle import oracle as orc
    threads = threads or host_cores()
    rng = np.random.default_rng(1234)
    sel = np.sort(rng.choice(N_RANGES, n_ranges_sample, replace=False))
    # contiguous calls of one range each would serialise; the oracle parallelises over a contiguous slice, so take
    # a contiguous run starting at a seeded offset (per-range cost is constant: no early exit, jit_bfencoder.py:41-59)
    start = int(sel[0]) % (N_RANGES - n_ranges_sample)
    t0 = time.perf_counter()
    orc.bf_encode(img, mode=1, range_begin=start, range_end=start + n_ranges_sample, nthreads=threads)
    dt = time.perf_counter() - t0
    return n_ranges_sample * N_POS * 8 / dt, dt, threads, start


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as orc
    orc.lib()
    img = load_image(0)
    cores = host_cores()
    # one short warm-up is enough for compiled C; it also gives the rate that sizes the per-step sample so that the
    # whole run stays near two minutes whatever --steps is
    _, dt0, _, _ = cpu_sample(img, max(cores, 4))
    rate = max(cores, 4) / max(dt0, 1e-6)                       # ranges per second
    sample = int(min(cores * 8, max(cores, 120.0 * rate / max(args.steps, 1))))
    sample = max(sample, 4)
    t_tot, m_tot = 0.0, 0
    for _ in range(args.steps):
        v, dt, cores, start = cpu_sample(img, sample)
        t_tot += dt
        m_tot += sample * N_POS * 8
    value = m_tot / t_tot
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "bundled image lena_512x512 (grayscale)",
        "config": {"workload": "bruteforce-jit 512x512 r8/d16 stride1 8iso", "sample": f"{sample} ranges per step"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{sample} contiguous ranges x all {N_POS * 8} candidates per step, "
                                   f"C restatement of jit_bfencoder.py (oracle/fic_oracle.c), OpenMP over ranges"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "encodes_per_s": value / MATCHES_PER_IMAGE,
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist

    import fic_b200

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()


    img = load_image(rank)
    h = fic_b200.Handle(local)
    # an explicit (non-default) stream: the library launches on it and torch's events are recorded on it
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0
    d_img = torch.from_numpy(img).to(dev)
    d_idx = torch.zeros((N_RANGES, 4), dtype=torch.int32, device=dev)
    # The images of a batch are independent, so their winner records are exchanged in chunks of G images (SURVEY 8(e):
    # "no collective except the final gather of records"): two chunk buffers, the all-gather of a full chunk overlaps the
    # encodes that fill the other one.  A collective after EVERY image locks all ranks to the pace of the slowest GPU plus
    # the launch jitter of eight processes (measured: 1.11 ms per image on every rank against 0.98-1.07 ms alone).
    G = max(1, int(os.environ.get("FIC_BENCH_GATHER_EVERY", "16")))
    d_so2 = [torch.zeros((G, N_RANGES, 2), dtype=torch.float64, device=dev) for _ in range(2)]
    d_so = d_so2[0][0]
    d_R = torch.zeros((N_RANGES,), dtype=torch.float64, device=dev)
    gathered = [[torch.zeros_like(d_so2[0]) for _ in range(world)] for _ in range(2)] if world > 1 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    flags = fic_b200.BF_EXACT_ONLY if args.exact_only else 0
    cursor = [0, 0]               # chunk being filled, next free slot in it
    last_slot = [0, 0]
    last_gather = [0, 0]          # chunk and image count of the most recent all-gather
    # The gather is enqueued by a helper thread while the main thread is already inside the next encode call (ctypes
    # releases the GIL): the NCCL launch overhead of image k then overlaps the kernels of image k+1 as well.
    import queue
    gq: "queue.Queue" = queue.Queue()
    submitted = [None, None]      # per buffer: (threading.Event set once enqueued, [work handle])

    def gather_worker():
        torch.cuda.set_device(local)
        side = torch.cuda.Stream(device=dev)
        with torch.cuda.stream(side):
            while True:
                item = gq.get()
                if item is None:
                    return
                k, count, ev, box = item
                box.append(dist.all_gather([g[:count] for g in gathered[k]], d_so2[k][:count], async_op=True))
                ev.set()

    worker = None
    if world > 1:
        worker = threading.Thread(target=gather_worker, daemon=True)
        worker.start()

    def wait_buffer(k):
        if submitted[k] is not None:
            ev, box = submitted[k]
            ev.wait()
            box[0].wait()              # stream-level: the encode about to overwrite the buffer waits for its gather
            submitted[k] = None

    def submit_gather(k, count):
        if world > 1 and count > 0 and not os.environ.get("FIC_BENCH_NO_GATHER"):   # (the env switch is a diagnostic)
            ev, box = threading.Event(), []
            submitted[k] = (ev, box)
            last_gather[0], last_gather[1] = k, count
            gq.put((k, count, ev, box))

    def step():
        k, slot = cursor
        if world > 1 and slot == 0:
            wait_buffer(k)        # the chunk about to be refilled has been gathered
        fic_b200.bf_encode_dev(h, d_img.data_ptr(), H, W, d_idx.data_ptr(), d_so2[k][slot].data_ptr(), d_R.data_ptr(),
                               flags=flags, stream=stream)
        last_slot[0], last_slot[1] = k, slot
        if slot + 1 == G:         # the only exchange of the path: winner records
            submit_gather(k, G)
            cursor[0], cursor[1] = k ^ 1, 0
        else:
            cursor[1] = slot + 1

    def drain():
        if world > 1:
            k, slot = cursor
            if slot > 0:          # a partly filled chunk: gather what is there
                wait_buffer(k)
                submit_gather(k, slot)
                cursor[0], cursor[1] = k ^ 1, 0
            wait_buffer(0)
            wait_buffer(1)

    for _ in range(max(args.warmup, 3)):
        step()
    drain()
    torch.cuda.synchronize()
    d_so = d_so2[last_slot[0]][last_slot[1]]   # the records of the last warm-up image
    if world > 1 and last_gather[1] and not os.environ.get("FIC_BENCH_NO_GATHER"):
        k, count = last_gather    # the exchange delivered this rank's records unchanged, and every rank's are there
        assert torch.equal(gathered[k][rank][:count], d_so2[k][:count]), "record gather corrupted this rank's records"
        assert all(bool(g[:count].abs().sum() > 0) for g in gathered[k]), "record gather left a rank's records empty"

    # parity spot check against the reference's own records (tests/golden, rank 0 image only)
    parity = "skipped"
    if rank == 0:
        g = np.load(os.path.join(ROOT, "tests", "golden", "bf_golden.npz"))
        idx = d_idx.cpu().numpy(); so = d_so.cpu().numpy(); R = d_R.cpu().numpy()
        ok = all(idx[k].tolist() == [int(v) for v in rec[:4]] and so[k, 0] == rec[4] and so[k, 1] == rec[5]
                 and R[k] == rec[6] for k, rec in zip(g["jit512_sel"], g["jit512_rec"]))
        parity = "bit-exact on golden ranges" if ok else "MISMATCH"

    # ---- device-resident timing -----------------------------------------------------------------
    sampler = ClockSampler(local)
    h.reset_counters()
    h.profile(True)
    h.profile_read()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    sampler.start()
    tail = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
    for a, b in evs:
        flush.zero_()
        a.record()
        step()
        b.record()
    tail[0].record()
    drain()               # the gather of the last chunk is timed in full
    tail[1].record()
    barrier()
    clocks = sampler.stop()
    total_ms = sum(a.elapsed_time(b) for a, b in evs) + tail[0].elapsed_time(tail[1])
    kern_ms, kern_launches = h.profile_read()
    h.profile(False)
    cnt = h.counters()
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    per_rank_ms = [total_ms / args.steps]
    if world > 1:
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        per_rank_ms = [float(x.item()) / args.steps for x in allt]
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * args.steps * MATCHES_PER_IMAGE / (total_ms * 1e-3)

    # ---- end to end through the C ABI with HOST buffers (pinned input) ----------------------------
    pin = torch.from_numpy(img).pin_memory()
    h_img = pin.numpy()
    for _ in range(2):
        fic_b200.bf_encode(h_img, mode=fic_b200.BF_MODE_JIT, flags=flags, handle=h)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fic_b200.bf_encode(h_img, mode=fic_b200.BF_MODE_JIT, flags=flags, handle=h)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    e2e_value = world * args.steps * MATCHES_PER_IMAGE / e2e_s

    if rank == 0:
        pk = peaks()
        launches_per_step = cnt["launches"] / args.steps
        roof = None
        if kern_launches:
            per_launch_s = kern_ms * 1e-3 / kern_launches
            matches_per_launch = MATCHES_PER_IMAGE * args.steps / kern_launches
            achieved = matches_per_launch * FLOP_PER_MATCH / per_launch_s / 1e12
            screen_path = not (args.exact_only or not cnt["screen_ranges"])
            traffic = None
            ncu_json = os.path.join(ROOT, "profiles", "r1_screen_ncu.json")
            if screen_path and os.path.exists(ncu_json):   # dram bytes per launch from the committed ncu --set full capture
                traffic = json.load(open(ncu_json)).get("traffic_bytes_per_launch_avg")
            roof = {"bound": "tensor", "achieved": achieved, "peak": pk["tflops"], "unit": "TFLOP/s",
                    "frac": achieved / pk["tflops"], "traffic": traffic,
                    "algorithmic": "2*64 flop per candidate match (SURVEY 8(d)); the kernel issues K=80 MMAs, i.e. 1.25x that",
                    "launches_per_step": kern_launches / args.steps,
                    "kernel": "bf_scan8_kernel (exact CUDA-core scan)" if args.exact_only or not cnt["screen_ranges"]
                    else "bf_screen_kernel (tcgen05 correlation GEMM + screen epilogue)",
                    "kernel_ms_per_launch": per_launch_s * 1e3, "kernel_share_of_step": kern_ms / total_ms,
                    "peak_source": pk["source"] + ", burst figure (sub-ms kernel timed alone with events)"}
        cpu = None
        if world == 1 and not args.no_cpu:
            from oracle import oracle as orc
            cores = host_cores()
            sample = max(24 * cores, 64)   # about 10-20 s of host work
            cpu_sample(img, max(cores, 4))
            v, dt, cores, start = cpu_sample(img, sample)
            cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"{sample} contiguous ranges from range {start} x all {N_POS * 8} candidates "
                             f"({dt:.1f} s), C restatement of jit_bfencoder.py, OpenMP over ranges"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "bundled image lena_512x512 (grayscale), rolled by 8*rank rows per rank",
            "config": {"workload": "bruteforce-jit 512x512 r8/d16 stride1 8iso, one image per GPU",
                       "l2": "256 MiB buffer rewritten between timed steps", "path": "exact-only" if args.exact_only else "screen+exact",
                       "exchange": f"winner records all-gathered over NCCL every {G} images, inside the timed region" if world > 1 else "none (one GPU)",
                       "parity": parity},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": H * W,
                    "d2h_bytes_per_step": N_RANGES * (16 + 16 + 8)},
            "gpu_launches": cnt["launches"], "gpu_launches_per_step": launches_per_step,
            "encodes_per_s": value / MATCHES_PER_IMAGE, "e2e_encodes_per_s": e2e_value / MATCHES_PER_IMAGE,
            "screen": {"ranges_screened": cnt["screen_ranges"], "ranges_rescanned": cnt["rescan_ranges"],
                       "survivors": cnt["survivors"]},
            "clocks": clocks, "ms_per_step_by_rank": per_rank_ms,
        }
        if roof:
            line["roofline"] = roof
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        gq.put(None)
        worker.join(timeout=5)
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--exact-only", action="store_true", help="skip the tensor-core screen (debug)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
