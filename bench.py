#!/usr/bin/env python
"""bench.py — gain-map hot path throughput on B200 (BASELINE.json metric: "Gain-map generate/apply
Mpix/s (4K, 1-8 B200) and % of HBM BW peak").

A step = one pass of the hot path over one batch of synthetic 4K image pairs on every GPU:
generateGainMap (P010 HLG BT.2100 + YUV420 BT.709 -> u8 map) followed by applyGainMap (YUV420 +
that map -> RGBA1010102 HLG, max_display_boost = FLT_MAX), i.e. BASELINE config 5 (whose
single-image form is configs 1-2 at 3840x2160). Images are independent, so N GPUs = N ranks, each
with its own batch, no collective on the data path (weak scaling).

  value      whole-job Mpix/s (full-resolution pixels through generate+apply), inputs resident
             in HBM, CUDA events on the launch stream, max over ranks.
  e2e        the same metric through the host-pointer C ABI (what the reference-side shim calls):
             pinned HOST planes in, host map + host RGBA1010102 out, copies inside the timing.
  roofline   the dominant kernel (apply: 5.5625 algorithmic B/px) against MEASURED_PEAKS.json.
  cpu_baseline  the reference's own CPU code (oracle/_ref, built from /root/reference by
             oracle/Makefile) on this box's host cores, bounded sample. Checker-side only.

`--impl reference` times that CPU code alone (rank 0 only) on the same config/metric.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

W, H = 3840, 2160
MW, MH = W // 4, H // 4
PIX = W * H
FLT_MAX = 3.4028234663852886e38
GEN_BPP = 4.5625    # P010 3 B + 420 1.5 B read, 1/16 B written
APPLY_BPP = 5.5625  # 420 1.5 B + map 1/16 B read, RGBA1010102 4 B written
METRIC = "gain-map generate+apply throughput (4K)"
UNIT = "Mpix/s"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------ CPU arms
def cpu_worker(n_images, seed, tag):
    """Child process: the reference CPU path (<=4 threads, ultrahdr.cpp:304,500) over n_images
    4K pairs: generate (HLG, 2100->709) + apply -> RGBA1010102 HLG. Prints elapsed seconds."""
    import numpy as np

    from libultrahdr_dev_b200 import abi, synth
    from oracle.pyoracle import CpuImpl

    impl = CpuImpl.reference() if tag == "reference" else CpuImpl.oracle()
    sdr_b, hdr_b = make_pair(seed)  # the pixels of image `seed` of the GPU arm's job
    sdr = abi.host_yuv420(sdr_b, W, H, abi.COLORGAMUT_BT709)
    hdr = abi.host_p010(hdr_b, W, H, abi.COLORGAMUT_BT2100)
    sys.stdout.write("ready\n")
    sys.stdout.flush()
    sys.stdin.readline()
    t0 = time.perf_counter()
    for _ in range(n_images):
        st, mp, md, _ = impl.generate_gainmap(sdr, hdr, abi.TF_HLG)
        assert st == 0
        st, out, _ = impl.apply_gainmap(sdr, abi.host_map(mp, MW, MH), md, abi.OUTPUT_HDR_HLG, FLT_MAX)
        assert st == 0
    sys.stdout.write(json.dumps({"s": time.perf_counter() - t0, "sum": int(out[:4096].sum())}) + "\n")
    sys.stdout.flush()


def make_pair(index):
    """Image `index` of the job: seeded synthetic natural pair, identical in both bench arms."""
    from libultrahdr_dev_b200 import abi, synth

    return synth.natural_pair(W, H, int(index), abi.TF_HLG, abi.COLORGAMUT_BT709, abi.COLORGAMUT_BT2100)


def make_pairs(indices, procs):
    """Generates the pairs in parallel worker processes (numpy, ~3.5 s per 4K pair per core). Called before CUDA is up."""
    if procs <= 1 or len(indices) <= 1:
        return [make_pair(i) for i in indices]
    import multiprocessing as mp

    with mp.get_context("spawn").Pool(min(procs, len(indices))) as pool:
        return pool.map(make_pair, list(indices))


def cpu_arm(images_per_worker, workers=None):
    """Runs `workers` concurrent reference processes; returns (Mpix/s, info dict)."""
    from oracle.pyoracle import CpuImpl

    tag = "reference" if CpuImpl.have_reference() else "port"
    cores = host_cores()
    threads_per = min(cores, 4) if tag == "reference" else 1
    if workers is None:
        workers = max(1, cores // threads_per)
    procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--cpu-worker", str(images_per_worker),
                               str(i), tag], stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True, cwd=ROOT)
             for i in range(workers)]  # worker i runs on image i of the job (the GPU arm's pair i)
    for p in procs:
        assert p.stdout.readline().strip() == "ready"
    t0 = time.perf_counter()
    for p in procs:
        p.stdin.write("go\n")
        p.stdin.flush()
    for p in procs:
        json.loads(p.stdout.readline())
    dt = time.perf_counter() - t0
    for p in procs:
        p.wait()
    n = workers * images_per_worker
    return n * PIX / dt / 1e6, {
        "kind": tag, "cores": workers * threads_per, "host_cores": cores,
        "sample": f"{n} passes over the job's first {workers} synthetic natural 3840x2160 pairs (the GPU arm's pixels; generate HLG 2100->709 + apply->RGBA1010102 HLG), "
                  f"{workers} concurrent processes x {threads_per} threads (reference caps a call at 4 threads), {dt:.1f} s",
        "seconds": dt, "images": n}


# ------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock, power and throttle reasons of one GPU every ~5 ms through NVML (falls back to
    `nvidia-smi -lms`) while the timed region runs; window(t0, t1) summarises the samples inside it."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.stop_flag, self.max_mhz = [], None, False, None
        try:
            import pynvml

            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nv = pynvml
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
            return
        except Exception:
            self.nv = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _poll(self):
        nv = self.nv
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8,
                "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        while not self.stop_flag:
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                self.rows.append((time.perf_counter(), [str(mhz), str(self.max_mhz), str(pw)] +
                                  ["Active" if r & bits[k] else "Not Active" for k in ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")]))
            except Exception:
                pass
            time.sleep(0.005)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def window(self, t0, t1):
        rows = [r for t, r in self.rows if t0 <= t <= t1] or [r for _, r in self.rows[-3:]]
        sm = sorted(float(r[0]) for r in rows if r and r[0].replace(".", "").isdigit())
        pw = sorted(float(r[2]) for r in rows if len(r) > 2 and r[2].replace(".", "").isdigit())
        reasons = set()
        for r in rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        mx = [float(r[1]) for r in rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None,
                "power_w_median": pw[len(pw) // 2] if pw else None, "reasons": sorted(reasons), "samples": len(rows)}

    def stop(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()


# ------------------------------------------------------------------------------------ GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="4K pairs per step per GPU (device-resident; weak scaling)")
    ap.add_argument("--job-images", type=int, default=0,
                    help="BASELINE config 5 as written: a FIXED job of this many 4K pairs (1024) image-sharded over the ranks "
                         "(strong scaling: rank r owns images r, r+N, ...); 0 = the default weak-scaling batch per GPU")
    ap.add_argument("--distinct", type=int, default=0, help="distinct seeded pairs materialised per rank (default 64 at N=1, 16 per rank at N>1); "
                                                          "the other device slots hold cyclic shifts of them, so no two slots are byte-identical")
    ap.add_argument("--e2e-images", type=int, default=64, help="4K pairs per e2e step per GPU (host planes)")
    ap.add_argument("--e2e-threads", type=int, default=4, help="host threads calling the synchronous host-pointer API concurrently")
    ap.add_argument("--e2e-join", default="auto", choices=["auto", "spin", "blocking"], help="how the host entry points wait (uhdrb200_set_sync_mode)")
    ap.add_argument("--e2e-host-memory", default="pageable", choices=["pinned", "pageable"],
                    help="host planes of the e2e leg: uhdrb200_host_alloc (page-locked) or plain malloc'ed numpy buffers")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-stages", action="store_true", help="skip the per-kernel rooflines of the other stages")
    ap.add_argument("--cpu-images", type=int, default=24, help="images per CPU worker process in the CPU arms")
    ap.add_argument("--lib", default=None, help="experiments: path of a variant libuhdr_b200.so")
    ap.add_argument("--no-e2e", action="store_true", help="experiments: skip the end-to-end leg (the line then has no valid e2e)")
    ap.add_argument("--profile-once", action="store_true", help="one launch per other-stage kernel, no warm-up (for ncu captures; timings meaningless)")
    ap.add_argument("--cpu-worker", nargs=3, metavar=("N", "SEED", "TAG"))
    a = ap.parse_args()
    if a.cpu_worker:
        return cpu_worker(int(a.cpu_worker[0]), int(a.cpu_worker[1]), a.cpu_worker[2])

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if a.job_images:
        a.batch = len(range(rank, a.job_images, world))  # shard_round_robin
    config = {"workload": (f"BASELINE config 5 as written: a fixed job of {a.job_images} synthetic 3840x2160 pairs, image i on rank i mod N (strong scaling), "
                           if a.job_images else "BASELINE config 5 flavour: a batch of synthetic 3840x2160 pairs per GPU (weak scaling), ") +
                          "P010(HLG,BT.2100)+YUV420(BT.709), generateGainMap then applyGainMap->RGBA1010102 HLG (max_display_boost=FLT_MAX), image-sharded",
              "width": W, "height": H, "map_scale": 4, "images_per_step_per_gpu": a.batch, "job_images": a.job_images or None,
              "l2": "inputs+outputs per step >> 126 MB L2 (no flush needed)", "mode": "analytic (reference as built)"}

    if a.impl == "reference":
        if rank != 0:
            return
        steps = max(1, a.steps)
        vals = []
        for _ in range(a.warmup and 1):
            cpu_arm(1)
        info = None
        t_all = 0.0
        n_all = 0
        for _ in range(steps):
            v, info = cpu_arm(a.cpu_images)
            vals.append(v)
            t_all += info["seconds"]
            n_all += info["images"]
            if t_all > 150:
                break
        value = n_all * PIX / t_all / 1e6
        info.update(value=value, unit=UNIT)
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus,
                          "steps": len(vals), "warmup": 1 if a.warmup else 0, "ms_per_step": 1e3 * t_all / len(vals),
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                          "data": "synthetic", "config": config, "cpu_baseline": info, "gpu_launches": 0,
                          "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import numpy as np

    # the distinct seeded pairs of this rank, generated in worker processes before CUDA is initialised
    K = a.distinct or (64 if world == 1 else 16)
    K = max(1, min(K, a.batch))
    my_first = list(range(rank, (a.job_images or world * a.batch), world))[:K]
    pairs = make_pairs(my_first, max(1, host_cores() // world))

    import torch

    from libultrahdr_dev_b200 import abi, api, refdata, synth

    lib = api.load_library(a.lib)  # raises when the CUDA library is missing: no CPU fallback
    if not torch.cuda.is_available() or lib.uhdrb200_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    uh = api.JpegR(lib)
    assert lib.uhdrb200_init(local) == 0, lib.uhdrb200_last_error()

    # N > 1: keep this rank's host threads (and the library's staging helpers) on the cores next to its GPU
    numa_cores = None
    if world > 1:
        try:
            import pynvml

            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[local]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else local
            ncpu = os.cpu_count() or 1
            words = pynvml.nvmlDeviceGetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(phys), (ncpu + 63) // 64)
            near = {i for i in range(ncpu) if (words[i // 64] >> (i % 64)) & 1} & os.sched_getaffinity(0)
            if near:
                os.sched_setaffinity(0, near)
                numa_cores = len(near)
        except Exception:
            numa_cores = None

    # ---- CPU baseline first (rank 0, N=1 only), before the GPU is busy
    cpu_info = None
    if world == 1 and not a.no_cpu_baseline:
        v, cpu_info = cpu_arm(a.cpu_images)
        cpu_info.update(value=v, unit=UNIT)

    # ---- synthetic inputs. The job is J images (world*B, or --job-images); image i belongs to rank i mod world
    # (shard.py). A rank materialises K distinct seeded pairs (seed = the global index of its first K images,
    # generated BEFORE this point in worker processes) and fills its B device slots with cyclic shifts of them
    # (slot j = pair j mod K rolled by 64 * (j div K) pixels in x and 36 * (j div K) in y, SDR and HDR alike), so that
    # every slot holds different bytes.
    from libultrahdr_dev_b200 import shard

    B = a.batch
    mine = shard.shard_round_robin(a.job_images or world * B, rank, world)
    assert len(mine.indices) == B
    K = len(pairs)
    sdr_len, hdr_len = abi.yuv420_size(W, H), abi.p010_size(W, H)
    d_sdr = torch.empty((B, sdr_len), dtype=torch.uint8, device=dev)
    d_hdr = torch.empty((B, hdr_len), dtype=torch.int16, device=dev)
    d_map = torch.zeros((B, MW * MH), dtype=torch.uint8, device=dev)
    d_out = torch.zeros((B, PIX), dtype=torch.int32, device=dev)

    def roll420(dst, src, dy, dx):  # packed I420: Y | U | V
        for off, pw, ph_, sc in ((0, W, H, 1), (W * H, W // 2, H // 2, 2), (W * H * 5 // 4, W // 2, H // 2, 2)):
            dst[off:off + pw * ph_].view(ph_, pw).copy_(torch.roll(src[off:off + pw * ph_].view(ph_, pw), (dy // sc, dx // sc), (0, 1)))

    def roll_p010(dst, src, dy, dx):  # Y plane | interleaved UV plane (dx even keeps the (u, v) pairing)
        dst[:W * H].view(H, W).copy_(torch.roll(src[:W * H].view(H, W), (dy, dx), (0, 1)))
        dst[W * H:].view(H // 2, W).copy_(torch.roll(src[W * H:].view(H // 2, W), (dy // 2, dx), (0, 1)))

    for i in range(min(K, B)):
        s = torch.from_numpy(pairs[i][0]).to(dev)
        h = torch.from_numpy(pairs[i][1].view(np.int16)).to(dev)
        for j in range(i, B, K):
            roll420(d_sdr[j], s, 36 * (j // K), 64 * (j // K))
            roll_p010(d_hdr[j], h, 36 * (j // K), 64 * (j // K))
    torch.cuda.synchronize()

    Img = abi.Image * B
    sdrs, hdrs, maps, outs = Img(), Img(), Img(), Img()
    mds = (abi.Metadata * B)()
    for i in range(B):
        for arr, base, fmt, gam in ((sdrs, d_sdr[i].data_ptr(), abi.PIX_FMT_YUV420, abi.COLORGAMUT_BT709),):
            arr[i].data, arr[i].chroma_data = base, base + W * H
            arr[i].width, arr[i].height, arr[i].colorGamut = W, H, gam
            arr[i].luma_stride, arr[i].chroma_stride, arr[i].pixelFormat = W, W // 2, fmt
        hb = d_hdr[i].data_ptr()
        hdrs[i].data, hdrs[i].chroma_data = hb, hb + 2 * W * H
        hdrs[i].width, hdrs[i].height, hdrs[i].colorGamut = W, H, abi.COLORGAMUT_BT2100
        hdrs[i].luma_stride, hdrs[i].chroma_stride, hdrs[i].pixelFormat = W, W, abi.PIX_FMT_P010
        maps[i].data = d_map[i].data_ptr()
        maps[i].width, maps[i].height, maps[i].luma_stride = MW, MH, MW
        maps[i].pixelFormat = abi.PIX_FMT_MONOCHROME
        outs[i].data = d_out[i].data_ptr()
    md = abi.default_metadata(1000.0 / 203.0)
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        assert uh.generateGainMapBatch(B, sdrs, hdrs, abi.TF_HLG, mds, maps, False, stream) == 0
        assert uh.applyGainMapBatch(B, sdrs, maps, md, abi.OUTPUT_HDR_HLG, FLT_MAX, outs, stream) == 0

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(a.warmup, 3)):
        step()
    barrier()
    clocks = ClockSampler(local) if rank == 0 else None
    time.sleep(0.25)
    n0 = lib.uhdrb200_launch_count()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(a.steps)]
    barrier()
    t0 = time.perf_counter()
    for k in range(a.steps):
        ev[k][0].record()
        assert uh.generateGainMapBatch(B, sdrs, hdrs, abi.TF_HLG, mds, maps, False, stream) == 0
        ev[k][1].record()
        assert uh.applyGainMapBatch(B, sdrs, maps, md, abi.OUTPUT_HDR_HLG, FLT_MAX, outs, stream) == 0
        ev[k][2].record()
    barrier()
    t1 = time.perf_counter()
    launches = lib.uhdrb200_launch_count() - n0
    total_ms = ev[0][0].elapsed_time(ev[-1][2])
    gen_ms = sum(e[0].elapsed_time(e[1]) for e in ev) / a.steps
    app_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / a.steps
    clk = clocks.window(t0, t1) if clocks else None

    # ---- the other kernels of the path (not part of `value`): per-kernel roofline on the same batch, rank 0 only
    other = {}
    jpeg_stage = jpeg_enc_stage = jpeg_flow_stage = None
    if rank == 0 and not a.no_other_stages:
        Bo = min(B, 64)
        d_tm = torch.zeros((Bo, sdr_len), dtype=torch.uint8, device=dev)
        tms = (abi.Image * Bo)()
        for i in range(Bo):
            tb = d_tm[i].data_ptr()
            tms[i].data, tms[i].chroma_data = tb, tb + W * H
            tms[i].width, tms[i].height, tms[i].colorGamut = W, H, abi.COLORGAMUT_UNSPECIFIED
            tms[i].luma_stride, tms[i].chroma_stride, tms[i].pixelFormat = W, W // 2, abi.PIX_FMT_YUV420
        d_big = torch.zeros((Bo, PIX * 8), dtype=torch.uint8, device=dev)
        bigs = (abi.Image * Bo)()
        for i in range(Bo):
            bigs[i].data = d_big[i].data_ptr()
        md_pq = abi.default_metadata(10000.0 / 203.0)

        def timed(fn, bpp, iters=8):
            if a.profile_once:
                iters = 1
            for _ in range(0 if a.profile_once else 2):
                assert fn() == 0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / iters
            gbs = Bo * PIX * bpp / (ms * 1e-3) / 1e9
            return {"mpix_s": Bo * PIX / ms / 1e3, "achieved_gbs": gbs, "algorithmic_bytes_per_px": bpp, "ms_per_launch": ms,
                    "images_per_launch": Bo}

        other["tonemap"] = timed(lambda: uh.toneMapBatch(Bo, hdrs, tms, stream), 4.5)
        other["api0_fused_tonemap_generate"] = timed(lambda: uh.toneMapGenerateGainMapBatch(Bo, hdrs, abi.TF_HLG, tms, mds, maps, stream), 4.5625)
        other["api0_fused_tonemap_generate_convert601"] = timed(lambda: uh.toneMapGenerateGainMapBatch(Bo, hdrs, abi.TF_HLG, tms, mds, maps, stream, sdr_to_601=True), 4.5625)
        other["generate_pq_p3"] = timed(lambda: uh.generateGainMapBatch(Bo, sdrs, hdrs, abi.TF_PQ, mds, maps, False, stream), GEN_BPP)
        assert uh.generateGainMapBatch(Bo, sdrs, hdrs, abi.TF_HLG, mds, maps, False, stream) == 0  # restore HLG maps
        other["apply_rgba_f16_linear"] = timed(lambda: uh.applyGainMapBatch(Bo, sdrs, maps, md, abi.OUTPUT_HDR_LINEAR, FLT_MAX, bigs, stream), 9.5625)
        other["apply_1010102_pq"] = timed(lambda: uh.applyGainMapBatch(Bo, sdrs, maps, md_pq, abi.OUTPUT_HDR_PQ, FLT_MAX, bigs, stream), APPLY_BPP)
        # an SDR-like display (max_display_boost = 1 < maxContentBoost): the gain factor maxContentBoost^(1/maxContentBoost)
        # exceeds 1, codes can pass 1023 and wrap through the reference's `& 0x3ff` (gainmapmath.cpp:722-727): the masked
        # packing path. From max_display_boost ~1.5 upwards the factor stays below 1 and the unmasked kernel runs.
        other["apply_1010102_hlg_display_boost_1"] = timed(lambda: uh.applyGainMapBatch(Bo, sdrs, maps, md, abi.OUTPUT_HDR_HLG, 1.0, bigs, stream), APPLY_BPP)
        other["apply_planar_10bit"] = timed(lambda: uh.applyGainMapBatch(Bo, sdrs, maps, md, abi.OUTPUT_HDR_LINEAR_RGB_10BIT, FLT_MAX, bigs, stream), 7.5625)
        other["sdr_rgba8888_pack"] = timed(lambda: uh.yuv420ToRgba8888Batch(Bo, sdrs, bigs, stream), 5.5)
        # one image per call through the device-pointer entry points (the reference API's own granularity): kernel time of
        # a single 4K image, images rotated so that L2 never holds the next one. Rates are per launch, not per batch.
        def timed_single(fn, bpp, iters=64):
            if a.profile_once:
                iters = 1
            for i in range(0 if a.profile_once else 4):
                assert fn(i) == 0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(iters):
                fn(i % Bo)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / iters
            return {"mpix_s": PIX / ms / 1e3, "achieved_gbs": PIX * bpp / (ms * 1e-3) / 1e9, "algorithmic_bytes_per_px": bpp,
                    "ms_per_launch": ms, "images_per_launch": 1}

        md1 = abi.Metadata()
        other["generate_single_image_call"] = timed_single(
            lambda i: uh.generateGainMapDevice(sdrs[i], hdrs[i], abi.TF_HLG, md1, maps[i], False, stream), GEN_BPP)
        other["apply_1010102_hlg_single_image_call"] = timed_single(
            lambda i: uh.applyGainMapDevice(sdrs[i], maps[i], md, abi.OUTPUT_HDR_HLG, FLT_MAX, bigs[i], stream), APPLY_BPP)
        # N1 convertYuv in place (jpegr.cpp:1132-1206), one launch per image: four 8160x6144 images (75 MB each) in rotation
        cw_, ch_ = 8160, 6144
        d_cv = torch.empty((4, cw_ * ch_ * 3 // 2), dtype=torch.uint8, device=dev)
        for i in range(4):  # the natural 4K SDR planes of image i, tiled 3 x 3 and cropped to 8160 x 6144
            src = d_sdr[i % K]
            for off_s, off_d, pw, ph_, tw, th in ((0, 0, W, H, cw_, ch_), (W * H, cw_ * ch_, W // 2, H // 2, cw_ // 2, ch_ // 2),
                                                  (W * H * 5 // 4, cw_ * ch_ * 5 // 4, W // 2, H // 2, cw_ // 2, ch_ // 2)):
                plane = src[off_s:off_s + pw * ph_].view(ph_, pw).repeat(3, 3)[:th, :tw]
                d_cv[i, off_d:off_d + tw * th] = plane.reshape(-1)
        cvs = (abi.Image * 4)()
        for i in range(4):
            cb = d_cv[i].data_ptr()
            cvs[i].data, cvs[i].chroma_data = cb, cb + cw_ * ch_
            cvs[i].width, cvs[i].height, cvs[i].colorGamut = cw_, ch_, abi.COLORGAMUT_BT709
            cvs[i].luma_stride, cvs[i].chroma_stride, cvs[i].pixelFormat = cw_, cw_ // 2, abi.PIX_FMT_YUV420
        cv_i = [0]

        def cv_fn():
            cv_i[0] += 1
            return uh.convertYuvDevice(cvs[cv_i[0] % 4], abi.COLORGAMUT_BT709, abi.COLORGAMUT_P3, stream)

        r = timed(cv_fn, 3.0, 8)
        scale = (cw_ * ch_) / (Bo * PIX)  # timed() assumes Bo 4K images per launch
        r.update(mpix_s=r["mpix_s"] * scale, achieved_gbs=r["achieved_gbs"] * scale, images_per_launch=1, image="8160x6144")
        other["convert_yuv_709_to_601_50mp"] = r
        del d_cv
        assert lib.uhdrb200_test_set_apply_mode(abi.APPLY_TWO_TABLE) == 0
        other["apply_1010102_hlg_two_table_kernel"] = timed(lambda: uh.applyGainMapBatch(Bo, sdrs, maps, md, abi.OUTPUT_HDR_HLG, FLT_MAX, bigs, stream), APPLY_BPP)
        assert lib.uhdrb200_test_set_apply_mode(abi.APPLY_AUTO) == 0
        # worst case for the q-domain kernel: i.i.d. uniform 8-bit samples (SURVEY 8d "uniform"), 19 % of the channels on
        # srgbInvOetf's linear toe -> almost every quad-row takes the exact (MUFU) branch
        d_rnd = torch.randint(0, 256, (Bo, sdr_len), dtype=torch.uint8, device=dev)
        rnds = (abi.Image * Bo)()
        for i in range(Bo):
            C.memmove(C.byref(rnds[i]), C.byref(sdrs[i]), C.sizeof(abi.Image))
            rb = d_rnd[i].data_ptr()
            rnds[i].data, rnds[i].chroma_data = rb, rb + W * H
        other["apply_1010102_hlg_uniform_random_input"] = timed(lambda: uh.applyGainMapBatch(Bo, rnds, maps, md, abi.OUTPUT_HDR_HLG, FLT_MAX, bigs, stream), APPLY_BPP)
        del d_rnd
        # dark natural scene: the natural SDR planes with luma / 8 and chroma pulled towards neutral -> every quad-row on the
        # srgbInvOetf toe branch, with the spatial coherence of a real image (no table bank conflicts, no divergence)
        d_drk = d_sdr[:Bo].clone()
        d_drk[:, :W * H] //= 8
        d_drk[:, W * H:] = (128 + (d_drk[:, W * H:].to(torch.int16) - 128) // 8).to(torch.uint8)
        drks = (abi.Image * Bo)()
        for i in range(Bo):
            C.memmove(C.byref(drks[i]), C.byref(sdrs[i]), C.sizeof(abi.Image))
            rb = d_drk[i].data_ptr()
            drks[i].data, drks[i].chroma_data = rb, rb + W * H
        other["apply_1010102_hlg_dark_natural_input"] = timed(lambda: uh.applyGainMapBatch(Bo, drks, maps, md, abi.OUTPUT_HDR_HLG, FLT_MAX, bigs, stream), APPLY_BPP)
        del d_drk
        # map scale factor 2 (half-resolution gain map, as other encoders write): the any-scale q-domain kernel
        d_map2 = d_map[:Bo].view(Bo, MH, MW).repeat_interleave(2, 1).repeat_interleave(2, 2).contiguous()
        maps2 = (abi.Image * Bo)()
        for i in range(Bo):
            maps2[i].data = d_map2[i].data_ptr()
            maps2[i].width, maps2[i].height, maps2[i].luma_stride = 2 * MW, 2 * MH, 2 * MW
            maps2[i].pixelFormat = abi.PIX_FMT_MONOCHROME
        other["apply_1010102_hlg_map_scale_2"] = timed(lambda: uh.applyGainMapBatch(Bo, sdrs, maps2, md, abi.OUTPUT_HDR_HLG, FLT_MAX, bigs, stream), 1.5 + 0.25 + 4.0)
        del d_map2
        # map scale factor 8 (an eighth-resolution map): the block-inside-one-cell specialisation of the any-scale kernel
        d_map8 = d_map[:Bo].view(Bo, MH, MW)[:, ::2, ::2].contiguous()
        maps8 = (abi.Image * Bo)()
        for i in range(Bo):
            maps8[i].data = d_map8[i].data_ptr()
            maps8[i].width, maps8[i].height, maps8[i].luma_stride = MW // 2, MH // 2, MW // 2
            maps8[i].pixelFormat = abi.PIX_FMT_MONOCHROME
        other["apply_1010102_hlg_map_scale_8"] = timed(lambda: uh.applyGainMapBatch(Bo, sdrs, maps8, md, abi.OUTPUT_HDR_HLG, FLT_MAX, bigs, stream), 1.5 + 1.0 / 64 + 4.0)
        del d_map8
        d_map1 = d_map[:Bo].view(Bo, MH, MW).repeat_interleave(4, 1).repeat_interleave(4, 2).contiguous()  # full-resolution map
        maps1 = (abi.Image * Bo)()
        for i in range(Bo):
            maps1[i].data = d_map1[i].data_ptr()
            maps1[i].width, maps1[i].height, maps1[i].luma_stride = W, H, W
            maps1[i].pixelFormat = abi.PIX_FMT_MONOCHROME
        other["apply_1010102_hlg_map_scale_1"] = timed(lambda: uh.applyGainMapBatch(Bo, sdrs, maps1, md, abi.OUTPUT_HDR_HLG, FLT_MAX, bigs, stream), 1.5 + 1.0 + 4.0)
        del d_map1
        assert lib.uhdrb200_test_set_apply_mode(abi.APPLY_GENERIC) == 0
        other["apply_1010102_hlg_generic_kernel"] = timed(lambda: uh.applyGainMapBatch(Bo, sdrs, maps, md, abi.OUTPUT_HDR_HLG, FLT_MAX, bigs, stream), APPLY_BPP, 2)
        assert lib.uhdrb200_test_set_apply_mode(abi.APPLY_AUTO) == 0
        # LUT numeric mode (UHDRB200_MODE_LUT, byte-identical to the reference's -DUSE_*_LUT=1 build): the per-pixel kernels
        # with the reference's nearest-entry tables and float order
        assert lib.uhdrb200_set_numeric_mode(abi.MODE_LUT) == 0
        other["lut_mode_generate"] = timed(lambda: uh.generateGainMapBatch(Bo, sdrs, hdrs, abi.TF_HLG, mds, maps, False, stream), GEN_BPP, 4)
        other["lut_mode_apply_1010102_hlg"] = timed(lambda: uh.applyGainMapBatch(Bo, sdrs, maps, md, abi.OUTPUT_HDR_HLG, FLT_MAX, bigs, stream), APPLY_BPP, 2)
        other["lut_mode_apply_rgba_f16_linear"] = timed(lambda: uh.applyGainMapBatch(Bo, sdrs, maps, md, abi.OUTPUT_HDR_LINEAR, FLT_MAX, bigs, stream), 9.5625, 2)
        assert lib.uhdrb200_set_numeric_mode(abi.MODE_ANALYTIC) == 0
        assert uh.generateGainMapBatch(Bo, sdrs, hdrs, abi.TF_HLG, mds, maps, False, stream) == 0  # restore the analytic maps
        # ---- N4: dequantise + ISLOW IDCT + range-limit of 4:2:0 coefficient planes (3 B per sample: 2 in, 1 out): one 8160x6144
        # image per call (150 MB of coefficients), four distinct images in rotation
        jw, jh = 8160, 6144
        jn = jw * jh * 3 // 2
        d_co = (torch.randint(-40, 40, (4, jn), dtype=torch.int16, device=dev) * (torch.rand((4, jn), device=dev) < 0.12).to(torch.int16))
        d_pl = torch.zeros((4, jn), dtype=torch.uint8, device=dev)
        d_q = torch.randint(1, 32, (64,), dtype=torch.int16, device=dev)
        jc = (abi.JpegComponent * 3)()

        def idct_fn(state=[0]):
            i = state[0] % 4
            state[0] += 1
            base, pl = d_co[i].data_ptr(), d_pl[i].data_ptr()
            for k, (off, cw, ch) in enumerate(((0, jw, jh), (jw * jh, jw // 2, jh // 2), (jw * jh * 5 // 4, jw // 2, jh // 2))):
                jc[k].coeffs, jc[k].quant = base + 2 * off, d_q.data_ptr()
                jc[k].blocks_w, jc[k].blocks_h = cw // 8, ch // 8
                jc[k].plane, jc[k].width, jc[k].height, jc[k].stride = pl + off, cw, ch, cw
            return lib.uhdrb200_jpeg_idct_dev(3, jc, stream)

        r = timed(idct_fn, 4.5, 8)  # per image pixel: 1.5 samples x (2 B coefficients + 1 B sample)
        scale = (jw * jh) / (Bo * PIX)  # timed() assumes Bo 4K images per launch
        r.update(mpix_s=r["mpix_s"] * scale, achieved_gbs=r["achieved_gbs"] * scale, images_per_launch=1, image="8160x6144 4:2:0")
        other["jpeg_idct_420_50mp"] = r
        del d_co, d_pl
        # ---- N4, whole decode: 4K 4:2:0 baseline JPEGs (quality 90, libjpeg-turbo's default tables, made here through
        # Pillow from the natural SDR planes + sensor-like noise) -> planes in device memory. Host parse and byte-stuffing
        # removal, upload of the compressed bytes, parallel Huffman decode and IDCT, wall clock per synchronous call; one
        # calling thread and four (each with its own context and stream). Latency-bound, not HBM-bound: no roofline.
        try:
            import io
            import threading

            from PIL import Image

            jp, noisy = [], []
            for i in range(4):
                hp = d_sdr[i % K].cpu().numpy()
                yy = hp[: W * H].reshape(H, W)
                uu = hp[W * H: W * H * 5 // 4].reshape(H // 2, W // 2).repeat(2, 0).repeat(2, 1)
                vv = hp[W * H * 5 // 4:].reshape(H // 2, W // 2).repeat(2, 0).repeat(2, 1)
                rr = np.random.Generator(np.random.PCG64(i))
                yy = np.clip(yy.astype(np.int16) + rr.integers(-6, 7, yy.shape), 0, 255).astype(np.uint8)
                noisy.append(np.concatenate([yy.ravel(), hp[W * H:]]))
                bio = io.BytesIO()
                Image.fromarray(np.stack([yy, uu, vv], -1), "YCbCr").save(bio, "JPEG", quality=90, subsampling=2)
                jp.append(bio.getvalue())
            jbufs = [(C.c_ubyte * len(b)).from_buffer_copy(b) for b in jp]

            def decode_loop(tid, iters, out):
                torch.cuda.set_device(dev)
                d_pl = torch.zeros(W * H * 3 // 2, dtype=torch.uint8, device=dev)
                ts = torch.cuda.Stream(device=dev).cuda_stream  # one stream per calling thread
                torch.cuda.synchronize()
                im = abi.Image()
                im.data = d_pl.data_ptr()
                for k in range(2):
                    assert lib.uhdrb200_jpeg_decode_dev(jbufs[k % 4], len(jp[k % 4]), C.byref(im), ts) == 0, lib.uhdrb200_last_error()
                t0 = time.perf_counter()
                for k in range(iters):
                    lib.uhdrb200_jpeg_decode_dev(jbufs[(k + tid) % 4], len(jp[(k + tid) % 4]), C.byref(im), ts)
                out[tid] = (time.perf_counter() - t0) / iters

            res1 = {}
            decode_loop(0, 12, res1)
            res4 = {}
            ths = [threading.Thread(target=decode_loop, args=(t, 12, res4)) for t in range(4)]
            for t in ths:
                t.start()
            for t in ths:
                t.join()
            t0 = time.perf_counter()
            for k in range(2):
                im = Image.open(io.BytesIO(jp[k]))
                im.draft("YCbCr", (W, H))
                im.load()
            cpu_ms = (time.perf_counter() - t0) / 2 * 1e3
            jpeg_stage = {"mpix_s": PIX / res1[0] / 1e6, "ms_per_image": res1[0] * 1e3, "images_per_launch": 1,
                          "mpix_s_4_host_threads": 4 * PIX / (sum(res4.values()) / 4) / 1e6, "jpeg_bytes": int(np.mean([len(b) for b in jp])),
                          "libjpeg_turbo_1_core_mpix_s": PIX / cpu_ms / 1e3, "libjpeg_turbo_1_core_ms": cpu_ms,
                          "content": "natural SDR planes + noise, quality 90, 4:2:0, through Pillow / libjpeg-turbo's encoder",
                          "bound": "latency (sequential synchronisation chains of the parallel Huffman decode), not HBM"}
            # the encode direction: the same planes, resident in device memory (where API-0 leaves the SDR image), -> JFIF
            # bytes in host memory at quality 90; forward DCT, quantisation and Huffman coding on the device
            d_noisy = torch.stack([torch.from_numpy(x) for x in noisy]).to(dev)
            e_src = (abi.Image * 4)()
            for k in range(4):
                eb = d_noisy[k].data_ptr()
                e_src[k].data, e_src[k].chroma_data = eb, eb + W * H
                e_src[k].width, e_src[k].height, e_src[k].luma_stride, e_src[k].chroma_stride, e_src[k].pixelFormat = W, H, W, W // 2, abi.PIX_FMT_YUV420

            def encode_loop(tid, iters, out):
                torch.cuda.set_device(dev)
                ob = np.zeros(W * H * 3, np.uint8)
                on = C.c_size_t(0)
                ts = torch.cuda.Stream(device=dev).cuda_stream  # one stream per calling thread
                torch.cuda.synchronize()
                for k in range(2):
                    assert lib.uhdrb200_jpeg_encode_dev(C.byref(e_src[k % 4]), 90, None, 0, ob.ctypes.data, ob.size, C.byref(on), ts) == 0, lib.uhdrb200_last_error()
                t0 = time.perf_counter()
                for k in range(iters):
                    lib.uhdrb200_jpeg_encode_dev(C.byref(e_src[(k + tid) % 4]), 90, None, 0, ob.ctypes.data, ob.size, C.byref(on), ts)
                out[tid] = ((time.perf_counter() - t0) / iters, on.value)

            enc1, enc4 = {}, {}
            encode_loop(0, 12, enc1)
            ths = [threading.Thread(target=encode_loop, args=(t, 12, enc4)) for t in range(4)]
            for t in ths:
                t.start()
            for t in ths:
                t.join()
            t0 = time.perf_counter()
            for k in range(2):
                bio = io.BytesIO()
                Image.open(io.BytesIO(jp[k])).save(bio, "JPEG", quality=90, subsampling=2)
            enc_cpu_ms = (time.perf_counter() - t0) / 2 * 1e3 - cpu_ms  # decode + encode minus the decode measured above
            jpeg_enc_stage = {"mpix_s": PIX / enc1[0][0] / 1e6, "ms_per_image": enc1[0][0] * 1e3, "images_per_launch": 1,
                              "mpix_s_4_host_threads": 4 * PIX / (sum(v[0] for v in enc4.values()) / 4) / 1e6, "jpeg_bytes": int(enc1[0][1]),
                              "libjpeg_turbo_1_core_mpix_s": PIX / max(enc_cpu_ms, 1e-3) / 1e3, "libjpeg_turbo_1_core_ms": enc_cpu_ms,
                              "parity": "file byte-identical to libjpeg-turbo's (tests/test_gpu_jpeg_encode.py)"}
            # the pixel side of encodeJPEGR API-0 and of decodeJPEGR (jpegr.cpp:167-261, 693-801) per 4K image and host thread,
            # wall clock: P010 in PAGEABLE host memory -> [upload, toneMap + generateGainMap + convertYuv in one kernel, JPEG
            # encode of the SDR image (quality 95) and of the gain map (85)] -> two byte strings; then those bytes -> [two
            # decodes, applyGainMap -> RGBA1010102 HLG, copy to pageable host memory]
            h_hdr = [d_hdr[i % K].cpu().numpy().copy() for i in range(2)]
            f_hdr = torch.empty(hdr_len, dtype=torch.int16, device=dev)
            f_sdr = torch.zeros(sdr_len, dtype=torch.uint8, device=dev)
            f_map = torch.zeros(MW * MH, dtype=torch.uint8, device=dev)
            f_out = torch.zeros(PIX, dtype=torch.int32, device=dev)
            h_out = np.zeros(PIX, np.int32)
            fh, fs, fm, fo = abi.Image(), abi.Image(), abi.Image(), abi.Image()
            C.memmove(C.byref(fh), C.byref(hdrs[0]), C.sizeof(abi.Image))
            fh.data, fh.chroma_data = f_hdr.data_ptr(), f_hdr.data_ptr() + 2 * W * H
            C.memmove(C.byref(fs), C.byref(sdrs[0]), C.sizeof(abi.Image))
            fs.data, fs.chroma_data, fs.colorGamut = f_sdr.data_ptr(), f_sdr.data_ptr() + W * H, abi.COLORGAMUT_UNSPECIFIED
            fm.data, fo.data = f_map.data_ptr(), f_out.data_ptr()
            fmd = abi.Metadata()
            ob1, ob2 = np.zeros(W * H * 3, np.uint8), np.zeros(W * H, np.uint8)
            n1, n2 = C.c_size_t(0), C.c_size_t(0)

            def encode_flow(i):
                f_hdr.copy_(torch.from_numpy(h_hdr[i % 2]), non_blocking=True)
                assert uh.toneMapGenerateGainMapDevice(fh, abi.TF_HLG, fs, fmd, fm, stream, True) == 0
                assert lib.uhdrb200_jpeg_encode_dev(C.byref(fs), 95, None, 0, ob1.ctypes.data, ob1.size, C.byref(n1), stream) == 0
                assert lib.uhdrb200_jpeg_encode_dev(C.byref(fm), 85, None, 0, ob2.ctypes.data, ob2.size, C.byref(n2), stream) == 0

            def decode_flow(b1, b2, l1, l2):
                ds, dm = abi.Image(), abi.Image()
                ds.data, dm.data = f_sdr.data_ptr(), f_map.data_ptr()
                assert lib.uhdrb200_jpeg_decode_dev(b1, l1, C.byref(ds), stream) == 0
                assert lib.uhdrb200_jpeg_decode_dev(b2, l2, C.byref(dm), stream) == 0
                ds.colorGamut = abi.COLORGAMUT_BT2100
                assert lib.uhdrb200_apply_gainmap_dev(C.byref(ds), C.byref(dm), C.byref(fmd), abi.OUTPUT_HDR_HLG, FLT_MAX, C.byref(fo), stream) == 0
                torch.from_numpy(h_out).copy_(f_out)

            for i in range(2):
                encode_flow(i)
            t0 = time.perf_counter()
            for i in range(8):
                encode_flow(i)
            enc_flow = (time.perf_counter() - t0) / 8
            b1 = (C.c_ubyte * n1.value).from_buffer_copy(ob1[: n1.value].tobytes())
            b2 = (C.c_ubyte * n2.value).from_buffer_copy(ob2[: n2.value].tobytes())
            for i in range(2):
                decode_flow(b1, b2, n1.value, n2.value)
            t0 = time.perf_counter()
            for i in range(8):
                decode_flow(b1, b2, n1.value, n2.value)
            dec_flow = (time.perf_counter() - t0) / 8
            jpeg_flow_stage = {"encode_side_ms_per_image": enc_flow * 1e3, "encode_side_mpix_s": PIX / enc_flow / 1e6,
                               "decode_side_ms_per_image": dec_flow * 1e3, "decode_side_mpix_s": PIX / dec_flow / 1e6,
                               "mpix_s": PIX / (enc_flow + dec_flow) / 1e6, "images_per_launch": 1, "sdr_jpeg_bytes": int(n1.value),
                               "map_jpeg_bytes": int(n2.value), "host_threads": 1,
                               "what": "P010 (pageable host) -> fused toneMap + generateGainMap + convertYuv -> 2 JPEG encodes -> bytes; "
                                       "bytes -> 2 JPEG decodes -> applyGainMap HLG -> pageable host; nothing but P010, JPEG bytes and the "
                                       "final pixels crosses PCIe"}
        except ImportError:
            pass
        # ---- real image content: the reference's own fixtures (tests/data, staged by build() into tests/_ref_data):
        # the 1280x720 raw P010 + YUV420 pair of ProfileGainMapFuncs (tests/jpegr_test.cpp:2250-2304) tiled 3 x 3 to 4K,
        # and the only photograph in the tree (minnie-320x240.yu12) tiled 12 x 9; every slot a different cyclic shift
        if refdata.available():
            fp, fw, fh = refdata.load("p010")
            fy, _, _ = refdata.load("yuv420")
            fm, mw0, mh0 = refdata.load("minnie")
            base_s = torch.from_numpy(refdata.tile_yuv420(fy, fw, fh, W // fw, H // fh)[0]).to(dev)
            base_h = torch.from_numpy(refdata.tile_p010(fp, fw, fh, W // fw, H // fh)[0].view(np.int16)).to(dev)
            base_m = torch.from_numpy(refdata.tile_yuv420(fm, mw0, mh0, W // mw0, H // mh0)[0]).to(dev)
            base_mm = torch.from_numpy(synth.natural_map(MW, MH, 71)).to(dev)
            d_ps = torch.empty((Bo, sdr_len), dtype=torch.uint8, device=dev)
            d_ph = torch.empty((Bo, hdr_len), dtype=torch.int16, device=dev)
            d_pm = torch.empty((Bo, sdr_len), dtype=torch.uint8, device=dev)
            pss, phs, pms = (abi.Image * Bo)(), (abi.Image * Bo)(), (abi.Image * Bo)()
            for j in range(Bo):
                dy, dx = 4 * ((5 * j) % MH), 4 * ((7 * j) % MW)
                roll420(d_ps[j], base_s, dy, dx)
                roll_p010(d_ph[j], base_h, dy, dx)
                roll420(d_pm[j], base_m, dy, dx)
                for arr, t in ((pss, d_ps), (pms, d_pm)):
                    C.memmove(C.byref(arr[j]), C.byref(sdrs[j]), C.sizeof(abi.Image))
                    arr[j].data, arr[j].chroma_data = t[j].data_ptr(), t[j].data_ptr() + W * H
                C.memmove(C.byref(phs[j]), C.byref(hdrs[j]), C.sizeof(abi.Image))
                phs[j].data, phs[j].chroma_data = d_ph[j].data_ptr(), d_ph[j].data_ptr() + 2 * W * H
            other["generate_photo_content"] = timed(lambda: uh.generateGainMapBatch(Bo, pss, phs, abi.TF_HLG, mds, maps, False, stream), GEN_BPP)
            other["generate_photo_content"]["content"] = "raw_p010_image.p010 + raw_yuv420_image.yuv420 (1280x720) tiled 3x3, 64 cyclic shifts"
            other["apply_1010102_hlg_photo_content"] = timed(lambda: uh.applyGainMapBatch(Bo, pss, maps, md, abi.OUTPUT_HDR_HLG, FLT_MAX, bigs, stream), APPLY_BPP)
            other["apply_1010102_hlg_photo_content"]["content"] = "raw_yuv420_image.yuv420 tiled 3x3 with its own generated gain map"
            for j in range(Bo):
                d_map[j].view(MH, MW).copy_(torch.roll(base_mm, ((5 * j) % MH, (7 * j) % MW), (0, 1)))
            other["apply_1010102_hlg_photo_minnie"] = timed(lambda: uh.applyGainMapBatch(Bo, pms, maps, md, abi.OUTPUT_HDR_HLG, FLT_MAX, bigs, stream), APPLY_BPP)
            other["apply_1010102_hlg_photo_minnie"]["content"] = "minnie-320x240.yu12 (photograph) tiled 12x9, smooth synthetic full-range gain map"
            other["apply_rgba_f16_photo_minnie"] = timed(lambda: uh.applyGainMapBatch(Bo, pms, maps, md, abi.OUTPUT_HDR_LINEAR, FLT_MAX, bigs, stream), 9.5625)
            del d_ps, d_ph, d_pm
        del d_tm, d_big

    # ---- e2e: host planes -> C ABI host entry points -> host results (copies inside the timing)
    E = a.e2e_images
    KH = max(4, a.e2e_threads)  # one private set of pinned buffers per in-flight image
    pageable_keep = []

    def make_bufs(kind):
        bufs = []
        for i in range(KH):
            if kind == "pinned":
                ps = lib.uhdrb200_host_alloc(sdr_len)
                ph = lib.uhdrb200_host_alloc(hdr_len * 2)
                po = lib.uhdrb200_host_alloc(PIX * 4)
                pm = lib.uhdrb200_host_alloc(MW * MH)
                assert ps and ph and po and pm, "pinned allocation failed"
            else:  # what an unmodified libultrahdr caller passes: pageable memory
                keep = [np.empty(n, dtype=np.uint8) for n in (sdr_len, hdr_len * 2, PIX * 4, MW * MH)]
                for k in keep:
                    k[:] = 0  # touch the pages
                pageable_keep.append(keep)
                ps, ph, po, pm = (k.ctypes.data for k in keep)
            C.memmove(ps, pairs[i % K][0].ctypes.data, sdr_len)
            C.memmove(ph, pairs[i % K][1].ctypes.data, hdr_len * 2)
            s_img, h_img, m_img, o_img = abi.Image(), abi.Image(), abi.Image(), abi.Image()
            s_img.data, s_img.chroma_data = ps, ps + W * H
            s_img.width, s_img.height, s_img.colorGamut = W, H, abi.COLORGAMUT_BT709
            s_img.luma_stride, s_img.chroma_stride, s_img.pixelFormat = W, W // 2, abi.PIX_FMT_YUV420
            h_img.data, h_img.chroma_data = ph, ph + 2 * W * H
            h_img.width, h_img.height, h_img.colorGamut = W, H, abi.COLORGAMUT_BT2100
            h_img.luma_stride, h_img.chroma_stride, h_img.pixelFormat = W, W, abi.PIX_FMT_P010
            o_img.data = po
            bufs.append((s_img, h_img, m_img, o_img, pm))
        return bufs

    host_bufs = make_bufs(a.e2e_host_memory)

    # Each host thread owns a CUDA stream + staging buffers inside the library (thread-local context), so T
    # threads calling the synchronous reference-style entry points overlap H2D, kernels and D2H of different
    # images — what a multi-threaded server on top of the drop-in API gets. ctypes releases the GIL in the calls.
    import concurrent.futures as cf

    T = max(1, a.e2e_threads)
    # more calling threads (all ranks) than cores: sleeping joins instead of spinning ones (include/uhdr_b200.h)
    blocking = world * (T + 1) > host_cores() if a.e2e_join == "auto" else a.e2e_join == "blocking"
    assert lib.uhdrb200_set_sync_mode(abi.SYNC_BLOCKING if blocking else abi.SYNC_SPIN) == 0
    pool = cf.ThreadPoolExecutor(T, initializer=lambda: lib.uhdrb200_init(local))

    cur_bufs = [host_bufs]

    def e2e_one(i):
        s_img, h_img, m_img, o_img, pm = cur_bufs[0][i % KH]
        mdh, m_loc = abi.Metadata(), abi.Image()
        st = lib.uhdrb200_generate_gainmap_host_into(C.byref(s_img), C.byref(h_img), abi.TF_HLG, C.byref(mdh),
                                                     C.byref(m_loc), 0, pm, MW * MH)
        assert st == 0, lib.uhdrb200_last_error()
        o_loc = abi.Image()
        o_loc.data = o_img.data
        st = lib.uhdrb200_apply_gainmap_host(C.byref(s_img), C.byref(m_loc), C.byref(mdh), abi.OUTPUT_HDR_HLG,
                                             FLT_MAX, C.byref(o_loc))
        assert st == 0, lib.uhdrb200_last_error()

    def e2e_step():
        list(pool.map(e2e_one, range(E)))

    e2e_steps = 0 if a.no_e2e else max(2, min(a.steps, 5))
    if not a.no_e2e:
        e2e_step()
    barrier()
    te0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - te0
    h2d = E * (sdr_len + hdr_len * 2 + sdr_len + MW * MH)  # generate: SDR+HDR planes; apply: SDR planes + map
    d2h = E * (MW * MH + PIX * 4)
    # the same leg with the OTHER kind of host memory, reported beside the headline: page-locked planes
    # (uhdrb200_host_alloc: what a caller that adopts the library's allocator gets) when the headline is pageable
    other_kind = "pinned" if a.e2e_host_memory == "pageable" else "pageable"
    other_mpix = None
    if not a.no_e2e:
        cur_bufs[0] = make_bufs(other_kind)
        e2e_step()
        barrier()
        tp0 = time.perf_counter()
        for _ in range(2):
            e2e_step()
        barrier()
        other_s = time.perf_counter() - tp0
        cur_bufs[0] = host_bufs

    # what the host <-> device links of this box give with ALL ranks copying at once (every N): pinned <-> device copies of
    # 256 MB alone and in both directions together. The e2e rate is judged against these, not against HBM.
    pcie_rates = [0.0, 0.0, 0.0]
    if not a.no_e2e:
        nb = 256 << 20
        h_a = torch.empty(nb, dtype=torch.uint8).pin_memory()
        h_b = torch.empty(nb, dtype=torch.uint8).pin_memory()
        d_a = torch.empty(nb, dtype=torch.uint8, device=dev)
        d_b = torch.empty(nb, dtype=torch.uint8, device=dev)
        s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

        def copy_rate(up, down, reps=4):
            barrier()
            t = time.perf_counter()
            for _ in range(reps):
                if up:
                    with torch.cuda.stream(s1):
                        d_a.copy_(h_a, non_blocking=True)
                if down:
                    with torch.cuda.stream(s2):
                        h_b.copy_(d_b, non_blocking=True)
            barrier()
            return reps * nb / (time.perf_counter() - t) / 1e9

        copy_rate(True, True, 1)
        pcie_rates = [copy_rate(True, False), copy_rate(False, True), copy_rate(True, True)]
        del h_a, h_b, d_a, d_b

    times = torch.tensor([total_ms, e2e_s * 1e3, gen_ms, app_ms, (other_s if not a.no_e2e else 0.0) * 1e3], dtype=torch.float64, device=dev)
    rates = torch.tensor(pcie_rates, dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.all_reduce(rates, op=dist.ReduceOp.SUM)  # every rank timed the same barrier-to-barrier window: the sum is the box aggregate
    total_ms, e2e_ms, gen_ms, app_ms, other_ms = [float(x) for x in times.tolist()]
    pcie = None
    if not a.no_e2e:
        r_up, r_down, r_both = [float(x) for x in rates.tolist()]
        pcie = {"ranks_copying_at_once": world, "h2d_alone_gbs": r_up, "d2h_alone_gbs": r_down, "each_direction_when_both_gbs": r_both}
        other_mpix = world * 2 * E * PIX / (other_ms * 1e-3) / 1e6
    if clocks:
        clocks.stop()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return

    peaks, peak_src = None, "fallback (B200_PROFILING.md)"
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
        peak, peak_src = float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except (OSError, KeyError, ValueError):
        peak = 6650.0
    # DRAM bytes per pixel of each kernel from the committed `ncu --set full` capture of this command
    # (profiles/ncu_traffic.json, written by scripts/ncu_traffic.py from the .ncu-rep raw page).
    traffic = {}
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            traffic = json.load(f)
    except (OSError, ValueError):
        pass

    def traffic_of(kernel):
        t = traffic.get(kernel)
        return t["dram_bytes_per_pixel"] * B * PIX if t else None

    app_gbs = B * PIX * APPLY_BPP / (app_ms * 1e-3) / 1e9
    gen_gbs = B * PIX * GEN_BPP / (gen_ms * 1e-3) / 1e9
    n_job = a.job_images or world * B  # images of the whole job per step
    value = n_job * PIX * a.steps / (total_ms * 1e-3) / 1e6
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
        "ms_per_step": total_ms / a.steps, "higher_is_better": True, "scaling": "strong" if a.job_images else "weak", "vs_baseline": None,
        "dtype": "f32", "data": f"synthetic ({K} distinct seeded natural pairs per rank, cyclically shifted into {B} byte-distinct device slots; image i of the {a.job_images or world * B}-image job on rank i mod {world})",
        "config": config,
        "roofline": {"bound": "hbm", "kernel": "apply_q_kernel (applyGainMap -> RGBA1010102 HLG)", "achieved": app_gbs,
                     "peak": peak, "unit": "GB/s", "frac": app_gbs / peak, "traffic": traffic_of("apply_q_kernel"), "peak_source": peak_src,
                     "ms_per_launch": app_ms, "algorithmic_bytes_per_launch": B * PIX * APPLY_BPP},
        "roofline_generate": {"bound": "hbm", "kernel": "generate_kernel", "achieved": gen_gbs, "peak": peak,
                              "unit": "GB/s", "frac": gen_gbs / peak, "traffic": traffic_of("generate_kernel"), "ms_per_launch": gen_ms,
                              "algorithmic_bytes_per_launch": B * PIX * GEN_BPP},
        "stage_mpix_s": {"generate": B * PIX / gen_ms / 1e3, "apply": B * PIX / app_ms / 1e3},
        "e2e": {"value": world * E * PIX * e2e_steps / (e2e_ms * 1e-3) / 1e6 if e2e_steps else None, "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "images_per_step_per_gpu": E, "steps": e2e_steps,
                "host_threads": T, "numa_local_cores": numa_cores, "join": "blocking" if blocking else "spin", "pcie_measured": pcie, "host_memory": a.e2e_host_memory, other_kind + "_planes_mpix_s": other_mpix,
                "api": "uhdrb200_generate_gainmap_host_into + uhdrb200_apply_gainmap_host (synchronous, reference-style); " +
                       ("pageable host planes, as an unmodified libultrahdr caller passes them" if a.e2e_host_memory == "pageable" else "page-locked host planes")},
        "other_stages": dict({k: dict(v, frac_of_hbm_peak=v["achieved_gbs"] / peak) for k, v in other.items()},
                             **({"jpeg_decode_4k_q90_to_device_planes": jpeg_stage} if jpeg_stage else {}),
                             **({"jpeg_encode_4k_q90_from_device_planes": jpeg_enc_stage} if jpeg_enc_stage else {}),
                             **({"jpegr_pixel_flow_4k_one_host_thread": jpeg_flow_stage} if jpeg_flow_stage else {})),
        "gpu_launches": int(launches),
        "clocks": clk,
    }
    if out["e2e"]["value"]:
        ips = out["e2e"]["value"] * 1e6 / PIX / world  # images per second per GPU
        out["e2e"]["h2d_gbs_per_gpu"] = ips * h2d / E / 1e9
        out["e2e"]["d2h_gbs_per_gpu"] = ips * d2h / E / 1e9
        if pcie and pcie["each_direction_when_both_gbs"] > 0:
            # ceiling of this traffic mix on the measured links: both directions busy, the up direction carries more bytes
            ceil_ips = pcie["each_direction_when_both_gbs"] * 1e9 / max(h2d / E, d2h / E)
            out["e2e"]["link_ceiling_mpix_s"] = ceil_ips * PIX / 1e6
            out["e2e"]["frac_of_link_ceiling"] = out["e2e"]["value"] / out["e2e"]["link_ceiling_mpix_s"]
    if cpu_info:
        out["cpu_baseline"] = cpu_info
    print(json.dumps(out))


if __name__ == "__main__":
    main()
