#!/usr/bin/env python
"""bench.py -- Msamples/s of the path-tracing render core on the Mario scene at 1080p (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--world 5] [--pp 32]
                    [--width 1920 --height 1080] [--shard tiles|samples]

A "step" is one whole frame of the workload: every pixel x pp*pp camera paths through the hot path
(camera rays -> closest hit -> shading -> accumulation -> resolve).  `value` = samples of all ranks
/ device time (scene resident in HBM, CUDA events on the launch stream, max over ranks); `e2e` =
the same metric through the public C-ABI with HOST buffers every step: pt_upload_scene (H2D of the
scene + BVH build) -> pt_render -> pt_readback (D2H of the BGRA frame into pinned memory).
For N > 1 the frame is strong-scaled: the one frame is split over the ranks by interleaved 32x32
tiles (gather to rank 0) or by sample ranges (NCCL reduce of the float accumulation buffers).
--impl reference times the reference's own CPU renderer (oracle/_ref, else the C port) on the host
cores on a bounded sample of the same workload.
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
sys.path.insert(0, ROOT)

# algorithmic flops per sample of the reference algorithm, SURVEY.md section 8d / Appendix D
F_PER_SAMPLE = {1: 620.0, 2: 7300.0, 3: 1030.0, 4: 27300.0, 5: 2140.0}
WORLD_NAMES = {1: "w1 default", 2: "w2 metal-roughness", 3: "w3 cornell", 4: "w4 rtow cover", 5: "w5 mario glb"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--world", type=int, default=5)
    ap.add_argument("--pp", type=int, default=32)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--shard", default="tiles", choices=["tiles", "samples", "tiles_p2p"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kernel", default="auto", choices=["auto", "mega", "wavefront"])
    return ap.parse_args()


def workload_name(a):
    return f"{WORLD_NAMES[a.world]} {a.width}x{a.height} -p{a.pp} ({a.pp * a.pp} spp)"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ---------------------------------------------------------------------------------------------------
# CPU arm: the reference's own renderer (oracle/_ref) or, where that was never built, the C port.
def cpu_render_rate(a, pp_sample, threads):
    """Renders one width x height frame at pp_sample on `threads` host threads; returns (samples/s, kind)."""
    from oracle import ref_harness
    samples = a.width * a.height * pp_sample * pp_sample
    if ref_harness.available():
        h = ref_harness.RefHarness()
        h.lib.ref_set_core_count(threads)
        h.load(a.world, a.width, a.height, pp_sample, 0, 7)
        h.seed(1)
        sec, _ = h.render_threads(threads)
        return samples / sec, "reference"
    from oracle import pt_oracle as O
    from pathtracer_b200 import host
    sc = host.HostWorld(a.world).scene(a.width, a.height)
    sec, _, _ = O.render_threads(sc, a.width, a.height, pp_sample, 1, threads)
    return samples / sec, "port"


def best_thread_count(a, pp_sample, cores):
    """The reference adds two busy-spinning service threads to its workers (master + viewer, src/win32_main.cpp:252-274,
    937-960), so `-t C` oversubscribes a C-core host by two: time -t C and -t C-2 and keep the better.
    Returns (threads, {threads: Msamples/s})."""
    tried = {}
    for t in sorted({cores, max(1, cores - 2)}, reverse=True):
        rate, _ = cpu_render_rate(a, pp_sample, t)
        tried[t] = rate / 1e6
    return max(tried, key=tried.get), tried


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    pp_s = 2          # -p1 under-reports the reference (per-tile hand-out overhead dominates): 0.91 vs 2.17 Msamples/s
    threads, tried = best_thread_count(a, pp_s, cores)          # also the first warm-up
    for _ in range(max(0, a.warmup - 1)):
        cpu_render_rate(a, pp_s, threads)
    t_tot, s_tot, kind = 0.0, 0.0, "port"
    for _ in range(a.steps):
        rate, kind = cpu_render_rate(a, pp_s, threads)
        n = a.width * a.height * pp_s * pp_s
        t_tot += n / rate
        s_tot += n
    v = s_tot / t_tot / 1e6
    sample = (f"each step = one {a.width}x{a.height} frame at -p{pp_s} ({pp_s * pp_s} spp) of the same world (the rate is spp-independent from -p2 up); "
              f"-t {threads} on {cores} host cores (tried: " + ", ".join(f"-t {t}: {r:.2f}" for t, r in tried.items()) + " Msamples/s)")
    print(json.dumps({
        "impl": "reference", "metric": "Msamples/s", "value": v, "unit": "Msamples/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": t_tot / a.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(a), "sample": sample},
        "cpu_baseline": {"value": v, "unit": "Msamples/s", "cores": cores, "threads": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


# ---------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop = index, [], threading.Event()
        self.proc = None
        self.t_begin = self.t_end = None        # host timestamps of the timed region; rows outside it are dropped

    def mark_begin(self):
        self.t_begin = time.monotonic()

    def mark_end(self):
        self.t_end = time.monotonic()

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")] + [time.monotonic()])
                if self._stop.is_set():
                    break
        except Exception:
            pass

    def stop(self):
        self._stop.set()
        if self.proc:
            self.proc.terminate()
        rows = [r for r in self.rows if len(r) >= 8 and r[0].isdigit()]
        if self.t_begin is not None and self.t_end is not None:
            # nvidia-smi is started before the warm-up (it needs ~0.5 s to come up); keep the samples of the timed region,
            # one sampling period of slack on the far side
            inside = [r for r in rows if self.t_begin <= r[-1] <= self.t_end + 0.1]
            rows = inside or [r for r in rows if r[-1] >= self.t_begin]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        sm = sorted(int(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": int(rows[0][1]), "reasons": reasons, "samples": len(rows)}


def ncu_record(world, width, height):
    """What the committed `ncu --set full` capture of this world's k_render variant measured (profiles/k_render_ncu.json,
    written by tools/ncu_summary.py, with the commit it profiled): DRAM bytes per launch (the accumulator is read and
    written once per launch, so the figure scales with the pixel count, not with spp) and EXECUTED FP32 flops per sample
    (thread-level FADD + FMUL + 2 x FFMA; a property of the code and the scene, not of the clock).  {} where none exists."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "k_render_ncu.json")))
        r = dict(t.get(f"w{world}", {}))
        if r and r.get("pixels"):
            r["dram_bytes_per_launch"] = r["dram_bytes_per_launch"] * (width * height) / r["pixels"]
        return r
    except Exception:
        return {}


def scene_bytes(sc, textures=True):
    n = sc.n_materials * 68 + sc.n_spheres * 20 + sc.n_planes * 20 + sc.n_quads * 40 + sc.n_aabbs * 52 + sc.n_triangles * 40
    for i in range(4):
        if textures and sc.textures[i].rgba8:
            n += sc.textures[i].width * sc.textures[i].height * 4
    return n + 116


def run_b200(a):
    import numpy as np
    import torch
    import torch.distributed as dist
    from pathtracer_b200 import _ctypes_api as T
    from pathtracer_b200 import dist as D
    from pathtracer_b200 import host, render

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # a non-default stream: torch events, NCCL and the library's kernels all order on it
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    kern = {"auto": T.KERNEL_AUTO, "mega": T.KERNEL_MEGA, "wavefront": T.KERNEL_WAVEFRONT}[a.kernel]
    core = render.RenderCore(local, kernel=kern, stream=stream.cuda_stream)
    hw = host.HostWorld(a.world)
    sc = hw.scene(a.width, a.height)
    core.upload_scene(sc)
    frame = D.ShardedFrame(core, a.width, a.height, a.shard)
    frame.phase_events = []
    samples_per_step = a.width * a.height * a.pp * a.pp

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident metric -------------------------------------------------------------
    sampler = ClockSampler(local)
    sampler.start()                                     # before the warm-up: it is up and sampling when the timed region starts
    for i in range(a.warmup):
        core.flush_l2()                                 # also allocates the scratch buffer outside the timed region
        frame.render(a.pp, seed=100 + i)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    launches = 0
    frame.phase_events = []
    barrier()
    sampler.mark_begin()
    e0.record()
    for i in range(a.steps):
        core.flush_l2()                                 # 256 MiB memset > 126 MB L2, between timed iterations (inside the region)
        k_ev[i][0].record()
        out = frame.render(a.pp, seed=1 + i)
        k_ev[i][1].record()
        launches += frame.launches
    e1.record()
    barrier()
    sampler.mark_end()
    clocks = sampler.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    step_ms = [round(k_ev[i][0].elapsed_time(k_ev[i][1]), 3) for i in range(a.steps)]
    value = samples_per_step * a.steps / (ms_total * 1e-3) / 1e6
    # per-phase device times of every rank: render launch (k_render of its share) and everything after it (resolve / pack /
    # collective / unpack), so that the max-over-ranks step can be attributed
    mid = frame.phase_events[:a.steps]
    frame.phase_events = None
    ph = torch.tensor([sum(k_ev[i][0].elapsed_time(mid[i]) for i in range(a.steps)) / a.steps,
                       sum(mid[i].elapsed_time(k_ev[i][1]) for i in range(a.steps)) / a.steps], device="cuda")
    if world > 1:
        allph = [torch.zeros_like(ph) for _ in range(world)]
        dist.all_gather(allph, ph)
    else:
        allph = [ph]
    phases = {"render_ms_per_rank": [round(float(x[0]), 3) for x in allph], "exchange_ms_per_rank": [round(float(x[1]), 3) for x in allph]}

    # the dominant kernel alone (k_render of this rank's share), events on its launch stream
    kms = []
    for i in range(a.steps):
        if a.shard in ("tiles", "tiles_p2p"):
            core.render(a.width, a.height, a.pp, seed=1 + i, tile_rank=rank, tile_count=world)
        else:
            lo, hi = D.sample_range(a.pp, rank, world)
            core.render(a.width, a.height, a.pp, seed=1 + i, sample_begin=lo, sample_end=hi)
        _, _, st = core.readback(False, False)
        kms.append(st.render_ms)
    kernel_ms = sum(kms) / len(kms)
    rays_per_sample = st.rays / max(1, st.samples)

    # ---- end to end through the C-ABI with host buffers ------------------------------------
    pinned = torch.empty((a.height, a.width), dtype=torch.int32).pin_memory()
    pinned_np = pinned.numpy().view(np.uint32)
    # Static texel assets are uploaded once (the reference decodes them once, at start-up: LoadTexture main.cpp:1711-1746);
    # every timed step re-sends everything else -- geometry, materials, camera -- and rebuilds the BVH.
    has_tex = any(sc.textures[i].rgba8 for i in range(4))
    core.upload_scene(sc)
    sc_step = T.Scene.from_buffer_copy(sc)
    sc_step.flags |= T.SCENE_KEEP_TEXTURES
    for i in range(2):
        core.upload_scene(sc_step)
        frame.render(a.pp, seed=50 + i)
        if rank == 0:
            core.readback(True, False, out_bgra=pinned_np)
    barrier()
    t0 = time.perf_counter()
    for i in range(a.steps):
        core.upload_scene(sc_step)                        # host scene -> device (+ BVH build)
        frame.render(a.pp, seed=1 + i)
        if rank == 0:
            core.readback(True, False, out_bgra=pinned_np)      # device frame -> pinned host memory
    barrier()
    t_e2e = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_value = samples_per_step * a.steps / float(t_e2e.item()) / 1e6
    # textured worlds: the same loop with NO texture-keep flag (every texel re-sent every step), reported beside e2e so that
    # what the flag hides is on the record
    e2e_full = None
    if has_tex:
        barrier()
        t0 = time.perf_counter()
        for i in range(a.steps):
            core.upload_scene(sc)
            frame.render(a.pp, seed=1 + i)
            if rank == 0:
                core.readback(True, False, out_bgra=pinned_np)
        barrier()
        t_full = torch.tensor([time.perf_counter() - t0], device="cuda")
        if world > 1:
            dist.all_reduce(t_full, op=dist.ReduceOp.MAX)
        e2e_full = samples_per_step * a.steps / float(t_full.item()) / 1e6

    if rank == 0:
        import json as _json
        peaks = {}
        p = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(p):
            peaks = _json.load(open(p))
        fp32_peak = core.measure_fp32_peak()
        my_share = 1.0 / world
        algo_flops = samples_per_step * my_share * F_PER_SAMPLE[a.world]
        achieved_tf = algo_flops / (kernel_ms * 1e-3) / 1e12
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        algo_bytes = a.width * a.height * my_share * 20.0
        rec = ncu_record(a.world, a.width, a.height)
        executed = None
        if rec.get("executed_flop_per_sample"):
            ex_tf = samples_per_step * my_share * rec["executed_flop_per_sample"] / (kernel_ms * 1e-3) / 1e12
            executed = {"achieved": ex_tf, "unit": "TFLOP/s", "frac": ex_tf / fp32_peak if fp32_peak else None,
                        "flops_per_sample": rec["executed_flop_per_sample"], "threads_per_inst": rec.get("threads_per_inst"),
                        "issue_active_pct": rec.get("issue_active_pct"), "warp_inst_per_sample": rec.get("warp_inst_per_sample"),
                        "source": f"ncu smsp__sass_thread_inst_executed_op_(fadd+fmul+2*ffma) per sample, profiles/k_render_ncu.json, commit {rec.get('commit')}; "
                                  "x this run's samples / this run's kernel time"}
        roofline = {
            "bound": "fp32", "kernel": "k_render", "achieved": achieved_tf, "peak": fp32_peak, "unit": "TFLOP/s",
            "frac": achieved_tf / fp32_peak if fp32_peak else None, "executed": executed,
            "peak_source": "measured live: FMA-chain probe pt_measure_fp32_peak (nominal 148 SM x 128 x 2 x 1.965 GHz = 74.4)",
            "flops_per_sample": F_PER_SAMPLE[a.world], "kernel_ms": kernel_ms, "rays_per_sample": rays_per_sample,
            "traffic": rec.get("dram_bytes_per_launch"), "traffic_source": (f"ncu --set full, commit {rec.get('commit')}" if rec else None),
            "hbm": {"bound": "hbm", "achieved": algo_bytes / (kernel_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "frac": algo_bytes / (kernel_ms * 1e-3) / 1e9 / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)",
                    "bytes_per_pixel": 20},
        }
        line = {
            "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_total / a.steps, "step_ms": step_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(a), "world": a.world, "width": a.width, "height": a.height, "pp": a.pp,
                       "shard": a.shard if world > 1 else "none", "rng": "philox4x32-10", "kernel": a.kernel,
                       "l2": "256 MiB memset between timed steps (inside the timed region)"},
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": scene_bytes(sc, textures=False),
                    "d2h_bytes_per_step": a.width * a.height * 4,
                    "textures": (f"{scene_bytes(sc) - scene_bytes(sc, textures=False)} texel bytes uploaded once before the timed region "
                                 "(PT_SCENE_KEEP_TEXTURES), not per step") if has_tex else "none in this world",
                    "value_with_texels_resent_every_step": e2e_full},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "phases": phases,
        }
        if world == 1 and not a.no_cpu_baseline:
            try:
                cores = host_cores()
                pp_s = 3
                threads, tried = best_thread_count(a, 2, cores)
                rate, kind = cpu_render_rate(a, pp_s, threads)
                line["cpu_baseline"] = {"value": rate / 1e6, "unit": "Msamples/s", "cores": cores, "threads": threads, "kind": kind,
                                        "sample": f"one {a.width}x{a.height} frame at -p{pp_s} ({pp_s * pp_s} spp) of the same world, -t {threads} on {cores} host cores "
                                                  "(-p2 probes: " + ", ".join(f"-t {t}: {r:.2f}" for t, r in tried.items()) + " Msamples/s)"}
            except Exception as e:      # the CPU leg must never take the GPU number down with it
                line["cpu_baseline"] = {"value": None, "unit": "Msamples/s", "cores": host_cores(), "kind": "port", "sample": f"failed: {e}"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
