#!/usr/bin/env python
"""Config-size image parity: the reference's own render against this renderer, through the repo's image_compare.

For each BASELINE.json configuration (C1..C5 and the headline 1080p frame): the reference (oracle/_ref, every host core,
oracle/ref_parallel.py) renders the frame twice with different mt19937 seeds (A, B), the GPU renders it at the SAME
resolution and spp, and bin/image_compare reports
    * legacy "Percentage Similarity" and RMSE of the 8-bit images:  reference A vs GPU,  next to  A vs B (the noise floor),
    * the per-pixel z-test of the HDR means against the sample standard error:  fraction of |z| > 3, mean z, global z,
      reference A vs GPU  next to  A vs B.  The standard error of a pixel's mean at this spp is measured, not assumed: the
      sample variance of that pixel over K further GPU renders with other seeds (K = 32).  It is used for BOTH pairs, so the
      two rows are the same statistic on the same scale.  (A pixel's own sum of squares at 16..256 spp is a poor error bar
      here: the estimator is heavy-tailed -- a 15 W/sr sun times weights 2/pdf -- so the sample variance is correlated with
      the sample mean and z is skewed even for the reference against itself.  For the same reason the table quotes the
      fraction beyond 3 sigma and the MEDIAN z; mean and rms of z are dominated by a few firefly pixels and are in the raw
      output only.)
    * the global z of the whole-image mean, where the sum over two million pixels makes each side's OWN variance estimate
      sound: the reference's from its per-pixel sums of squares, the GPU's from the K renders;
    * a bias test with all the power the GPU can buy: reference A vs the MEAN of the K GPU renders (global z).
Where the reference at the configuration's spp would take longer than --budget seconds of CPU per render, the largest
-p that fits is used FOR BOTH SIDES and stated.  Test/measurement infrastructure: runs on a GPU box.

    python tools/parity_report.py [--configs C1,C2,...] [--budget 45] [--k 16] [--out gpurun_out/parity]
"""
import argparse, json, os, re, subprocess, sys, time
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)

CONFIGS = {   # name: (world, W, H, pp of the configuration, thin lens)
    "C1": (1, 1280, 720, 16, False), "C2": (2, 1280, 720, 16, False), "C3": (3, 1280, 720, 32, False),
    "C4": (4, 1920, 1080, 23, True), "C5": (5, 3840, 2160, 32, False), "HEAD": (5, 1920, 1080, 32, False)}


def write_hdr(path, mean_sum, count, var_of_mean):
    H, W, _ = mean_sum.shape
    a = np.zeros((H, W, 4), np.float32)
    a[..., :3] = mean_sum
    a[..., 3] = count
    v = np.zeros((H, W, 4), np.float32)
    v[..., :3] = var_of_mean
    with open(path, "wb") as f:
        f.write(np.array([W, H], np.int32).tobytes()); f.write(a.tobytes()); f.write(v.tobytes())


def image_compare(*args):
    exe = os.path.join(ROOT, "pathtracer_b200", "bin", "image_compare")
    return subprocess.run([exe, *args], capture_output=True, text=True, check=True).stdout


def parse_ldr(txt):
    return dict(similarity=float(re.search(r"Percentage Similarity: ([\d.]+)", txt).group(1)),
                rmse=float(re.search(r"RMSE \(0-255, RGB\): ([\d.]+)", txt).group(1)))


def parse_hdr(txt):
    m = re.search(r"\|z\| > [\d.]+ on ([\d.]+) % .*median z: (-?[\d.]+)\s+mean z: (-?[\d.]+)\s+rms z: ([\d.]+)\s+global z of the image mean: (-?[\d.]+)", txt)
    return dict(frac_z3=float(m.group(1)), median_z=float(m.group(2)), mean_z=float(m.group(3)), rms_z=float(m.group(4)), global_z=float(m.group(5)),
                hdr_rmse=float(re.search(r"HDR RMSE: ([\d.]+)", txt).group(1)))


def run_config(name, budget, K, out_dir, rc_cache):
    from oracle import ref_parallel as RP
    from pathtracer_b200 import host, render
    world, W, H, pp_cfg, lens = CONFIGS[name]
    flags = 1 if lens else 0
    # probe the reference's rate (a -p2 frame, minus the workers' start-up measured on an 8x8 frame), then take the largest
    # -p <= the configuration's that fits the budget
    t0 = RP.render(world, 8, 8, 1, flags, 7, 98)["seconds"]
    t = RP.render(world, W, H, 2, flags, 7, 99)["seconds"]
    rate = W * H * 4 / max(0.05, t - t0)
    pp = max(2, min(pp_cfg, int((budget * rate / (W * H)) ** 0.5)))
    A = RP.render(world, W, H, pp, flags, 7, 1)
    B = RP.render(world, W, H, pp, flags, 7, 2)
    spp = pp * pp
    rc = rc_cache.setdefault("rc", render.RenderCore(0))
    hw = host.HostWorld(world, scene_seed=7)
    rc.upload_scene(hw.scene(W, H, lens=lens))
    acc = np.zeros((H, W, 3), np.float64)
    acc2 = np.zeros((H, W, 3), np.float64)
    ms = []
    rc.render(W, H, pp, seed=1)                     # THE GPU image of this configuration
    bgra1, a, _ = rc.readback(True, True)
    bgra1, g1 = bgra1.copy(), a[..., :3].astype(np.float64) / spp
    for k in range(K):                              # K more renders: the per-pixel variance of such a mean, and a precise mean
        rc.render(W, H, pp, seed=1001 + k)
        _, a, st = rc.readback(False, True)
        assert (a[..., 3] == spp).all()
        m = a[..., :3].astype(np.float64) / spp
        acc += m; acc2 += m * m
        ms.append(st.render_ms)
    gm = acc / K
    V = np.maximum(0.0, acc2 / K - gm * gm) * K / max(1, K - 1)          # variance of ONE render's pixel mean
    d = os.path.join(out_dir, name)
    os.makedirs(d, exist_ok=True)
    p = lambda f: os.path.join(d, f)
    host.write_bmp(p("ref_a.bmp"), A["ldr"]); host.write_bmp(p("ref_b.bmp"), B["ldr"]); host.write_bmp(p("gpu.bmp"), bgra1)
    write_hdr(p("ref_a.hdr.bin"), A["sum"], spp, V); write_hdr(p("ref_b.hdr.bin"), B["sum"], spp, V)
    write_hdr(p("gpu.hdr.bin"), g1 * spp, spp, V); write_hdr(p("gpu_k.hdr.bin"), gm * (K * spp), K * spp, V / K)
    write_hdr(p("ref_a_own.hdr.bin"), A["sum"], spp, RP.variance_of_mean(A)); write_hdr(p("ref_b_own.hdr.bin"), B["sum"], spp, RP.variance_of_mean(B))
    own = {"ref_gpu": parse_hdr(image_compare("--hdr", p("ref_a_own.hdr.bin"), p("gpu.hdr.bin")))["global_z"],
           "ref_ref": parse_hdr(image_compare("--hdr", p("ref_a_own.hdr.bin"), p("ref_b_own.hdr.bin")))["global_z"],
           "ref_gpumean": parse_hdr(image_compare("--hdr", p("ref_a_own.hdr.bin"), p("gpu_k.hdr.bin")))["global_z"]}
    raw = {"ldr_ref_gpu": image_compare(p("ref_a.bmp"), p("gpu.bmp")), "ldr_ref_ref": image_compare(p("ref_a.bmp"), p("ref_b.bmp")),
           "hdr_ref_gpu": image_compare("--hdr", p("ref_a.hdr.bin"), p("gpu.hdr.bin")),
           "hdr_ref_ref": image_compare("--hdr", p("ref_a.hdr.bin"), p("ref_b.hdr.bin")),
           "hdr_ref_gpumean (bias test)": image_compare("--hdr", p("ref_a.hdr.bin"), p("gpu_k.hdr.bin")),
           "hdr_ref_ref, each side's own sum of squares as the error bar (skewed, see the header of tools/parity_report.py)":
               image_compare("--hdr", p("ref_a_own.hdr.bin"), p("ref_b_own.hdr.bin"))}
    for f in ("ref_a.hdr.bin", "ref_b.hdr.bin", "gpu.hdr.bin", "gpu_k.hdr.bin", "ref_a_own.hdr.bin", "ref_b_own.hdr.bin", "ref_b.bmp"):
        os.remove(os.path.join(d, f))                # the dumps are large; the numbers and two BMPs stay
    if W * H > 1280 * 720:
        for f in ("ref_a.bmp", "gpu.bmp"):
            os.remove(os.path.join(d, f))
    open(os.path.join(d, "image_compare.txt"), "w").write("\n".join(f"== {k}\n{v}" for k, v in raw.items()))
    row = dict(config=name, world=world, width=W, height=H, pp_config=pp_cfg, pp_compared=pp, spp_compared=spp, lens=lens,
               reference_kind=A["kind"], reference_seconds=[A["seconds"], B["seconds"]], host_workers=A["workers"],
               gpu_renders=K, gpu_ms_per_render=float(np.median(ms)),
               ref_vs_gpu=dict(**parse_ldr(raw["ldr_ref_gpu"]), **parse_hdr(raw["hdr_ref_gpu"]), global_z_own=own["ref_gpu"]),
               ref_vs_gpu_mean_of_k=dict(median_z=parse_hdr(raw["hdr_ref_gpumean (bias test)"])["median_z"], global_z_own=own["ref_gpumean"]),
               ref_vs_ref=dict(**parse_ldr(raw["ldr_ref_ref"]), **parse_hdr(raw["hdr_ref_ref"]), global_z_own=own["ref_ref"]))
    return row


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="C1,C2,C3,C4,C5,HEAD")
    ap.add_argument("--budget", type=float, default=45.0, help="seconds of wall time per reference render")
    ap.add_argument("--k", type=int, default=32)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "parity"))
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    rows, cache = [], {}
    for name in a.configs.split(","):
        t0 = time.time()
        row = run_config(name, a.budget, a.k, a.out, cache)
        row["seconds_total"] = time.time() - t0
        rows.append(row)
        print(json.dumps(row), flush=True)
        json.dump(rows, open(os.path.join(a.out, "parity_report.json"), "w"), indent=1)
    md = ["| Config | compared at | pair | Percentage Similarity | RMSE (0-255) | HDR RMSE | \\|z\\| > 3 | median z | global z |", "|---|---|---|---|---|---|---|---|---|"]
    for r in rows:
        for pair, k in (("reference vs GPU", "ref_vs_gpu"), ("reference vs reference (noise floor)", "ref_vs_ref")):
            v = r[k]
            md.append(f"| {r['config']} w{r['world']} {r['width']}x{r['height']}{' -d' if r['lens'] else ''} | -p{r['pp_compared']} ({r['spp_compared']} spp; config -p{r['pp_config']}) | {pair} | "
                      f"{v['similarity']:.4f} % | {v['rmse']:.4f} | {v['hdr_rmse']:.5f} | {v['frac_z3']:.3f} % | {v['median_z']:+.4f} | {v['global_z_own']:+.2f} |")
        b = r["ref_vs_gpu_mean_of_k"]
        md.append(f"| | | reference vs mean of {r['gpu_renders']} GPU renders (bias) | | | | | {b['median_z']:+.4f} | {b['global_z_own']:+.2f} |")
    open(os.path.join(a.out, "parity_report.md"), "w").write("\n".join(md) + "\n")
    print("\n".join(md))


if __name__ == "__main__":
    main()
