// image_compare -- successor of the reference's ImageCompare.exe (src/image_compare.c).
//
//   image_compare a.bmp b.bmp            legacy "Percentage Similarity" + RMSE + PSNR + max abs difference
//   image_compare --hdr a.hdr.bin b.hdr.bin [--sigma 3]
//                                        two float accumulator dumps (pathtracer --hdr): HDR RMSE, bias, outliers against the
//                                        pooled difference; and, when BOTH dumps carry the optional second plane (per pixel
//                                        and channel: the variance of the mean, i.e. the squared standard error), the
//                                        per-pixel z-test of BASELINE.json's north_star: z = (a - b) / sqrt(se_a^2 + se_b^2),
//                                        fraction of |z| > sigma, mean z, and the global z of the image mean
//                                        (tools/parity_report.py writes such dumps for the reference and for this renderer)
//
// The legacy number is kept BUG-COMPATIBLE and labelled so: the reference's Magnitude() (image_compare.c:121-124)
// is sqrt(a*a + b*b + g*g + a*a) -- red is never compared and alpha is counted twice -- on channels scaled to
// [0,1] (ExtractColorVector :126-134), and similarity = mean(1 - distance) * 100 (CompareImages :79-104).
// RMSE over the three colour channels is the number BASELINE.json's north_star asks for.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "pt_host.h"

static void usage() { std::printf("usage: image_compare image_file1 image_file2 [--hdr] [--sigma S]\n"); }   // image_compare.c:146-149

// dump layout: int32 width, height; width*height float4 (sum r, g, b, sample count); optionally a second plane of
// width*height float4 (variance of the mean of r, g, b; unused)
static bool read_hdr(const char *path, int &w, int &h, std::vector<float> &v, std::vector<float> &var) {
    FILE *f = std::fopen(path, "rb");
    if (!f) return false;
    int32_t hd[2];
    bool ok = std::fread(hd, sizeof(hd), 1, f) == 1 && hd[0] > 0 && hd[1] > 0;
    if (ok) {
        w = hd[0]; h = hd[1];
        v.resize((size_t)w * h * 4);
        ok = std::fread(v.data(), sizeof(float), v.size(), f) == v.size();
        if (ok) {
            var.resize(v.size());
            if (std::fread(var.data(), sizeof(float), var.size(), f) != var.size()) var.clear();
        }
    }
    std::fclose(f);
    return ok;
}

int main(int argc, char **argv) {
    const char *p1 = nullptr, *p2 = nullptr;
    bool hdr = false;
    double sigma = 3.0;
    for (int i = 1; i < argc; i++) {
        if (!std::strcmp(argv[i], "--hdr")) hdr = true;
        else if (!std::strcmp(argv[i], "--sigma") && i + 1 < argc) sigma = std::atof(argv[++i]);
        else if (!p1) p1 = argv[i];
        else if (!p2) p2 = argv[i];
    }
    if (!p1 || !p2) { usage(); return 1; }

    if (hdr) {
        // two independent renders of the same scene: every pixel-channel mean is an estimate with standard error
        // ~ |mean| / sqrt(n) at best; without per-sample variance in the dump the test uses the pooled difference
        // variance of the image as the noise scale and reports the fraction of outliers beyond `sigma`.
        int w1, h1, w2, h2;
        std::vector<float> a, b, va, vb;
        if (!read_hdr(p1, w1, h1, a, va)) { std::printf("Error: '%s' is an invalid file path\n", p1); return -1; }
        if (!read_hdr(p2, w2, h2, b, vb)) { std::printf("Error: '%s' is an invalid file path\n", p2); return -1; }
        if (w1 != w2) { std::printf("Error: Images must have same width but they are %d and %d\n", w1, w2); return -1; }
        if (h1 != h2) { std::printf("Error: Images must have same height but they are %d and %d\n", h1, h2); return -1; }
        const size_t n = (size_t)w1 * h1;
        double sum_d = 0, sum_d2 = 0, sum_a = 0, sum_b = 0;
        std::vector<double> diff(n * 3);
        for (size_t i = 0; i < n; i++)
            for (int c = 0; c < 3; c++) {
                double ma = a[4 * i + 3] > 0 ? a[4 * i + c] / a[4 * i + 3] : 0.0, mb = b[4 * i + 3] > 0 ? b[4 * i + c] / b[4 * i + 3] : 0.0;
                double d = ma - mb;
                diff[3 * i + c] = d;
                sum_d += d; sum_d2 += d * d; sum_a += ma; sum_b += mb;
            }
        const double N = (double)n * 3, mean_d = sum_d / N, sd = std::sqrt(std::fmax(0.0, sum_d2 / N - mean_d * mean_d));
        size_t out = 0;
        for (double d : diff) if (std::fabs(d - mean_d) > sigma * sd) out++;
        std::printf("HDR mean A: %.6f  mean B: %.6f  relative difference of means: %.4f %%\n", sum_a / N, sum_b / N,
                    sum_a != 0 ? 100.0 * (sum_a - sum_b) / sum_a : 0.0);
        std::printf("HDR RMSE: %.6f  bias (mean difference): %.3e = %.2f standard errors\n", std::sqrt(sum_d2 / N), mean_d,
                    sd > 0 ? mean_d / (sd / std::sqrt(N)) : 0.0);
        std::printf("Outliers beyond %.1f sigma: %.4f %%\n", sigma, 100.0 * (double)out / N);
        if (!va.empty() && !vb.empty()) {
            // per-pixel, per-channel z-test against the sample standard errors of the two estimates; a relative 1e-5 floor
            // keeps pixels that are constant on both sides (sky) from dividing a float rounding difference by zero
            size_t beyond = 0;
            double sz = 0, sz2 = 0, sd_all = 0, sv_all = 0;
            std::vector<float> zs;
            zs.reserve(n * 3);
            for (size_t i = 0; i < n; i++)
                for (int c = 0; c < 3; c++) {
                    double ma = a[4 * i + 3] > 0 ? a[4 * i + c] / a[4 * i + 3] : 0.0, mb = b[4 * i + 3] > 0 ? b[4 * i + c] / b[4 * i + 3] : 0.0;
                    double fl = 1e-5 * std::fmax(std::fabs(ma), std::fabs(mb));
                    double v = (double)va[4 * i + c] + (double)vb[4 * i + c] + fl * fl + 1e-14;
                    double z = (ma - mb) / std::sqrt(v);
                    if (std::fabs(z) > sigma) beyond++;
                    zs.push_back((float)z);
                    sz += z; sz2 += z * z; sd_all += ma - mb; sv_all += v;
                }
            // the estimator is heavy-tailed (fireflies): mean and rms of z are dominated by a few pixels, the median is not
            std::nth_element(zs.begin(), zs.begin() + zs.size() / 2, zs.end());
            std::printf("Per-pixel z-test (sample standard errors): |z| > %.1f on %.4f %% of pixel-channels   median z: %.4f   mean z: %.4f   rms z: %.4f   "
                        "global z of the image mean: %.3f\n", sigma, 100.0 * (double)beyond / N, (double)zs[zs.size() / 2], sz / N, std::sqrt(sz2 / N),
                        sd_all / std::sqrt(sv_all));
        }
        return 0;
    }

    uint32_t *d1 = nullptr, *d2 = nullptr;
    int x1, y1, x2, y2;
    if (pt_host_read_bmp(p1, &d1, &x1, &y1)) { std::printf("Error: '%s' is an invalid file path\n", p1); return -1; }
    if (pt_host_read_bmp(p2, &d2, &x2, &y2)) { std::printf("Error: '%s' is an invalid file path\n", p2); return -1; }
    if (x1 != x2) { std::printf("Error: Images must have same width but they are %d and %d\n", x1, x2); return -1; }
    if (y1 != y2) { std::printf("Error: Images must have same height but they are %d and %d\n", y1, y2); return -1; }
    const size_t n = (size_t)x1 * y1;
    double legacy = 0, se = 0;
    int max_abs = 0;
    for (size_t i = 0; i < n; i++) {
        // BGRA words: stb_image would hand the reference R,G,B,A bytes = (w>>16, w>>8, w, w>>24)
        int dr = (int)((d1[i] >> 16) & 255) - (int)((d2[i] >> 16) & 255), dg = (int)((d1[i] >> 8) & 255) - (int)((d2[i] >> 8) & 255);
        int db = (int)(d1[i] & 255) - (int)(d2[i] & 255), da = (int)(d1[i] >> 24) - (int)(d2[i] >> 24);
        double fa = da / 255.0, fb = db / 255.0, fg = dg / 255.0;
        legacy += 1.0 - std::sqrt(fa * fa + fb * fb + fg * fg + fa * fa);          // sic: no red, alpha twice
        se += (double)dr * dr + (double)dg * dg + (double)db * db;
        max_abs = std::max(max_abs, std::max(std::abs(dr), std::max(std::abs(dg), std::abs(db))));
    }
    const double rmse = std::sqrt(se / ((double)n * 3));
    std::printf("Percentage Similarity: %f %%\n", legacy / (double)n * 100.0);     // image_compare.c:71 (legacy, bug-compatible)
    std::printf("RMSE (0-255, RGB): %f   normalised: %f   PSNR: %.2f dB   max abs diff: %d\n", rmse, rmse / 255.0,
                rmse > 0 ? 20.0 * std::log10(255.0 / rmse) : INFINITY, max_abs);
    std::free(d1);
    std::free(d2);
    return 0;
}
