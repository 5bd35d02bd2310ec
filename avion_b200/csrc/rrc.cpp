// T0: random-resized-crop box sampler (host).  Same arithmetic and the same draw order as decord's
// FFMPEGThreadedDecoder::SetCodecContext, use_rrc branch (third_party/decord/src/video/ffmpeg/threaded_decoder.cc:33-93),
// but with an explicit seed: clip i uses std::mt19937_64(seed + i), where the reference seeds each reader from
// std::random_device.  One box per clip, shared by all of its frames.
#include <cmath>
#include <cstdint>
#include <random>

#include "../../include/avion_b200.h"

namespace {
struct Box { int x, y, w, h, hflip; };

Box center_fallback(int W, int H, double rmin, double rmax) {
    Box b{};
    const double in_ratio = double(W) / double(H);
    if (in_ratio < rmin) { b.w = W; b.h = int(std::round(b.w / rmin)); }
    else if (in_ratio > rmax) { b.h = H; b.w = int(std::round(b.h * rmax)); }
    else { b.w = W; b.h = H; }
    b.y = (H - b.h) / 2;
    b.x = (W - b.w) / 2;
    return b;
}

Box sample_one(std::mt19937_64& gen, int W, int H, double smin, double smax, double rmin, double rmax, double hflip_prob) {
    int hflip = 0;
    if (hflip_prob > 0.) {
        std::uniform_real_distribution<double> coin(0, 1);
        if (coin(gen) < hflip_prob) hflip = 1;
    }
    const int area = W * H;
    std::uniform_real_distribution<double> scale(smin, smax);
    std::uniform_real_distribution<double> logr(std::log(rmin), std::log(rmax));
    for (int attempt = 0; attempt < 10; ++attempt) {
        const double target = area * scale(gen);
        const double ar = std::exp(logr(gen));
        const int w = int(std::round(std::sqrt(target * ar)));
        const int h = int(std::round(std::sqrt(target / ar)));
        if (w > 0 && w <= W && h > 0 && h <= H) {
            // inclusive upper bounds H-h+1 / W-w+1 are the reference's (one too large; ffmpeg's crop filter clamps)
            std::uniform_int_distribution<> ys(0, H - h + 1);
            std::uniform_int_distribution<> xs(0, W - w + 1);
            Box b{};
            b.y = ys(gen);
            b.x = xs(gen);
            b.w = w; b.h = h; b.hflip = hflip;
            return b;
        }
    }
    Box b = center_fallback(W, H, rmin, rmax);
    b.hflip = hflip;
    return b;
}
}  // namespace

extern "C" int avn_sample_rrc_boxes(uint64_t seed, int n, int src_w, int src_h, double scale_min, double scale_max,
                                    double ratio_min, double ratio_max, double hflip_prob, int32_t* out) {
    if (n <= 0 || src_w <= 0 || src_h <= 0 || out == nullptr || !(scale_min > 0) || scale_max < scale_min ||
        !(ratio_min > 0) || ratio_max < ratio_min)
        return 1;
    for (int i = 0; i < n; ++i) {
        std::mt19937_64 gen(seed + (uint64_t)i);
        const Box b = sample_one(gen, src_w, src_h, scale_min, scale_max, ratio_min, ratio_max, hflip_prob);
        out[5 * i + 0] = b.x; out[5 * i + 1] = b.y; out[5 * i + 2] = b.w; out[5 * i + 3] = b.h; out[5 * i + 4] = b.hflip;
    }
    return 0;
}
