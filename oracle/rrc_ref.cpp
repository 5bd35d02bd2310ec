// ORACLE (test infrastructure only): line-by-line restatement of the use_rrc branch of
// FFMPEGThreadedDecoder::SetCodecContext (third_party/decord/src/video/ffmpeg/threaded_decoder.cc:33-93), with the
// generator seed passed in (the reference: std::mt19937_64 generator(rd())).  Output: x, y, out_w, out_h, hflip.
// ffmpeg itself is absent, so the golden vectors for this function are produced by this restatement compiled with
// the same libstdc++ <random> the reference is built against; parity of the draw sequence is by construction.
#include <cmath>
#include <cstdint>
#include <random>
using namespace std;

extern "C" void rrc_ref(uint64_t seed, int orig_width, int orig_height, double scale_min, double scale_max,
                        double ratio_min, double ratio_max, double hflip_prob, int32_t* out5) {
    std::mt19937_64 generator(seed);                                          // :34-35
    std::uniform_real_distribution<double> hflip_sampler(0, 1);               // :36
    bool hflip = false;
    if ((hflip_prob > 0.) && (hflip_sampler(generator) < hflip_prob)) {       // :43
        hflip = true;
    }
    int area = orig_width * orig_height;                                      // :50
    double log_ratio_min = log(ratio_min);
    double log_ratio_max = log(ratio_max);
    int out_w = -1;
    int out_h = -1;
    int x = -1;
    int y = -1;
    std::uniform_real_distribution<double> scale_sampler(scale_min, scale_max);
    std::uniform_real_distribution<double> ratio_sampler(log_ratio_min, log_ratio_max);
    bool success = false;
    for (int i = 0; (i < 10) && (!success); ++i) {                            // :60
        double target_area = area * scale_sampler(generator);
        double aspect_ratio = exp(ratio_sampler(generator));
        out_w = int(round(sqrt(target_area * aspect_ratio)));
        out_h = int(round(sqrt(target_area / aspect_ratio)));
        if ((out_w > 0) && (out_w <= orig_width) && (out_h > 0) && (out_h <= orig_height)) {
            std::uniform_int_distribution<> y_sampler(0, orig_height - out_h + 1);
            std::uniform_int_distribution<> x_sampler(0, orig_width - out_w + 1);
            y = y_sampler(generator);
            x = x_sampler(generator);
            success = true;
        }
    }
    if (!success) {                                                           // :73
        double in_ratio = double(orig_width) / double(orig_height);
        if (in_ratio < ratio_min) {
            out_w = orig_width;
            out_h = int(round(out_w / ratio_min));
        } else if (in_ratio > ratio_max) {
            out_h = orig_height;
            out_w = int(round(out_h * ratio_max));
        } else {
            out_h = orig_height;
            out_w = orig_width;
        }
        y = (orig_height - out_h) / 2;
        x = (orig_width - out_w) / 2;
    }
    out5[0] = x; out5[1] = y; out5[2] = out_w; out5[3] = out_h; out5[4] = hflip ? 1 : 0;
}
