// Test helper for the dataset loaders (no GPU, no library): loads every frame of a dataset directory through
// img_loader.h and writes
//     <out>.f32   the depth frames, float32 metres, concatenated
//     stdout      one line per frame: "<depth timestamp> <rgb timestamp> <width> <height>", then "K k00 ... k22"
//   loader_probe <tum|synth|printed|rw> <dataset dir with trailing /> <out prefix> [threads [slots]]
// With threads > 0 the frames go through FramePrefetcher (parallel decode, in-order delivery) instead of load_next.
#include <cstdio>

#include "prefetch.h"

int main(int argc, char** argv) {
    if (argc < 4 || argc > 6) { std::fprintf(stderr, "usage: loader_probe <type> <dir/> <out prefix> [threads [slots]]\n"); return 2; }
    const std::string type = argv[1], dir = argv[2], out = argv[3];
    const int threads = argc > 4 ? std::atoi(argv[4]) : 0, slots = argc > 5 ? std::atoi(argv[5]) : 6;
    ImageLoader* loader = nullptr;
    if (type == "tum") loader = new TumrgbdLoader(dir);
    else if (type == "synth") loader = new SynthLoader(dir);
    else if (type == "printed") loader = new Printed3dLoader(dir);
    else if (type == "rw" || type == "redwood") loader = new RedwoodLoader(dir);
    else { std::fprintf(stderr, "unknown dataset type\n"); return 2; }
    FILE* f = std::fopen(out == "-" ? "/dev/null" : (out + ".f32").c_str(), "wb");  // "-": timing runs, nothing kept
    if (!f) return 2;
    std::streambuf* keep = std::cout.rdbuf(std::cerr.rdbuf());  // the loaders chat on cout like the reference's: keep stdout clean
    cv::Mat color, depth;
    std::vector<std::string> lines;
    if (threads > 0) {
        FramePrefetcher pf(loader, threads, slots, (size_t)1 << 22);
        std::vector<const FramePrefetcher::Frame*> held;  // four frames are held back, as the GPU pipeline holds its frames in flight
        while (const FramePrefetcher::Frame* fr = pf.next()) {
            std::fwrite(fr->data, sizeof(float), (size_t)fr->width * fr->height, f);
            lines.push_back(fr->stamp_depth + " " + fr->stamp_rgb + " " + std::to_string(fr->width) + " " + std::to_string(fr->height));
            held.push_back(fr);
            if ((int)held.size() >= std::min(4, slots - 1)) { pf.release(held.front()); held.erase(held.begin()); }
        }
        if (!pf.error().empty()) std::cerr << pf.error() << std::endl;
    } else
    while (loader->load_next(color, depth)) {
        std::fwrite(depth.data, sizeof(float), (size_t)depth.rows * depth.cols, f);
        lines.push_back(loader->depth_timestamp() + " " + loader->rgb_timestamp() + " " + std::to_string(depth.cols) + " " +
                        std::to_string(depth.rows));
    }
    std::fclose(f);
    std::cout.rdbuf(keep);
    for (const auto& l : lines) std::printf("%s\n", l.c_str());
    if (loader->load_intrinsics()) {
        const Mat3f K = loader->K();
        std::printf("K");
        for (int i = 0; i < 9; ++i) std::printf(" %.9g", K.m[i]);
        std::printf("\n");
    }
    delete loader;
    return 0;
}
