// Minimal PNG reader for depth images: 8- or 16-bit greyscale, non-interlaced, all five scanline filters, zlib for
// the inflate.  It stands in for what the reference's loaders get from
//     cv::imread(path, cv::IMREAD_ANYDEPTH | cv::IMREAD_ANYCOLOR)            (img_loader/ImageLoader.h:167)
// on the depth files of its four dataset conventions (16-bit single-channel PNGs).  Anything else (colour types,
// palettes, interlacing, bit depths below 8) is rejected with a message: the reference's `convertTo(depth, CV_32FC1)`
// would not yield a depth map from such a file either.  CRCs of the chunks are checked.
#ifndef GSDF_HOST_PNG16_H_
#define GSDF_HOST_PNG16_H_

#include <zlib.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace gsdf_host {

struct GrayImage {
    int width = 0, height = 0, bit_depth = 0;  // bit_depth 8 or 16
    std::vector<uint16_t> px;                  // row-major sample values (an 8-bit file is widened, not rescaled)
    // scratch kept between calls (file bytes, IDAT stream, inflated scanlines): a decoder thread that reuses one
    // GrayImage allocates nothing per frame — fresh 600 kB vectors per frame are mmap/munmap pairs, and a pool of
    // decoding threads then serialises on the process's address-space lock instead of inflating
    std::vector<unsigned char> file, idat, raw, zero;
};

inline uint32_t be32(const unsigned char* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

// Returns an empty string on success, else what is wrong with the file.
inline std::string read_png_gray(const std::string& path, GrayImage& out) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return "cannot open " + path;
    std::vector<unsigned char>& file = out.file;
    file.clear();
    unsigned char buf[1 << 16];
    size_t got;
    while ((got = std::fread(buf, 1, sizeof(buf), f)) > 0) file.insert(file.end(), buf, buf + got);
    std::fclose(f);
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (file.size() < 8 + 25 || std::memcmp(file.data(), sig, 8) != 0) return path + " is not a PNG file";
    size_t pos = 8;
    bool have_ihdr = false, done = false;
    int colour_type = -1, interlace = -1;
    std::vector<unsigned char>& idat = out.idat;
    idat.clear();
    while (!done) {
        if (pos + 12 > file.size()) return path + ": truncated chunk";
        const uint32_t len = be32(&file[pos]);
        const unsigned char* type = &file[pos + 4];
        if (pos + 12 + (size_t)len > file.size()) return path + ": truncated chunk data";
        const unsigned char* data = &file[pos + 8];
        const uint32_t crc = be32(&file[pos + 8 + len]);
        if ((uint32_t)crc32(crc32(0L, Z_NULL, 0), type, 4 + len) != crc) return path + ": chunk CRC mismatch";
        if (std::memcmp(type, "IHDR", 4) == 0) {
            if (len != 13) return path + ": bad IHDR";
            out.width = (int)be32(data);
            out.height = (int)be32(data + 4);
            out.bit_depth = data[8];
            colour_type = data[9];
            interlace = data[12];
            if (data[10] != 0 || data[11] != 0) return path + ": unknown compression / filter method";
            have_ihdr = true;
        } else if (std::memcmp(type, "IDAT", 4) == 0) {
            idat.insert(idat.end(), data, data + len);
        } else if (std::memcmp(type, "IEND", 4) == 0) {
            done = true;
        }
        pos += 12 + (size_t)len;
    }
    if (!have_ihdr || out.width <= 0 || out.height <= 0) return path + ": no image header";
    if (colour_type != 0) return path + ": not a single-channel greyscale PNG (colour type " + std::to_string(colour_type) + ")";
    if (out.bit_depth != 8 && out.bit_depth != 16) return path + ": bit depth " + std::to_string(out.bit_depth) + " not supported";
    if (interlace != 0) return path + ": interlaced PNGs are not supported";

    const size_t bpp = (size_t)out.bit_depth / 8;       // bytes per pixel = filter distance
    const size_t stride = (size_t)out.width * bpp;
    std::vector<unsigned char>& raw = out.raw;
    raw.resize((stride + 1) * (size_t)out.height);
    uLongf raw_len = (uLongf)raw.size();
    const int zr = uncompress(raw.data(), &raw_len, idat.data(), (uLong)idat.size());
    if (zr != Z_OK || raw_len != raw.size()) return path + ": inflate failed (" + std::to_string(zr) + ")";

    // undo the scanline filters in place (PNG specification, section 9: None, Sub, Up, Average, Paeth)
    std::vector<unsigned char>& zero = out.zero;
    zero.assign(stride, 0);
    for (int y = 0; y < out.height; ++y) {
        unsigned char* line = &raw[(stride + 1) * (size_t)y];
        const int ft = line[0];
        unsigned char* cur = line + 1;
        const unsigned char* up = y ? &raw[(stride + 1) * (size_t)(y - 1) + 1] : zero.data();
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = up[i], c = i >= bpp ? up[i - bpp] : 0;
            int pred;
            switch (ft) {
                case 0: pred = 0; break;
                case 1: pred = a; break;
                case 2: pred = b; break;
                case 3: pred = (a + b) >> 1; break;
                case 4: {
                    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
                    pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
                    break;
                }
                default: return path + ": unknown scanline filter " + std::to_string(ft);
            }
            cur[i] = (unsigned char)(cur[i] + pred);
        }
    }
    out.px.resize((size_t)out.width * out.height);
    for (int y = 0; y < out.height; ++y) {
        const unsigned char* cur = &raw[(stride + 1) * (size_t)y + 1];
        uint16_t* dst = &out.px[(size_t)y * out.width];
        if (bpp == 2)
            for (int x = 0; x < out.width; ++x) dst[x] = (uint16_t)((cur[2 * x] << 8) | cur[2 * x + 1]);  // big-endian samples
        else
            for (int x = 0; x < out.width; ++x) dst[x] = cur[x];
    }
    return "";
}

}  // namespace gsdf_host

#endif  // GSDF_HOST_PNG16_H_
