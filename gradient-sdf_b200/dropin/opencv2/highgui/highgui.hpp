// cv::imread for the reference's loaders (img_loader/ImageLoader.h:167,184,204): 8/16-bit greyscale PNG through the library's
// own zlib reader (host/png16.h).  A file that exists but is not a greyscale PNG (the datasets' colour images) comes back as a
// non-empty CV_8UC3 image of unspecified content — the Scan3D path never reads colour.
#ifndef GSDF_DROPIN_OPENCV_HIGHGUI_HPP_
#define GSDF_DROPIN_OPENCV_HIGHGUI_HPP_

#include <fstream>
#include <string>

#include "../core/core.hpp"
#include "../../../host/png16.h"

namespace cv {

enum { IMREAD_UNCHANGED = -1, IMREAD_GRAYSCALE = 0, IMREAD_COLOR = 1, IMREAD_ANYDEPTH = 2, IMREAD_ANYCOLOR = 4 };

inline Mat imread(const std::string& filename, int flags = IMREAD_COLOR) {
    (void)flags;
    gsdf_host::GrayImage img;
    if (gsdf_host::read_png_gray(filename, img).empty()) {
        const size_t n = (size_t)img.width * img.height;
        Mat m(img.height, img.width, img.bit_depth == 16 ? CV_16U : CV_8U);
        if (img.bit_depth == 16) std::memcpy(m.data, img.px.data(), n * 2);
        else for (size_t i = 0; i < n; ++i) m.data[i] = (unsigned char)img.px[i];
        return m;
    }
    std::ifstream f(filename.c_str(), std::ios::binary);
    if (!f.good()) return Mat();      // missing file: empty, as cv::imread
    return Mat(1, 1, CV_8UC3);        // present but not decodable here (colour): placeholder
}

}  // namespace cv

#endif
