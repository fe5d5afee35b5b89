// The reference's dataset loaders (cpp/include/img_loader/*.h) without OpenCV: same class names, same file
// conventions, same metric units, same timestamps, for the depth stream the Scan3D path consumes.
//   ImageLoader       base: unit, path, K, load_intrinsics, load_depth, load_pose        (ImageLoader.h:50-262)
//   TumrgbdLoader     associated.txt lines "ts_rgb rgb_file ts_depth depth_file", 1/5000 m (TumrgbdLoader.h:44-116)
//   SynthLoader       depth/%03d.png + rgb/%03d.png from 001, 1/1000 m                    (SynthLoader.h:43-113)
//   Printed3dLoader   depth_%06d.png + color_%06d.png from 000000, 1/1000 m               (Printed3dLoader.h:43-113)
//   RedwoodLoader     sorted depth/*.png + rgb/*.jpg, 1/1000 m                            (RedwoodLoader.h:45-142)
// Depth: 16-bit PNG -> float metres as `convertTo(depth, CV_32FC1, unit_)` does: (float)sample * unit_, unit_ a float.
// Colour: Map*PixelSdf::update ignores it, so the file is only required to exist and be non-empty (the reference stops
// at the first frame whose colour image cannot be read); no JPEG/PNG colour decoding here.
#ifndef GSDF_HOST_IMG_LOADER_H_
#define GSDF_HOST_IMG_LOADER_H_

#include <dirent.h>

#include <algorithm>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "mat.h"
#include "png16.h"

#ifndef CV_32FC1
#define CV_32FC1 5
#endif

// One frame of a dataset as the loaders enumerate it: file names relative to the dataset directory and the two
// timestamps.  Enumeration (next_entry) is sequential and cheap; decoding (ImageLoader::decode) touches no loader
// state, so a thread pool may decode entries in parallel (host/prefetch.h).
struct FrameEntry {
    std::string depth_file, color_file, stamp_depth, stamp_rgb;
};

class ImageLoader {
protected:
    Mat3f K_;
    const float unit_;        // metres per depth sample
    const std::string path_;  // directory containing the data, with trailing '/'
    std::string timestamp_rgb_, timestamp_depth_;
    bool consecutive_numbers_;
    std::vector<std::string> timestamps_depth_, timestamps_rgb_;
    std::vector<float> depth_buf_;  // storage behind the cv::Mat handed to the caller (valid until the next load)

    static bool readable(const std::string& file) {
        std::ifstream f(file, std::ios::binary);
        return f.is_open() && f.peek() != std::ifstream::traits_type::eof();
    }

public:
    ImageLoader(float unit, const std::string& path, bool consecutive_numbers)
        : unit_(unit), path_(path), consecutive_numbers_(consecutive_numbers) {}
    virtual ~ImageLoader() {}

    Mat3f K() const { return K_; }
    std::string rgb_timestamp() { return timestamp_rgb_; }
    std::string depth_timestamp() { return timestamp_depth_; }
    std::vector<std::string> getDepthTimestamps() { return timestamps_depth_; }
    std::vector<std::string> getRgbTimestamps() { return timestamps_rgb_; }
    bool consecutive_numbers() { return consecutive_numbers_; }

    bool load_intrinsics(const std::string& filename = "intrinsics.txt") {  // ImageLoader.h:138-157
        if (filename.empty()) return false;
        std::ifstream infile(path_ + filename);
        if (!infile.is_open()) return false;
        for (int i = 0; i < 9; ++i) {
            float tmp = 0;
            infile >> tmp;
            K_.m[i] = tmp;
        }
        return true;
    }

    // Decode one entry's depth image into `out` (float metres) and check its colour image: thread-safe, no loader state.
    // `out` must hold at least capacity floats if capacity > 0 (a fixed, e.g. pinned, buffer); else `vec` is resized.
    bool decode(const FrameEntry& e, float* out, size_t capacity, std::vector<float>* vec, int& w, int& h, std::string& err) const {
        if (e.depth_file.empty()) { err = "Error: missing filename"; return false; }
        static thread_local gsdf_host::GrayImage img;  // per decoding thread: its buffers are reused from frame to frame
        const std::string perr = gsdf_host::read_png_gray(path_ + e.depth_file, img);
        if (!perr.empty()) { err = "Error: empty depth image " + path_ + e.depth_file + " (" + perr + ")"; return false; }
        if (vec) { vec->resize(img.px.size()); out = vec->data(); }
        else if (img.px.size() > capacity) { err = "Error: depth image " + path_ + e.depth_file + " is larger than the frame buffer"; return false; }
        for (size_t i = 0; i < img.px.size(); ++i) out[i] = (float)img.px[i] * unit_;
        w = img.width; h = img.height;
        if (e.color_file.empty()) { err = "Error: missing filename"; return false; }
        if (!readable(path_ + e.color_file)) { err = "Error: empty color image " + path_ + e.color_file; return false; }
        return true;
    }

    bool load_depth(const std::string& filename, cv::Mat& depth) {  // ImageLoader.h:159-175
        FrameEntry e;
        e.depth_file = filename;
        e.color_file = filename;  // existence of the depth file itself: always true once it decoded
        int w = 0, h = 0;
        std::string err;
        if (!decode(e, nullptr, 0, &depth_buf_, w, h, err)) { std::cerr << err << std::endl; return false; }
        depth = cv::Mat(h, w, CV_32FC1, depth_buf_.data());
        return true;
    }

    // The next frame of the sequence without decoding anything; false at the end.  load_next = next_entry + decode.
    virtual bool next_entry(FrameEntry& e) = 0;

    bool load_next(cv::Mat& color, cv::Mat& depth) {
        FrameEntry e;
        if (!next_entry(e)) return false;
        return load_depth(e.depth_file, depth) && load_color(e.color_file, color);
    }

    bool load_color(const std::string& filename, cv::Mat& /*color*/) {  // ImageLoader.h:194-216 (existence only, see above)
        if (filename.empty()) { std::cerr << "Error: missing filename" << std::endl; return false; }
        if (!readable(path_ + filename)) { std::cerr << "Error: empty color image " << path_ + filename << std::endl; return false; }
        return true;
    }

    // "timestamp tx ty tz qx qy qz qw" per line (ImageLoader.h:231-259)
    bool load_pose(const std::string& filename, std::vector<SE3>& poses, std::vector<std::string>* stamps = nullptr) {
        std::ifstream file(filename.c_str());
        if (!file.is_open()) { std::cout << "can't load poses!" << std::endl; return false; }
        std::string line;
        while (std::getline(file, line)) {
            std::istringstream s(line);
            std::string ts;
            float t[3], q[4];
            if (!(s >> ts >> t[0] >> t[1] >> t[2] >> q[0] >> q[1] >> q[2] >> q[3])) continue;
            if (q[3] * q[3] + q[0] * q[0] + q[1] * q[1] + q[2] * q[2] < 0.99f)
                std::cerr << "pose " << ts << " has invalid rotation" << std::endl;
            poses.push_back(SE3(q[3], q[0], q[1], q[2], Vec3f(t[0], t[1], t[2])));
            if (stamps) stamps->push_back(ts);
        }
        return true;
    }

    virtual void reset() = 0;
};

class TumrgbdLoader : public ImageLoader {
    std::ifstream assoc_file_;

public:
    explicit TumrgbdLoader(const std::string& path) : ImageLoader(1.f / 5000, path, false) { assoc_file_.open(path_ + "associated.txt"); }

    bool next_entry(FrameEntry& e) override {  // TumrgbdLoader.h:78-101
        std::string line = "#";
        while (line.empty() || line.at(0) == '#')
            if (!std::getline(assoc_file_, line)) return false;
        std::istringstream dss(line);
        dss >> timestamp_rgb_ >> e.color_file >> timestamp_depth_ >> e.depth_file;
        std::cout << "load image " << timestamp_rgb_ << std::endl;
        timestamps_depth_.push_back(timestamp_depth_);
        timestamps_rgb_.push_back(timestamp_rgb_);
        e.stamp_depth = timestamp_depth_; e.stamp_rgb = timestamp_rgb_;
        return true;
    }

    void reset() override {
        assoc_file_.close();
        assoc_file_.clear();
        assoc_file_.open(path_ + "associated.txt");
        timestamps_depth_.clear();
        timestamps_rgb_.clear();
    }
};

class SynthLoader : public ImageLoader {
    size_t counter = 1;

public:
    explicit SynthLoader(const std::string& path) : ImageLoader(1.f / 1000, path, true) {}

    bool next_entry(FrameEntry& e) override {  // SynthLoader.h:61-82 (the counter advances per enumerated frame)
        std::stringstream ss;
        ss << std::setfill('0') << std::setw(3) << counter;
        timestamp_rgb_ = timestamp_depth_ = ss.str();
        e.depth_file = "depth/" + timestamp_rgb_ + ".png";
        e.color_file = "rgb/" + timestamp_rgb_ + ".png";
        e.stamp_depth = e.stamp_rgb = timestamp_rgb_;
        ++counter;
        return true;
    }

    void reset() override { counter = 1; timestamps_depth_.clear(); timestamps_rgb_.clear(); }
};

class Printed3dLoader : public ImageLoader {
    size_t counter = 0;

public:
    explicit Printed3dLoader(const std::string& path) : ImageLoader(1.f / 1000, path, true) {}

    bool next_entry(FrameEntry& e) override {  // Printed3dLoader.h:61-82
        std::stringstream ss;
        ss << std::setfill('0') << std::setw(6) << counter;
        timestamp_rgb_ = timestamp_depth_ = ss.str();
        e.depth_file = "depth_" + timestamp_rgb_ + ".png";
        e.color_file = "color_" + timestamp_rgb_ + ".png";
        e.stamp_depth = e.stamp_rgb = timestamp_rgb_;
        ++counter;
        return true;
    }

    void reset() override { counter = 0; timestamps_depth_.clear(); timestamps_rgb_.clear(); }
};

class RedwoodLoader : public ImageLoader {
    std::vector<std::string> depth_filenames_, rgb_filenames_;
    size_t counter = 0;

    static std::vector<std::string> list(const std::string& dir, const std::string& ext) {
        std::vector<std::string> out;
        if (DIR* d = opendir(dir.c_str())) {
            while (dirent* e = readdir(d)) {
                const std::string n = e->d_name;
                if (n.length() > 3 && n.substr(n.length() - 3) == ext) out.push_back(n);
            }
            closedir(d);
        }
        std::sort(out.begin(), out.end());
        return out;
    }
    void init() {  // RedwoodLoader.h:57-96: the *.png of depth/ and the *.jpg of rgb/, each sorted by name
        depth_filenames_ = list(path_ + "depth/", "png");
        rgb_filenames_ = list(path_ + "rgb/", "jpg");
    }

public:
    explicit RedwoodLoader(const std::string& path) : ImageLoader(1.f / 1000, path, true) { init(); }

    bool next_entry(FrameEntry& e) override {  // RedwoodLoader.h:98-122
        if (counter >= depth_filenames_.size() || counter >= rgb_filenames_.size()) return false;
        timestamp_depth_ = depth_filenames_[counter];
        timestamp_rgb_ = rgb_filenames_[counter];
        timestamps_depth_.push_back(timestamp_depth_);
        timestamps_rgb_.push_back(timestamp_rgb_);
        std::cout << "------load depth " << timestamp_depth_ << ". " << std::endl;
        std::cout << "------load rgb " << timestamp_rgb_ << ". " << std::endl;
        e.depth_file = "depth/" + timestamp_depth_;
        e.color_file = "rgb/" + timestamp_rgb_;
        e.stamp_depth = timestamp_depth_; e.stamp_rgb = timestamp_rgb_;
        ++counter;
        return true;
    }

    void reset() override { init(); timestamps_depth_.clear(); timestamps_rgb_.clear(); counter = 0; }
};

#endif  // GSDF_HOST_IMG_LOADER_H_
