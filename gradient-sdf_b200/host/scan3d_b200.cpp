// Scan3D on the B200 path: the loop of cpp/depth_scanning/src/main_scan_3d.cpp:62-319 over the forwarding shells of
// sdf_tracker.h.  Same options, same ordering contract (track against the map, fuse only if converged; first frame
// fused at identity or at the first ground-truth pose; with a pose file every frame is fused without tracking),
// same output files (<results>_poses.txt, <results>gradient_sdf_cloud_final.ply, <results>gradient_sdf_grid_info.txt,
// ..._sdf_d.txt, ..._sdf_weight.txt, ..._sdf_n{0,1,2}.txt).  --data-type tum | synth | printed | rw | redwood reads the
// reference's dataset conventions through img_loader.h (16-bit depth PNGs decoded with zlib, no OpenCV); in addition
// --data-type raw reads
//     <input>/intrinsics.txt   3x3 row-major, as ImageLoader::load_intrinsics (ImageLoader.h:138-157)
//     <input>/frames.f32       n frames of height*width float32 depths in metres (what convertTo yields, :172)
//     <input>/meta.txt         "width height n_frames"
//     <input>/<pose-file>      optional, "timestamp tx ty tz qx qy qz qw" per line (ImageLoader.h:231-259)
// --pipelined uses gsdf_scan_submit/collect (whole frames enqueued, no host round trip inside a frame).
// --threads T (with --pipelined and a PNG dataset) decodes frames ahead on T host threads into pinned buffers
// (prefetch.h): on real data the PNG decode, not the GPU, sets the frame rate.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <deque>
#include <fstream>
#include <iomanip>
#include <sstream>

#include "img_loader.h"
#include "prefetch.h"
#include "sdf_tracker.h"

namespace {
struct Timer {  // Timer.h:59-78
    std::chrono::steady_clock::time_point t0;
    void tic() { t0 = std::chrono::steady_clock::now(); }
    double toc(const std::string& what) {
        const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        std::cout << what << ": " << s * 1e3 << " ms" << std::endl;
        return s;
    }
};
}  // namespace

int main(int argc, char* argv[]) {
    Timer T;
    std::string input = "", output = "../results/", pose_file_name = "pose.txt", stype = "map-gp", dtype = "";
    size_t first = 0, last = std::numeric_limits<size_t>::max();
    float voxel_size = 0.01f, truncation_factor = 5.f;
    bool save_sdf = false, pipelined = false;
    int log2_capacity = 17, threads = 0;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto next = [&](const char* name) -> std::string {
            if (i + 1 >= argc) { std::cerr << name << " needs a value" << std::endl; std::exit(1); }
            return argv[++i];
        };
        if (a == "--input") input = next("--input");
        else if (a == "--results") output = next("--results");
        else if (a == "--pose-file") pose_file_name = next("--pose-file");
        else if (a == "--first") first = std::stoul(next("--first"));
        else if (a == "--last") last = std::stoul(next("--last"));
        else if (a == "--scan-type") stype = next("--scan-type");
        else if (a == "--data-type") dtype = next("--data-type");
        else if (a == "--voxel-size") voxel_size = std::stof(next("--voxel-size"));
        else if (a == "--trunc") truncation_factor = std::stof(next("--trunc"));
        else if (a == "--capacity-log2") log2_capacity = std::stoi(next("--capacity-log2"));
        else if (a == "--save-sdf") save_sdf = true;
        else if (a == "--pipelined") pipelined = true;
        else if (a == "--threads") threads = std::stoi(next("--threads"));
        else { std::cerr << "unknown option " << a << std::endl; return 1; }
    }
    if (stype != "grad-sdf" && stype != "base-sdf") {  // main_scan_3d.cpp:105-114
        std::cerr << "Your specified scan type is not supported (yet)." << std::endl;
        return 1;
    }
    if (!input.empty() && input.back() != '/') input += '/';
    ImageLoader* loader = nullptr;  // main_scan_3d.cpp:116-157
    if (dtype == "tum") loader = new TumrgbdLoader(input);
    else if (dtype == "synth") loader = new SynthLoader(input);
    else if (dtype == "printed") loader = new Printed3dLoader(input);
    else if (dtype == "rw" || dtype == "redwood") loader = new RedwoodLoader(input);
    else if (dtype != "raw") {
        std::cerr << "Your specified dataset type is not supported (yet)." << std::endl;
        return 1;
    }

    // intrinsics (ImageLoader::load_intrinsics)
    Mat3f K;
    {
        std::ifstream f(input + "intrinsics.txt");
        if (!f.is_open()) { std::cerr << "No intrinsics file found in " << input << "!" << std::endl; return 1; }
        for (int i = 0; i < 9; ++i) f >> K.m[i];
    }
    std::cout << "K: " << std::endl;
    for (int i = 0; i < 3; ++i) std::cout << K(i, 0) << " " << K(i, 1) << " " << K(i, 2) << std::endl;
    int width = 0, height = 0;
    size_t n_frames = 0;
    std::ifstream frames;
    if (!loader) {
        std::ifstream f(input + "meta.txt");
        if (!f.is_open() || !(f >> width >> height >> n_frames)) { std::cerr << "No meta.txt in " << input << "!" << std::endl; return 1; }
        frames.open(input + "frames.f32", std::ios::binary);
        if (!frames.is_open()) { std::cerr << "No frames.f32 in " << input << "!" << std::endl; return 1; }
    }

    // ground-truth poses (ImageLoader::load_pose)
    std::vector<SE3> poses;
    std::vector<std::string> stamps;
    bool GT_pose = false;
    {
        std::ifstream f(input + pose_file_name);
        std::string line;
        while (f.is_open() && std::getline(f, line)) {
            std::istringstream ss(line);
            std::string ts;
            float t[3], q[4];
            if (ss >> ts >> t[0] >> t[1] >> t[2] >> q[0] >> q[1] >> q[2] >> q[3]) {
                poses.push_back(SE3(q[3], q[0], q[1], q[2], Vec3f(t[0], t[1], t[2])));
                stamps.push_back(ts);
            }
        }
        if (poses.empty()) std::cerr << "No GT poses are avaible!" << std::endl;
        else { std::cout << poses.size() << " GT poses are loaded!" << std::endl; GT_pose = true; }
    }

    // the estimator needs the frame size: with a PNG dataset it is created when the first frame has been decoded
    // (the reference hard-codes 640x480, main_scan_3d.cpp:182)
    cv::NormalEstimator<float>* NEst = nullptr;
    auto make_estimator = [&]() {
        T.tic();
        NEst = new cv::NormalEstimator<float>(width, height, K, cv::Size(2 * 5 + 1, 2 * 5 + 1));
        T.toc("Init normal estimation");
    };
    if (!loader) make_estimator();
    const float truncation = truncation_factor * voxel_size;  // main_scan_3d.cpp:191

    Sdf* tSDF = nullptr;
    RigidPointOptimizer* pOpt = nullptr;
    std::ofstream pose_file(output + "_poses.txt");
    std::vector<float> depth_buf((size_t)width * height);
    cv::Mat color, depth(height, width, CV_32FC1, depth_buf.data());
    std::string cur_stamp;  // loader->depth_timestamp(), what main_scan_3d.cpp:276 writes in front of every pose
    auto load_next = [&]() -> bool {
        if (loader) {
            if (!loader->load_next(color, depth)) return false;
            cur_stamp = loader->depth_timestamp();
            if (!NEst) { width = depth.cols; height = depth.rows; make_estimator(); }
            if (depth.cols != width || depth.rows != height) { std::cerr << "frame size changed" << std::endl; return false; }
            return true;
        }
        frames.read((char*)depth_buf.data(), (std::streamsize)(depth_buf.size() * sizeof(float)));
        return (size_t)frames.gcount() == depth_buf.size() * sizeof(float);
    };
    const bool prefetched = loader && pipelined && threads > 0 && !GT_pose;
    if (!prefetched)
        for (size_t i = 0; i < first; ++i) load_next();

    double t_track = 0., t_fuse = 0.;
    size_t n_tracked = 0, n_fused = 0;
    auto write_pose = [&](size_t i, const SE3& p) {  // main_scan_3d.cpp:268-280
        // the reference goes through the 4x4 matrix and back: Eigen::Quaternion<float> q(R) (main_scan_3d.cpp:274-276)
        const float* d = p.data();
        float q[4];
        SE3::quaternionFromMatrix(p.rotationMatrix(), q);
        pose_file << (loader ? cur_stamp : (i < stamps.size() ? stamps[i] : std::to_string(i))) << " " << d[0] << " " << d[1] << " " << d[2] << " "
                  << q[0] << " " << q[1] << " " << q[2] << " " << q[3] << "\n";
    };
    int pending = 0;
    std::vector<size_t> pending_idx;
    auto collect_one = [&]() {
        gsdf_frame_result r;
        if (gsdf_scan_collect(tSDF->gpu(), &r) != GSDF_OK) { std::cerr << gsdf_last_error(tSDF->gpu()) << std::endl; std::exit(2); }
        write_pose(pending_idx.front(), SE3::fromStored(r.pose7));
        if (r.converged) std::cout << "... Convergence after " << r.iters << " iterations!" << std::endl;
        if (r.fused) { std::cout << r.points << " points integrated into tSDF volume!" << std::endl; ++n_fused; }
        pending_idx.erase(pending_idx.begin());
        --pending;
    };

    if (prefetched) {
        // Frames are decoded ahead on `threads` host threads into pinned buffers and handed to the GPU pipeline in order;
        // a buffer goes back to the pool when its frame has been collected (its upload is done by then).
        const size_t max_pixels = (size_t)1 << 21;  // up to 1920 x 1080
        FramePrefetcher pf(loader, threads, threads + 6, max_pixels,
                           [](size_t n) { void* p = nullptr; return gsdf_host_alloc(n * sizeof(float), &p) == GSDF_OK ? (float*)p : nullptr; },
                           [](float* p) { gsdf_host_free(p); });
        std::deque<const FramePrefetcher::Frame*> inflight;
        auto collect_frame = [&]() {
            cur_stamp = inflight.front()->stamp_depth;
            collect_one();
            pf.release(inflight.front());
            inflight.pop_front();
        };
        auto t_begin = std::chrono::steady_clock::now();  // restarted once the first frame has created and initialised the map
        size_t n_in = 0;
        for (size_t i = 0; i <= last; ++i) {
            const FramePrefetcher::Frame* f = pf.next();
            if (!f) { if (!pf.error().empty()) std::cerr << pf.error() << std::endl; std::cerr << " -> Frame " << i << " could not be loaded!" << std::endl; break; }
            if (i < first) { pf.release(f); continue; }
            ++n_in;
            if (i == first) {
                width = f->width; height = f->height;
                make_estimator();
                depth = cv::Mat(height, width, CV_32FC1, f->data);
                if (stype == "grad-sdf") tSDF = new MapGradPixelSdf(voxel_size, truncation);
                else tSDF = new MapPixelSdf(voxel_size, truncation);
                tSDF->set_capacity_log2(log2_capacity);
                tSDF->setup(color, depth, K, NEst);
                ++n_fused;
                pOpt = new RigidPointOptimizer(tSDF);
                cur_stamp = f->stamp_depth;
                write_pose(i, pOpt->pose());
                const SE3 start = pOpt->pose();
                gsdf_scan_set_pose(tSDF->gpu(), start.data());
                gsdf_scan_set_tracker(tSDF->gpu(), 25, 1e-3f, 1.f);
                pf.release(f);
                t_begin = std::chrono::steady_clock::now();
                n_in = 0;
                continue;
            }
            if (f->width != width || f->height != height) { std::cerr << "frame size changed" << std::endl; pf.release(f); break; }
            if (pending == 4) collect_frame();
            if (gsdf_scan_submit(tSDF->gpu(), f->data, nullptr, GSDF_SCAN_TRACK_FUSE) != GSDF_OK) {
                std::cerr << gsdf_last_error(tSDF->gpu()) << std::endl;
                return 2;
            }
            inflight.push_back(f);
            pending_idx.push_back(i);
            ++pending;
            ++n_tracked;
        }
        while (pending) collect_frame();
        const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
        std::cout << "prefetched loop: " << n_in << " frames in " << secs << " s = " << n_in / secs << " frames/s after the first frame (decode on "
                  << threads << " threads + upload + track + fuse)" << std::endl;
    } else
    for (size_t i = first; i <= last; ++i) {
        std::cout << "Working on frame: " << i << std::endl;
        if (!load_next()) { std::cerr << " -> Frame " << i << " could not be loaded!" << std::endl; break; }
        if (i == first) {
            T.tic();
            if (stype == "grad-sdf") tSDF = new MapGradPixelSdf(voxel_size, truncation);
            else tSDF = new MapPixelSdf(voxel_size, truncation);
            tSDF->set_capacity_log2(log2_capacity);
            T.toc("Create Sdf");
            T.tic();
            if (GT_pose) tSDF->update(color, depth, K, poses[0], NEst);
            else tSDF->setup(color, depth, K, NEst);
            t_fuse += T.toc("Integrate depth data into Sdf");
            ++n_fused;
            pOpt = new RigidPointOptimizer(tSDF);
            write_pose(i, GT_pose ? poses[0] : pOpt->pose());
            if (pipelined && !GT_pose) {
                const SE3 start = pOpt->pose();
                gsdf_scan_set_pose(tSDF->gpu(), start.data());
                gsdf_scan_set_tracker(tSDF->gpu(), 25, 1e-3f, 1.f);
            }
        } else if (GT_pose) {
            if (i >= poses.size()) break;
            T.tic();
            tSDF->update(color, depth, K, poses[i], NEst);
            t_fuse += T.toc("Integrate depth data into Sdf");
            ++n_fused;
            write_pose(i, poses[i]);
        } else if (pipelined) {
            // the host buffer is reused for the next frame: wait for the copy stream before overwriting it
            if (pending == 4) collect_one();
            if (gsdf_scan_submit(tSDF->gpu(), (const float*)depth.data, nullptr, GSDF_SCAN_TRACK_FUSE) != GSDF_OK) {
                std::cerr << gsdf_last_error(tSDF->gpu()) << std::endl;
                return 2;
            }
            pending_idx.push_back(i);
            ++pending;
            ++n_tracked;
            while (pending) collect_one();  // pageable single buffer: drain before load_next() overwrites it
        } else {
            T.tic();
            const bool conv = pOpt->optimize(depth, K);
            t_track += T.toc("Point optimization");
            ++n_tracked;
            if (conv) {
                T.tic();
                tSDF->update_current(pOpt->pose());  // same depth frame, already on the device
                t_fuse += T.toc("Integrate depth data into Sdf");
                ++n_fused;
            }
            write_pose(i, pOpt->pose());
        }
    }
    while (pending) collect_one();
    pose_file.close();
    std::cout << "tracked " << n_tracked << " frames, fused " << n_fused << "; track " << t_track << " s, fuse " << t_fuse << " s" << std::endl;
    if (!tSDF) return 1;

    const std::string file_prefix = "gradient_sdf";
    T.tic();
    if (!tSDF->extract_mesh(output + file_prefix + "_mesh_final.ply"))
        std::cerr << "Could not save mesh to " << output + file_prefix + "_mesh_final.ply" << "!" << std::endl;
    T.toc("Save mesh to disk");
    T.tic();
    if (!tSDF->extract_pc(output + file_prefix + "_cloud_final.ply"))
        std::cerr << "Could not save point cloud to " << output + file_prefix + "_cloud_final.ply" << "!" << std::endl;
    T.toc("Save point cloud to disk");
    if (save_sdf) {
        T.tic();
        if (!tSDF->save_sdf(output + file_prefix)) std::cerr << "could not save voxel grid info file " << output + file_prefix << "!" << std::endl;
        T.toc("Save sdf txt files to disk");
    }
    delete pOpt;
    delete tSDF;
    delete NEst;
    return 0;
}
