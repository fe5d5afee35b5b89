// Decoding a 640x480 16-bit depth PNG takes a few milliseconds of host time; the GPU needs ~45 microseconds for the
// frame.  On a real dataset the loaders, not the kernels, set the frame rate, so the
// driver decodes ahead on a pool of host threads: frames are ENUMERATED in order by the loader (next_entry, cheap and
// stateful) and DECODED in parallel (ImageLoader::decode, stateless) into a ring of frame buffers that the consumer
// receives in order.  No counterpart in the reference (its loop decodes one frame, then processes it).
//
//   FramePrefetcher pf(loader, threads, slots, alloc, free);   // slots buffers of max_pixels floats each
//   while (const FramePrefetcher::Frame* f = pf.next()) { ... use f->data (f->width x f->height) ...; pf.release(f); }
//
// next() blocks until the next frame IN SEQUENCE ORDER is decoded and returns nullptr at the end of the sequence or at
// the first frame that fails to load (error() then says why, as the loaders' messages do).  A frame's buffer stays
// valid until release(): the GPU pipeline may still be copying from it while later frames are being decoded.
#ifndef GSDF_HOST_PREFETCH_H_
#define GSDF_HOST_PREFETCH_H_

#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>

#include "img_loader.h"

class FramePrefetcher {
public:
    struct Frame {
        float* data = nullptr;
        int width = 0, height = 0;
        size_t index = 0;  // position in the sequence, from 0
        std::string stamp_depth, stamp_rgb;
        // internal
        bool ok = false, ready = false, free_ = true;
        std::string err;
        FrameEntry entry;
    };

    // alloc/release let the caller supply pinned memory (gsdf_host_alloc / gsdf_host_free) without this header knowing CUDA
    FramePrefetcher(ImageLoader* loader, int threads, int slots, size_t max_pixels,
                    std::function<float*(size_t)> alloc = nullptr, std::function<void(float*)> release = nullptr)
        : loader_(loader), max_pixels_(max_pixels), free_fn_(release), slots_((size_t)std::max(2, slots)) {
        for (Frame& f : slots_) f.data = alloc ? alloc(max_pixels) : new float[max_pixels];
        own_ = !alloc;
        for (int i = 0; i < std::max(1, threads); ++i) workers_.emplace_back([this] { work(); });
    }

    ~FramePrefetcher() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_jobs_.notify_all();
        for (std::thread& t : workers_) t.join();
        for (Frame& f : slots_) {
            if (own_) delete[] f.data;
            else if (free_fn_) free_fn_(f.data);
        }
    }

    // The next frame in order, or nullptr (end of sequence, or a frame could not be loaded: see error()).
    const Frame* next() {
        std::unique_lock<std::mutex> lk(mu_);
        fill(lk);
        if (pending_.empty()) return nullptr;
        Frame* f = pending_.front();
        cv_done_.wait(lk, [&] { return f->ready; });
        pending_.pop_front();
        if (!f->ok) {
            err_ = f->err;
            f->free_ = true;
            failed_ = true;  // nothing after a failed frame is delivered (the reference's loop breaks there)
            return nullptr;
        }
        return f;
    }

    void release(const Frame* f) {
        std::unique_lock<std::mutex> lk(mu_);
        const_cast<Frame*>(f)->free_ = true;
        fill(lk);
    }

    const std::string& error() const { return err_; }

private:
    // hand every free buffer to the pool, in sequence order (called with the lock held)
    void fill(std::unique_lock<std::mutex>&) {
        while (!ended_ && !failed_) {
            Frame* slot = nullptr;
            for (Frame& f : slots_)
                if (f.free_) { slot = &f; break; }
            if (!slot) break;
            FrameEntry e;
            if (!loader_->next_entry(e)) { ended_ = true; break; }
            slot->free_ = false; slot->ready = false; slot->ok = false; slot->err.clear();
            slot->entry = e; slot->index = issued_++;
            slot->stamp_depth = e.stamp_depth; slot->stamp_rgb = e.stamp_rgb;
            pending_.push_back(slot);
            jobs_.push_back(slot);
            cv_jobs_.notify_one();
        }
    }

    void work() {
        for (;;) {
            Frame* f = nullptr;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_jobs_.wait(lk, [&] { return stop_ || !jobs_.empty(); });
                if (stop_) return;
                f = jobs_.front();
                jobs_.pop_front();
            }
            std::string err;
            int w = 0, h = 0;
            const bool ok = loader_->decode(f->entry, f->data, max_pixels_, nullptr, w, h, err);
            {
                std::lock_guard<std::mutex> lk(mu_);
                f->ok = ok; f->err = err; f->width = w; f->height = h; f->ready = true;
            }
            cv_done_.notify_all();
        }
    }

    ImageLoader* loader_;
    size_t max_pixels_;
    std::function<void(float*)> free_fn_;
    bool own_ = true;
    std::vector<Frame> slots_;
    std::deque<Frame*> pending_, jobs_;  // in sequence order: issued and not yet delivered / not yet taken by a worker
    std::vector<std::thread> workers_;
    std::mutex mu_;
    std::condition_variable cv_jobs_, cv_done_;
    size_t issued_ = 0;
    bool stop_ = false, ended_ = false, failed_ = false;
    std::string err_;
};

#endif  // GSDF_HOST_PREFETCH_H_
