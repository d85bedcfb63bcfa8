// cassmantle_b200 -- which way a mid-size batch travels through a cm_pipe (cm_pipe.cu).
//
// Plain C++ (no CUDA) so that tests/c/route_tuner_test.cpp can drive it with a fake clock.
//
// A pipe can score a batch "in place" (the kernel reads/writes the page-locked slot across PCIe)
// or through the copy engine (H2D, kernel, D2H).  Which is faster for batches of a few hundred
// KB depends on how busy the PCIe path is -- in place wins by 5 % on one B200, loses by 8-19 %
// when eight GPUs pull 26 GB/s each through one host -- so the pipe measures it:
//   warm      the first kWarmSubmits mid-size submits go in place, untimed
//   trial     kSettle + kWindow submits in place, then the same through the copy engine; the
//             first kSettle of each are not timed (batches submitted on the other route are
//             still draining), the remaining kWindow are, on the caller's clock
//   hold      the winner serves kHoldSubmits submits, then the trial is repeated
// The copy engine has to win by 3 %: when the producer is the bottleneck both windows measure
// the producer, and in place costs the host one call per step instead of three.
// The windows are wall time between submits, so they mean "route cost" only while the caller keeps
// the pipe busy.  A gap of more than kIdleGapS between two submits (two orders of magnitude above
// a mid-size step) is an idle caller, not a slow route: the current trial window starts again, so
// bursty or light traffic never completes a trial and stays on the route it had (in place at first).
#pragma once

namespace cm {

class RouteTuner {
public:
    enum Route { kAuto, kInPlace, kCopy };
    static constexpr int kWarmSubmits = 64, kSettle = 32, kWindow = 256, kHoldSubmits = 16384;
    static constexpr double kIdleGapS = 2e-3;

    void pin(Route r) { pinned_ = r; }
    Route pinned() const { return pinned_; }

    // Called once per mid-size submit with the current time in seconds (any monotonic origin).
    // Returns true when this batch is to go in place, false for the copy engine.
    bool next(double now_s) {
        if (pinned_ != kAuto) return pinned_ == kInPlace;
        const double gap = now_s - last_s_;
        last_s_ = now_s;
        if ((phase_ == kTryInPlace || phase_ == kTryCopy) && count_ > 0 && gap > kIdleGapS) count_ = 0;
        const int n = ++count_;
        switch (phase_) {
            case kWarm:
                if (n >= kWarmSubmits) enter(kTryInPlace);
                return true;
            case kTryInPlace:
            case kTryCopy: {
                const bool in_place = phase_ == kTryInPlace;
                if (n == kSettle) window_start_ = now_s;
                if (n == kSettle + kWindow) {
                    window_s_[in_place ? 0 : 1] = now_s - window_start_;
                    if (in_place) {
                        enter(kTryCopy);
                    } else {
                        best_in_place_ = !(window_s_[1] < 0.97 * window_s_[0]);
                        enter(kHold);
                    }
                }
                return in_place;
            }
            case kHold:
                if (n >= kHoldSubmits) enter(kTryInPlace);
                return best_in_place_;
        }
        return true;
    }

    bool best_in_place() const { return best_in_place_; }
    double window_seconds(bool in_place) const { return window_s_[in_place ? 0 : 1]; }

private:
    enum Phase { kWarm, kTryInPlace, kTryCopy, kHold };
    void enter(Phase p) { phase_ = p; count_ = 0; }

    Route pinned_ = kAuto;
    Phase phase_ = kWarm;
    int count_ = 0;
    bool best_in_place_ = true;
    double window_start_ = 0.0, last_s_ = 0.0;
    double window_s_[2] = {0.0, 0.0};          // [in place, copy engine]
};

}  // namespace cm
