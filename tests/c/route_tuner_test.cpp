// CPU unit test of cm::RouteTuner (cassmantle_b200/csrc/cm_route_tuner.h) with a fake clock.
// Built and run by tests/test_route_tuner.py; exit code 0 = all checks passed.
#include <cstdio>
#include <cstdlib>

#include "cm_route_tuner.h"

using cm::RouteTuner;

#define CHECK(cond)                                                                  \
    do {                                                                             \
        if (!(cond)) { std::printf("FAILED line %d: %s\n", __LINE__, #cond); std::exit(1); } \
    } while (0)

// Drives the tuner for `submits` mid-size submits; a submit that goes in place takes `t_in`
// seconds, one through the copy engine `t_copy`.  Returns how many went in place.
static long drive(RouteTuner &t, double &clock, long submits, double t_in, double t_copy) {
    long in_place = 0;
    for (long i = 0; i < submits; ++i) {
        const bool r = t.next(clock);
        in_place += r ? 1 : 0;
        clock += r ? t_in : t_copy;
    }
    return in_place;
}

int main() {
    constexpr long kTrial = RouteTuner::kSettle + RouteTuner::kWindow;
    constexpr long kCycle = 2 * kTrial + RouteTuner::kHoldSubmits;
    {   // pinned routes never change and never look at the clock
        RouteTuner t; t.pin(RouteTuner::kCopy);
        double c = 0; CHECK(drive(t, c, 5000, 1.0, 1.0) == 0);
        RouteTuner u; u.pin(RouteTuner::kInPlace);
        CHECK(drive(u, c, 5000, 1.0, 1.0) == 5000);
    }
    {   // start-up and the first trial: warm in place, a window in place, a window through the copy engine
        RouteTuner t; double c = 100.0;
        CHECK(drive(t, c, RouteTuner::kWarmSubmits, 1e-5, 2e-5) == RouteTuner::kWarmSubmits);
        CHECK(drive(t, c, kTrial, 1e-5, 2e-5) == kTrial);
        CHECK(drive(t, c, kTrial, 1e-5, 2e-5) == 0);
        // in place was twice as fast: it serves the whole hold period
        CHECK(t.best_in_place());
        CHECK(drive(t, c, RouteTuner::kHoldSubmits, 1e-5, 2e-5) == RouteTuner::kHoldSubmits);
        // the timed windows cover kWindow submits each (the first one is stamped at submit
        // kSettle and closed at submit kSettle + kWindow), whatever the route costs
        const double w_in = t.window_seconds(true), w_cp = t.window_seconds(false);
        CHECK(w_in > 255.5e-5 && w_in < 256.5e-5);
        CHECK(w_cp > 511e-5 && w_cp < 513e-5);
    }
    {   // the copy engine is 20 % faster: chosen, kept for the hold period, then tried again
        RouteTuner t; double c = 0.0;
        drive(t, c, RouteTuner::kWarmSubmits + 2 * kTrial, 12e-6, 10e-6);
        CHECK(!t.best_in_place());
        CHECK(drive(t, c, RouteTuner::kHoldSubmits, 12e-6, 10e-6) == 0);
        CHECK(drive(t, c, kTrial, 12e-6, 10e-6) == kTrial);         // re-trial starts in place
        CHECK(drive(t, c, kTrial, 12e-6, 10e-6) == 0);
        CHECK(!t.best_in_place());
    }
    {   // a 2 % edge is not enough (hysteresis), 4 % is
        RouteTuner t; double c = 0.0;
        drive(t, c, RouteTuner::kWarmSubmits + 2 * kTrial, 10.0e-6, 9.8e-6);
        CHECK(t.best_in_place());
        RouteTuner u; c = 0.0;
        drive(u, c, RouteTuner::kWarmSubmits + 2 * kTrial, 10.0e-6, 9.6e-6);
        CHECK(!u.best_in_place());
    }
    {   // conditions change between trials (more GPUs start pulling on the same PCIe path)
        RouteTuner t; double c = 0.0;
        drive(t, c, RouteTuner::kWarmSubmits + kCycle, 10e-6, 11e-6);
        CHECK(t.best_in_place());
        drive(t, c, 2 * kTrial, 14e-6, 11e-6);
        CHECK(!t.best_in_place());
        long in_place = drive(t, c, RouteTuner::kHoldSubmits, 14e-6, 11e-6);
        CHECK(in_place == 0);
    }
    {   // a producer-bound caller (both routes take the producer's time): stays in place
        RouteTuner t; double c = 0.0;
        drive(t, c, RouteTuner::kWarmSubmits + 2 * kTrial, 50e-6, 50e-6);
        CHECK(t.best_in_place());
    }
    {   // idle gaps (bursty traffic) never decide a trial: a window that contains one starts again
        RouteTuner t; double c = 0.0;
        drive(t, c, RouteTuner::kWarmSubmits, 10e-6, 10e-6);
        // in-place window with a long pause in the middle, copy window without: without the
        // restart the copy engine would "win" by the length of the pause
        drive(t, c, kTrial / 2, 10e-6, 10e-6);
        c += 0.5;
        CHECK(drive(t, c, kTrial, 10e-6, 10e-6) == kTrial);       // the in-place window ran again in full
        CHECK(drive(t, c, kTrial, 10e-6, 10e-6) == 0);
        CHECK(t.best_in_place());
        CHECK(t.window_seconds(true) < 300e-5);
        // a caller that pauses every 100 submits never finishes a trial and keeps its route
        RouteTuner u; c = 0.0;
        long in_place = 0;
        for (int burst = 0; burst < 200; ++burst) { in_place += drive(u, c, 100, 10e-6, 5e-6); c += 0.01; }
        CHECK(in_place == 200 * 100);
    }
    {   // over a long run the losing route gets only the trial windows
        RouteTuner t; double c = 0.0;
        const long total = RouteTuner::kWarmSubmits + 5 * kCycle;
        const long in_place = drive(t, c, total, 10e-6, 13e-6);
        CHECK(total - in_place == 5 * kTrial);
    }
    std::printf("route tuner: all checks passed\n");
    return 0;
}
