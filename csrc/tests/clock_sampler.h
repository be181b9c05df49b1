// SM clock / throttle-reason record for the standalone timing harnesses: a background thread samples NVML
// (loaded with dlopen, so the binaries do not link against it) while a timed section runs, and the section
// prints one "[clocks] ..." line next to its numbers.  Without NVML (CPU box, stripped container) the line
// says so; nothing fails.
#pragma once
#include <dlfcn.h>
#include <stdio.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

namespace clocks {

typedef int (*nvml_init_t)();
typedef int (*nvml_handle_t)(unsigned int, void**);
typedef int (*nvml_clock_t)(void*, int, unsigned int*);
typedef int (*nvml_reasons_t)(void*, unsigned long long*);

struct Nvml {
  void* lib = nullptr;
  void* dev = nullptr;
  nvml_clock_t clock = nullptr, max_clock = nullptr;
  nvml_reasons_t reasons = nullptr;
  bool ok = false;
  explicit Nvml(unsigned int index) {
    lib = dlopen("libnvidia-ml.so.1", RTLD_NOW);
    if (!lib) return;
    nvml_init_t init = reinterpret_cast<nvml_init_t>(dlsym(lib, "nvmlInit_v2"));
    nvml_handle_t handle = reinterpret_cast<nvml_handle_t>(dlsym(lib, "nvmlDeviceGetHandleByIndex_v2"));
    clock = reinterpret_cast<nvml_clock_t>(dlsym(lib, "nvmlDeviceGetClockInfo"));
    max_clock = reinterpret_cast<nvml_clock_t>(dlsym(lib, "nvmlDeviceGetMaxClockInfo"));
    reasons = reinterpret_cast<nvml_reasons_t>(dlsym(lib, "nvmlDeviceGetCurrentClocksEventReasons"));
    if (!reasons) reasons = reinterpret_cast<nvml_reasons_t>(dlsym(lib, "nvmlDeviceGetCurrentClocksThrottleReasons"));
    if (!init || !handle || !clock || !max_clock) return;
    if (init() != 0 || handle(index, &dev) != 0) return;
    ok = true;
  }
};

// NVML clocks-event reason bits that count as a slowdown (nvml.h): sw power cap 0x4, hw slowdown 0x8,
// sw thermal 0x20, hw thermal 0x40, hw power brake 0x80.  Idle (0x1) and application-clock settings (0x2) do not.
constexpr unsigned long long kSlowdownBits = 0x4ull | 0x8ull | 0x20ull | 0x40ull | 0x80ull;

class Sampler {
 public:
  explicit Sampler(unsigned int device_index = 0) : nvml_(device_index) {}
  void start() {
    if (!nvml_.ok) return;
    stop_.store(false);
    samples_.clear();
    reason_bits_ = 0;
    thr_ = std::thread([this] {
      while (!stop_.load()) {
        unsigned int mhz = 0;
        if (nvml_.clock(nvml_.dev, /*NVML_CLOCK_SM*/ 1, &mhz) == 0) samples_.push_back(mhz);
        unsigned long long r = 0;
        if (nvml_.reasons && nvml_.reasons(nvml_.dev, &r) == 0) reason_bits_ |= r;
        std::this_thread::sleep_for(std::chrono::milliseconds(10));
      }
    });
  }
  // Stops sampling and prints the record of the section that just ran.
  void stop_and_print(const char* what) {
    if (!nvml_.ok) {
      printf("[clocks] %s: NVML not available, no clock record\n", what);
      return;
    }
    stop_.store(true);
    if (thr_.joinable()) thr_.join();
    unsigned int mx = 0;
    nvml_.max_clock(nvml_.dev, 1, &mx);
    unsigned int med = 0, lo = 0;
    if (!samples_.empty()) {
      std::sort(samples_.begin(), samples_.end());
      med = samples_[samples_.size() / 2];
      lo = samples_.front();
    }
    const unsigned long long slow = reason_bits_ & kSlowdownBits;
    printf("[clocks] %s: sm_mhz median %u min %u of max %u, %zu samples, reasons 0x%llx%s\n", what, med, lo, mx,
           samples_.size(), reason_bits_,
           slow ? ((slow & ~0x4ull) ? "  <-- SLOWDOWN, re-measure" : "  (sw_power_cap)") : "  (none that slow the SMs)");
  }

 private:
  Nvml nvml_;
  std::thread thr_;
  std::atomic<bool> stop_{true};
  std::vector<unsigned int> samples_;
  unsigned long long reason_bits_ = 0;
};

}  // namespace clocks
