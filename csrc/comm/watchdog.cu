// Host side of the signal-wait watchdog (device side: comm/symm.cuh).  Each kernel module that spins on peer
// flags registers a reader here; `llmctl_comm_watchdog_poll` gathers their records so one call tells the host
// whether any wait on this device gave up, and on what.
//
// Reference: the reference has no collective code and no hang detection.
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

#include "../common/util.cuh"
#include "symm.cuh"

namespace llmctl {
namespace {
struct Module {
  const char* name;
  WatchdogReadFn rd;
  WatchdogTimeoutFn st;
};
std::vector<Module>& modules() {
  static std::vector<Module> m;
  return m;
}
std::mutex& mu() {
  static std::mutex m;
  return m;
}
}  // namespace

void register_watchdog_module(const char* name, WatchdogReadFn rd, WatchdogTimeoutFn st) {
  std::lock_guard<std::mutex> g(mu());
  modules().push_back({name, rd, st});
}
}  // namespace llmctl

using namespace llmctl;

// Synchronous (it copies from device symbols): call it where the host waits for the device anyway.
// out[0] = total timed-out waits across modules since the last clearing poll; for the first module that tripped:
// out[1] expected epoch, out[2] last value seen, out[3] blockIdx.x, out[4] flag address, out[5] ns waited,
// out[6] index of that module in registration order.  `module_name` (may be NULL) receives its name.
LLMCTL_API int llmctl_comm_watchdog_poll(unsigned long long* out, char* module_name, int name_cap, int clear) {
  LLMCTL_BEGIN();
  unsigned long long total = 0;
  bool have = false;
  std::lock_guard<std::mutex> g(mu());
  for (size_t i = 0; i < modules().size(); ++i) {
    symm::WatchdogRecord r{};
    const int rc = modules()[i].rd(&r, clear);
    if (rc) LLMCTL_FAIL(rc, "watchdog record read");
    total += r.tripped;
    if (r.tripped && !have) {
      have = true;
      out[1] = r.expected;
      out[2] = r.seen;
      out[3] = r.block;
      out[4] = r.flag_addr;
      out[5] = r.waited_ns;
      out[6] = i;
      if (module_name && name_cap > 0) {
        std::strncpy(module_name, modules()[i].name, name_cap - 1);
        module_name[name_cap - 1] = 0;
      }
    }
  }
  out[0] = total;
  return 0;
}

// Deadline of every peer-flag wait on the current device, in milliseconds (default 30 s).
LLMCTL_API int llmctl_comm_watchdog_set_timeout_ms(unsigned long long ms) {
  LLMCTL_BEGIN();
  std::lock_guard<std::mutex> g(mu());
  for (auto& m : modules()) {
    const int rc = m.st(ms * 1000000ull);
    if (rc) LLMCTL_FAIL(rc, "watchdog timeout set");
  }
  return 0;
}

LLMCTL_API int llmctl_comm_watchdog_modules() {
  std::lock_guard<std::mutex> g(mu());
  return static_cast<int>(modules().size());
}
