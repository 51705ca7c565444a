// Library-wide state: error string, launch counter, version.
#include "common.cuh"

#include <mutex>

namespace xg {

std::atomic<uint64_t> g_launches{0};

namespace {
std::mutex g_err_mu;
std::string g_err;
thread_local std::string t_err_copy;
}  // namespace

void set_error(const std::string &msg) {
  std::lock_guard<std::mutex> lk(g_err_mu);
  g_err = msg;
}

}  // namespace xg

extern "C" {

const char *xg_last_error(void) {
  std::lock_guard<std::mutex> lk(xg::g_err_mu);
  xg::t_err_copy = xg::g_err;
  return xg::t_err_copy.c_str();
}

int xg_version(void) { return 100; }

uint64_t xg_launch_count(void) { return xg::g_launches.load(std::memory_order_relaxed); }

}  // extern "C"
