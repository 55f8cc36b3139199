// Metric allow-list (reference: src/cpp/monitoring/stackdriver_config.{h,cc}): a process singleton
// built from a comma-separated environment variable, with defaults when it is unset.  The
// reference's defaults are TensorFlow runtime metrics (graph optimisation time, graph run-time
// histogram, tf.data fetch statistics, stackdriver_config.cc:31-50); this framework's defaults
// are its own hot-path metrics.
#include <cstdlib>
#include <sstream>

#include "monitoring.h"

namespace cb2mon {

std::vector<std::string> Config::DefaultMetrics() {
  return {
      "/cloud_b200/train/step_time_usecs",         // histogram: device time of one training step
      "/cloud_b200/comm/allreduce_usecs",          // histogram: fused all-reduce+optimizer kernels
      "/cloud_b200/comm/bytes_reduced",            // counter
      "/cloud_b200/comm/bucket_launches",          // counter
      "/cloud_b200/data/wait_usecs",               // histogram: host waited for the input pipeline
      "/cloud_b200/train/images_per_sec",          // gauge
  };
}

Config::Config(const std::string& csv) {
  if (csv.empty()) {
    allowlist_ = DefaultMetrics();
    return;
  }
  std::stringstream ss(csv);
  std::string item;
  while (std::getline(ss, item, ',')) {
    size_t a = item.find_first_not_of(" \t");
    size_t b = item.find_last_not_of(" \t");
    if (a != std::string::npos) allowlist_.push_back(item.substr(a, b - a + 1));
  }
}

const Config* Config::Get() {
  static const Config* config = [] {
    const char* v = std::getenv("TFC_MONITORING_ALLOWLIST");
    if (!v) v = std::getenv("TF_MONITORING_STACKDRIVER_METRICS_WHITELIST");   // legacy name
    return new Config(v ? v : "");
  }();
  return config;
}

bool Config::IsAllowlisted(const std::string& metric_name) const {
  for (const auto& m : allowlist_)
    if (m == metric_name) return true;
  return false;
}

std::string Config::DebugString() const {
  std::string out = "allowlist: [";
  for (size_t i = 0; i < allowlist_.size(); ++i) {
    if (i) out += ", ";
    out += allowlist_[i];
  }
  return out + "]";
}

}  // namespace cb2mon
