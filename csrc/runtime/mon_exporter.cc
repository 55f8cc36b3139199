// Periodic exporter (reference: src/cpp/monitoring/stackdriver_exporter.{h,cc}).
// PeriodicallyExportMetrics() is idempotent, thread-safe and environment-gated (default off);
// every period (10 s upstream, stackdriver_exporter.cc:28) ExportMetrics() collects the registry,
// keeps allow-listed NON-EMPTY point sets (int64 != 0 or histogram num > 0, :38-70), creates each
// metric descriptor once (ALREADY_EXISTS counts as success, :105-126) and writes the time series.
// Also holds the C ABI used by the Python binding (cloud_b200/monitoring).
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <strings.h>
#include <thread>

#include "monitoring.h"
#include "cloud_b200_runtime.h"   // the public declarations must match the definitions below
#include "rt_common.h"

namespace cb2mon {

namespace {
bool EnvBool(const char* name, bool dflt) {
  const char* v = std::getenv(name);
  if (!v) return dflt;
  return !strcasecmp(v, "1") || !strcasecmp(v, "true") || !strcasecmp(v, "yes") || !strcasecmp(v, "on");
}
int IntervalMs() {
  const char* v = std::getenv("TFC_MONITORING_INTERVAL_MS");
  const int ms = v ? atoi(v) : 10000;
  return ms > 0 ? ms : 10000;
}
}  // namespace

bool ShouldExport(const Config& config, const PointSet& ps) {
  if (!config.IsAllowlisted(ps.metric_name)) return false;
  if (ps.points.empty()) return false;
  const Point& p = ps.points.front();
  if (p.value_type == ValueType::kInt64) return p.int64_value != 0;
  if (p.value_type == ValueType::kHistogram) return p.histogram_value.num > 0;
  return false;
}

Exporter* Exporter::Get() {
  static Exporter* e = new Exporter();
  return e;
}

bool Exporter::PeriodicallyExportMetrics() {
  const bool enabled = EnvBool("TFC_MONITORING_ENABLED",
                               EnvBool("TF_MONITORING_STACKDRIVER_EXPORTER_ENABLED", false));
  if (!enabled) return false;
  std::lock_guard<std::mutex> l(mu_);
  if (running_) return true;
  stop_.store(false);
  const int interval = IntervalMs();
  auto* th = new std::thread([this, interval] {
    while (!stop_.load()) {
      for (int waited = 0; waited < interval && !stop_.load(); waited += 10)
        std::this_thread::sleep_for(std::chrono::milliseconds(10));
      if (stop_.load()) break;
      ExportMetrics();
    }
  });
  thread_ = th;
  running_ = true;
  return true;
}

void Exporter::Stop() {
  std::thread* th = nullptr;
  {
    std::lock_guard<std::mutex> l(mu_);
    if (!running_) return;
    stop_.store(true);
    th = static_cast<std::thread*>(thread_);
    thread_ = nullptr;
    running_ = false;
  }
  if (th) { th->join(); delete th; }
}

void Exporter::ExportMetrics() {
  PointSetMap all, filtered;
  DescriptorMap descriptors, wanted;
  Registry::Default()->CollectMetrics(&all, &descriptors);
  const Config* config = Config::Get();
  for (const auto& kv : all) {
    if (ShouldExport(*config, kv.second)) {
      filtered[kv.first] = kv.second;
      wanted[kv.first] = descriptors[kv.first];
    }
  }
  ExportMetricDescriptors(wanted);
  Status s = SinkClient::Get()->CreateTimeSeries(filtered);
  if (!s.ok() && s.code != Code::kCancelled)
    cb2rt::set_error("CreateTimeSeries failed: %s", s.message.c_str());
  exports_.fetch_add(1);
}

void Exporter::ExportMetricDescriptors(const DescriptorMap& descriptors) {
  std::lock_guard<std::mutex> l(mu_);
  for (const auto& kv : descriptors) {
    bool done = false;
    for (const auto& n : exported_descriptors_) if (n == kv.first) { done = true; break; }
    if (done) continue;
    Status s = SinkClient::Get()->CreateMetricDescriptor(kv.second);
    if (s.ok() || s.code == Code::kAlreadyExists) exported_descriptors_.push_back(kv.first);
    else cb2rt::set_error("CreateMetricDescriptor failed for %s: %s", kv.first.c_str(), s.message.c_str());
  }
}

void Exporter::ForgetDescriptors() {
  std::lock_guard<std::mutex> l(mu_);
  exported_descriptors_.clear();
}

}  // namespace cb2mon

// ---- C ABI ------------------------------------------------------------------------------------
using namespace cb2mon;

extern "C" {

void cb2mon_define_counter(const char* name, const char* desc) { Registry::Default()->DefineCounter(name, desc ? desc : ""); }
void cb2mon_define_gauge(const char* name, const char* desc) { Registry::Default()->DefineGauge(name, desc ? desc : ""); }
void cb2mon_define_histogram(const char* name, const char* desc, const double* limits, int n) {
  std::vector<double> v(limits, limits + (limits ? n : 0));
  Registry::Default()->DefineHistogram(name, desc ? desc : "", v);
}
void cb2mon_counter_inc(const char* name, long long delta) { Registry::Default()->IncrementCounter(name, delta); }
void cb2mon_gauge_set(const char* name, long long value) { Registry::Default()->SetGauge(name, value); }
void cb2mon_observe(const char* name, double value) { Registry::Default()->Observe(name, value); }
void cb2mon_reset() { Registry::Default()->Reset(); Exporter::Get()->ForgetDescriptors(); }
int cb2mon_start() { return Exporter::Get()->PeriodicallyExportMetrics() ? 1 : 0; }
void cb2mon_stop() { Exporter::Get()->Stop(); }
void cb2mon_export_now() { Exporter::Get()->ExportMetrics(); }
int cb2mon_export_count() { return Exporter::Get()->exports(); }
int cb2mon_is_allowlisted(const char* name) { return Config::Get()->IsAllowlisted(name) ? 1 : 0; }

static int copy_str(const std::string& s, char* buf, size_t n) {
  if (s.size() + 1 > n) return -static_cast<int>(s.size() + 1);
  memcpy(buf, s.c_str(), s.size() + 1);
  return static_cast<int>(s.size());
}
int cb2mon_config_debug_string(char* buf, size_t n) { return copy_str(Config::Get()->DebugString(), buf, n); }

// kind: 0 = jsonl file, 1 = prometheus text file, 2 = in-memory (tests)
int cb2mon_set_sink(int kind, const char* path) {
  std::shared_ptr<Sink> s;
  if (kind == 0) s = std::make_shared<JsonlFileSink>(path ? path : "/tmp/cloud_b200_metrics.jsonl");
  else if (kind == 1) s = std::make_shared<PrometheusTextSink>(path ? path : "/tmp/cloud_b200_metrics.prom");
  else s = std::make_shared<MemorySink>();
  SinkClient::Get()->set_sink(s);
  return 0;
}

// newline-separated requests seen by the in-memory sink; which: 0 = time series, 1 = descriptors
int cb2mon_memory_sink_dump(int which, char* buf, size_t n) {
  auto* m = dynamic_cast<MemorySink*>(SinkClient::Get()->sink().get());
  if (!m) return -1;
  std::string out;
  std::lock_guard<std::mutex> l(m->mu);
  for (const auto& r : (which == 0 ? m->time_series : m->descriptors)) out += r + "\n";
  return copy_str(out, buf, n);
}

}  // extern "C"
