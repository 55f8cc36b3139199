// cloud_b200 runtime metrics: registry + allow-list config + sink client + periodic exporter.
//
// B200-native counterpart of the reference's only native subsystem, the TensorFlow -> Stackdriver
// metrics exporter (src/cpp/monitoring/: stackdriver_client.{h,cc}, stackdriver_config.{h,cc},
// stackdriver_exporter.{h,cc}).  The reference pulls TF's metric registry every 10 s, filters it
// through an allow-list, creates each metric descriptor once and pushes time series over gRPC.
// Here the registry is part of the runtime (the training engine feeds it: step time, all-reduce
// time, bytes reduced, ...), and the sink is local: JSON lines, Prometheus text, or an in-memory
// mock for tests.  The request documents keep the Stackdriver `CreateTimeSeries` /
// `CreateMetricDescriptor` shape, including the histogram -> distribution conversion
// (stackdriver_client.cc:69-98).
#pragma once

#include <atomic>
#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

namespace cb2mon {

// ---- data model (what TF's monitoring::PointSet / MetricDescriptor gave the reference) ---------
enum class MetricKind { kGauge = 0, kCumulative = 1 };
enum class ValueType { kInt64 = 0, kString = 1, kBool = 2, kHistogram = 3, kPercentiles = 4 };

struct Histogram {
  double min = 0, max = 0, num = 0, sum = 0, sum_squares = 0;
  std::vector<double> bucket_limit;   // upper bounds; the last one is +inf (DBL_MAX), like TF
  std::vector<double> bucket;         // counts, same length
};

struct Point {
  ValueType value_type = ValueType::kInt64;
  int64_t int64_value = 0;
  std::string string_value;
  bool bool_value = false;
  Histogram histogram_value;
  uint64_t start_timestamp_millis = 0;
  uint64_t end_timestamp_millis = 0;
};

struct MetricDescriptor {
  std::string name, description;
  std::vector<std::string> label_names;
  MetricKind metric_kind = MetricKind::kGauge;
  ValueType value_type = ValueType::kInt64;
};

struct PointSet {
  std::string metric_name;
  std::vector<Point> points;
};

using PointSetMap = std::map<std::string, PointSet>;
using DescriptorMap = std::map<std::string, MetricDescriptor>;

enum class Code { kOk = 0, kCancelled = 1, kAlreadyExists = 6, kInternal = 13 };
struct Status {
  Code code = Code::kOk;
  std::string message;
  bool ok() const { return code == Code::kOk; }
  static Status OK() { return Status(); }
};

// ---- registry (registry.cc) -----------------------------------------------------------------
// Lock-free updates on the hot path (atomics); the mutex only guards metric creation/collection.
class Registry {
 public:
  static Registry* Default();
  void DefineCounter(const std::string& name, const std::string& description);
  void DefineGauge(const std::string& name, const std::string& description);
  void DefineHistogram(const std::string& name, const std::string& description,
                       const std::vector<double>& bucket_limits);
  void IncrementCounter(const std::string& name, int64_t delta);
  void SetGauge(const std::string& name, int64_t value);
  void Observe(const std::string& name, double value);   // histogram sample
  void CollectMetrics(PointSetMap* points, DescriptorMap* descriptors) const;
  void Reset();                                           // tests
  static std::vector<double> ExponentialBuckets(double scale, double growth, int count);

 private:
  struct Cell;
  Cell* Find(const std::string& name) const;
  Cell* FindOrCreate(const std::string& name, ValueType type, MetricKind kind);
  mutable std::mutex mu_;
  std::map<std::string, std::unique_ptr<Cell>> cells_;
};

// ---- config (config.cc) -----------------------------------------------------------------------
class Config {
 public:
  static const Config* Get();                    // process singleton built from the environment
  explicit Config(const std::string& comma_separated_or_empty);
  bool IsAllowlisted(const std::string& metric_name) const;
  std::string DebugString() const;
  static std::vector<std::string> DefaultMetrics();

 private:
  std::vector<std::string> allowlist_;
};

// ---- sink client (sink_client.cc) -----------------------------------------------------------------
class Sink {   // the "stub": what gRPC MetricService was for the reference
 public:
  virtual ~Sink() = default;
  virtual Status CreateTimeSeries(const std::string& request_json) = 0;
  virtual Status CreateMetricDescriptor(const std::string& request_json) = 0;
};

class MemorySink : public Sink {
 public:
  Status CreateTimeSeries(const std::string& r) override;
  Status CreateMetricDescriptor(const std::string& r) override;
  std::vector<std::string> time_series, descriptors;
  Status next_descriptor_status;   // fault injection
  std::mutex mu;
};

class JsonlFileSink : public Sink {
 public:
  explicit JsonlFileSink(std::string path) : path_(std::move(path)) {}
  Status CreateTimeSeries(const std::string& r) override;
  Status CreateMetricDescriptor(const std::string& r) override;

 private:
  Status Append(const char* kind, const std::string& r);
  std::string path_;
  std::mutex mu_;
};

class PrometheusTextSink : public Sink {   // rewrites <path> with the text exposition format
 public:
  explicit PrometheusTextSink(std::string path) : path_(std::move(path)) {}
  Status CreateTimeSeries(const std::string& r) override;
  Status CreateMetricDescriptor(const std::string&) override { return Status::OK(); }
  void SetLatest(const PointSetMap& points);

 private:
  std::string path_;
  std::mutex mu_;
  PointSetMap latest_;
};

class SinkClient {
 public:
  struct Options { std::string project_id; };
  static SinkClient* Get();   // singleton from env (TFC_MONITORING_PROJECT_ID, TFC_MONITORING_SINK)
  SinkClient(Options options, std::shared_ptr<Sink> sink);    // dependency injection for tests
  Status CreateTimeSeries(const PointSetMap& point_set_map) const;
  Status CreateMetricDescriptor(const MetricDescriptor& descriptor) const;
  std::shared_ptr<Sink> sink() const { return sink_; }
  void set_sink(std::shared_ptr<Sink> s) { sink_ = std::move(s); }
  // exposed for tests
  static std::string TimeSeriesRequestJson(const std::string& project, const PointSetMap& m);
  static std::string DescriptorRequestJson(const std::string& project, const MetricDescriptor& d);

 private:
  Options options_;
  std::shared_ptr<Sink> sink_;
};

// ---- exporter (exporter.cc) --------------------------------------------------------------------
class Exporter {
 public:
  static Exporter* Get();
  // Idempotent and thread-safe; does nothing unless TFC_MONITORING_ENABLED (or the legacy
  // TF_MONITORING_STACKDRIVER_EXPORTER_ENABLED) is true.  Returns true if a thread is running.
  bool PeriodicallyExportMetrics();
  void Stop();
  void ExportMetrics();   // one pull -> filter -> descriptors-once -> push cycle
  int exports() const { return exports_.load(); }
  void ForgetDescriptors();

 private:
  void ExportMetricDescriptors(const DescriptorMap& descriptors);
  std::mutex mu_;
  bool running_ = false;
  std::atomic<bool> stop_{false};
  std::atomic<int> exports_{0};
  std::vector<std::string> exported_descriptors_;
  void* thread_ = nullptr;
};

bool ShouldExport(const Config& config, const PointSet& ps);

}  // namespace cb2mon
