// Sink client (reference: src/cpp/monitoring/stackdriver_client.{h,cc}): converts point sets and
// metric descriptors into CreateTimeSeries / CreateMetricDescriptor request documents and hands
// them to a pluggable sink.  Conversions follow stackdriver_client.cc:63-183: timestamps ms ->
// (seconds, nanos); histogram -> distribution with count, mean = sum/num,
// sum_of_squared_deviation = (sumsq*num - sum^2)/num, explicit bucket bounds WITHOUT the last
// limit, bucket counts; metric type = "custom.googleapis.com" + name; resource "global"; only
// the first point of a point set; labels are STRING; an empty map is CANCELLED.
#include <cinttypes>
#include <cstdio>
#include <cstdlib>

#include "monitoring.h"

namespace cb2mon {

namespace {
constexpr char kMetricTypePrefix[] = "custom.googleapis.com";
constexpr char kDefaultResourceType[] = "global";

std::string Esc(const std::string& s) {
  std::string o;
  for (char c : s) {
    if (c == '"' || c == '\\') { o += '\\'; o += c; }
    else if (c == '\n') o += "\\n";
    else o += c;
  }
  return o;
}
std::string Num(double v) {
  char buf[64];
  if (v == static_cast<int64_t>(v) && v > -1e15 && v < 1e15) snprintf(buf, sizeof(buf), "%.1f", v);
  else snprintf(buf, sizeof(buf), "%.17g", v);
  return buf;
}
std::string Timestamp(uint64_t millis) {
  char buf[96];
  snprintf(buf, sizeof(buf), "{\"seconds\": %" PRIu64 ", \"nanos\": %" PRIu64 "}", millis / 1000,
           (millis % 1000) * 1000000ull);
  return buf;
}
std::string Distribution(const Histogram& h) {
  std::string o = "{\"count\": " + std::to_string(static_cast<int64_t>(h.num));
  double mean = 0.0, ssd = 0.0;
  if (h.num != 0.0) {
    mean = h.sum / h.num;
    ssd = (h.sum_squares * h.num - h.sum * h.sum) / h.num;
  }
  o += ", \"mean\": " + Num(mean) + ", \"sum_of_squared_deviation\": " + Num(ssd);
  o += ", \"bucket_options\": {\"explicit_buckets\": {\"bounds\": [";
  for (size_t i = 0; i + 1 < h.bucket_limit.size(); ++i) {   // the last limit is omitted
    if (i) o += ", ";
    o += Num(h.bucket_limit[i]);
  }
  o += "]}}, \"bucket_counts\": [";
  for (size_t i = 0; i < h.bucket.size(); ++i) {
    if (i) o += ", ";
    o += std::to_string(static_cast<int64_t>(h.bucket[i]));
  }
  return o + "]}";
}
std::string PointJson(const Point& p) {
  std::string v;
  switch (p.value_type) {
    case ValueType::kInt64: v = "{\"int64_value\": " + std::to_string(p.int64_value) + "}"; break;
    case ValueType::kString: v = "{\"string_value\": \"" + Esc(p.string_value) + "\"}"; break;
    case ValueType::kBool: v = std::string("{\"bool_value\": ") + (p.bool_value ? "true" : "false") + "}"; break;
    case ValueType::kHistogram: v = "{\"distribution_value\": " + Distribution(p.histogram_value) + "}"; break;
    case ValueType::kPercentiles: v = "{}"; break;   // ignored, like the reference
  }
  return "{\"interval\": {\"end_time\": " + Timestamp(p.end_timestamp_millis) + "}, \"value\": " + v + "}";
}
}  // namespace

std::string SinkClient::TimeSeriesRequestJson(const std::string& project, const PointSetMap& m) {
  std::string o = "{\"name\": \"projects/" + Esc(project) + "\", \"time_series\": [";
  bool first = true;
  for (const auto& kv : m) {
    if (!first) o += ", ";
    first = false;
    o += "{\"metric\": {\"type\": \"" + std::string(kMetricTypePrefix) + Esc(kv.second.metric_name) +
         "\"}, \"resource\": {\"type\": \"" + kDefaultResourceType + "\"}";
    if (!kv.second.points.empty()) o += ", \"points\": " + PointJson(kv.second.points.front());
    o += "}";
  }
  return o + "]}";
}

std::string SinkClient::DescriptorRequestJson(const std::string& project, const MetricDescriptor& d) {
  const char* kind = d.metric_kind == MetricKind::kGauge ? "GAUGE" : "CUMULATIVE";
  const char* vt = "VALUE_TYPE_UNSPECIFIED";
  switch (d.value_type) {
    case ValueType::kInt64: vt = "INT64"; break;
    case ValueType::kString: vt = "STRING"; break;
    case ValueType::kBool: vt = "BOOL"; break;
    case ValueType::kHistogram: vt = "DISTRIBUTION"; break;
    case ValueType::kPercentiles: break;
  }
  std::string o = "{\"name\": \"projects/" + Esc(project) + "\", \"metric_descriptor\": {\"type\": \"" +
                  std::string(kMetricTypePrefix) + Esc(d.name) + "\", \"description\": \"" +
                  Esc(d.description) + "\", \"labels\": [";
  for (size_t i = 0; i < d.label_names.size(); ++i) {
    if (i) o += ", ";
    o += "{\"key\": \"" + Esc(d.label_names[i]) + "\", \"value_type\": \"STRING\"}";
  }
  o += std::string("], \"metric_kind\": \"") + kind + "\", \"value_type\": \"" + vt + "\"}}";
  return o;
}

SinkClient::SinkClient(Options options, std::shared_ptr<Sink> sink)
    : options_(std::move(options)), sink_(std::move(sink)) {}

SinkClient* SinkClient::Get() {
  static SinkClient* client = [] {
    Options o;
    const char* p = std::getenv("TFC_MONITORING_PROJECT_ID");
    if (!p) p = std::getenv("TF_MONITORING_STACKDRIVER_PROJECT_ID");   // legacy name
    o.project_id = p ? p : "local";
    const char* sink = std::getenv("TFC_MONITORING_SINK");   // jsonl:<path> | prometheus:<path> | memory
    std::string spec = sink ? sink : "jsonl:/tmp/cloud_b200_metrics.jsonl";
    std::shared_ptr<Sink> s;
    if (spec.rfind("prometheus:", 0) == 0) s = std::make_shared<PrometheusTextSink>(spec.substr(11));
    else if (spec == "memory") s = std::make_shared<MemorySink>();
    else s = std::make_shared<JsonlFileSink>(spec.rfind("jsonl:", 0) == 0 ? spec.substr(6) : spec);
    return new SinkClient(o, s);
  }();
  return client;
}

Status SinkClient::CreateTimeSeries(const PointSetMap& point_set_map) const {
  if (point_set_map.empty()) return Status{Code::kCancelled, "no time series to write"};
  if (auto* prom = dynamic_cast<PrometheusTextSink*>(sink_.get())) prom->SetLatest(point_set_map);
  return sink_->CreateTimeSeries(TimeSeriesRequestJson(options_.project_id, point_set_map));
}

Status SinkClient::CreateMetricDescriptor(const MetricDescriptor& descriptor) const {
  return sink_->CreateMetricDescriptor(DescriptorRequestJson(options_.project_id, descriptor));
}

// ---- sinks ------------------------------------------------------------------------------------
Status MemorySink::CreateTimeSeries(const std::string& r) {
  std::lock_guard<std::mutex> l(mu);
  time_series.push_back(r);
  return Status::OK();
}
Status MemorySink::CreateMetricDescriptor(const std::string& r) {
  std::lock_guard<std::mutex> l(mu);
  descriptors.push_back(r);
  Status s = next_descriptor_status;
  next_descriptor_status = Status::OK();
  return s;
}

Status JsonlFileSink::Append(const char* kind, const std::string& r) {
  std::lock_guard<std::mutex> l(mu_);
  FILE* f = fopen(path_.c_str(), "a");
  if (!f) return Status{Code::kInternal, "cannot open " + path_};
  fprintf(f, "{\"call\": \"%s\", \"request\": %s}\n", kind, r.c_str());
  fclose(f);
  return Status::OK();
}
Status JsonlFileSink::CreateTimeSeries(const std::string& r) { return Append("CreateTimeSeries", r); }
Status JsonlFileSink::CreateMetricDescriptor(const std::string& r) { return Append("CreateMetricDescriptor", r); }

void PrometheusTextSink::SetLatest(const PointSetMap& points) {
  std::lock_guard<std::mutex> l(mu_);
  for (const auto& kv : points) latest_[kv.first] = kv.second;
}
Status PrometheusTextSink::CreateTimeSeries(const std::string&) {
  std::lock_guard<std::mutex> l(mu_);
  const std::string tmp = path_ + ".tmp";
  FILE* f = fopen(tmp.c_str(), "w");
  if (!f) return Status{Code::kInternal, "cannot open " + tmp};
  for (const auto& kv : latest_) {
    std::string name = kv.first;
    for (char& c : name) if (!((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9'))) c = '_';
    while (!name.empty() && name[0] == '_') name.erase(0, 1);
    if (kv.second.points.empty()) continue;
    const Point& p = kv.second.points.front();
    if (p.value_type == ValueType::kHistogram) {
      const Histogram& h = p.histogram_value;
      fprintf(f, "# TYPE %s histogram\n", name.c_str());
      double cum = 0;
      for (size_t i = 0; i < h.bucket.size(); ++i) {
        cum += h.bucket[i];
        if (i + 1 < h.bucket.size()) fprintf(f, "%s_bucket{le=\"%g\"} %.0f\n", name.c_str(), h.bucket_limit[i], cum);
        else fprintf(f, "%s_bucket{le=\"+Inf\"} %.0f\n", name.c_str(), cum);
      }
      fprintf(f, "%s_sum %g\n%s_count %.0f\n", name.c_str(), h.sum, name.c_str(), h.num);
    } else {
      fprintf(f, "# TYPE %s gauge\n%s %lld\n", name.c_str(), name.c_str(), static_cast<long long>(p.int64_value));
    }
  }
  fclose(f);
  if (rename(tmp.c_str(), path_.c_str()) != 0) return Status{Code::kInternal, "rename failed"};
  return Status::OK();
}

}  // namespace cb2mon
