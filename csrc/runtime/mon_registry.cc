// Metric registry: the piece TensorFlow's monitoring::CollectionRegistry provided to the reference
// exporter (stackdriver_exporter.cc:87-91).  Counters and gauges are single atomics; histograms
// keep per-bucket atomic counts plus CAS-updated sum / sum of squares, so the training hot path
// never takes a lock.
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstring>

#include "monitoring.h"
#include "cloud_b200_runtime.h"   // the public declarations must match the definitions below

namespace cb2mon {

namespace {
uint64_t NowMillis() {
  return std::chrono::duration_cast<std::chrono::milliseconds>(
             std::chrono::system_clock::now().time_since_epoch()).count();
}
void AtomicAddDouble(std::atomic<uint64_t>* cell, double delta) {
  uint64_t old_bits = cell->load(std::memory_order_relaxed);
  for (;;) {
    double old_val;
    memcpy(&old_val, &old_bits, 8);
    const double new_val = old_val + delta;
    uint64_t new_bits;
    memcpy(&new_bits, &new_val, 8);
    if (cell->compare_exchange_weak(old_bits, new_bits, std::memory_order_relaxed)) return;
  }
}
double LoadDouble(const std::atomic<uint64_t>& cell) {
  const uint64_t bits = cell.load(std::memory_order_relaxed);
  double v;
  memcpy(&v, &bits, 8);
  return v;
}
}  // namespace

struct Registry::Cell {
  MetricDescriptor descriptor;
  std::atomic<int64_t> value{0};
  std::vector<double> limits;
  std::unique_ptr<std::atomic<uint64_t>[]> buckets;   // counts as integers
  std::atomic<uint64_t> num{0}, sum_bits{0}, sumsq_bits{0};
  uint64_t created_ms = NowMillis();
};

Registry* Registry::Default() {
  static Registry* r = new Registry();
  return r;
}

std::vector<double> Registry::ExponentialBuckets(double scale, double growth, int count) {
  std::vector<double> out;
  double v = scale;
  for (int i = 0; i < count - 1; ++i) { out.push_back(v); v *= growth; }
  out.push_back(DBL_MAX);
  return out;
}

Registry::Cell* Registry::Find(const std::string& name) const {
  std::lock_guard<std::mutex> l(mu_);
  auto it = cells_.find(name);
  return it == cells_.end() ? nullptr : it->second.get();
}

Registry::Cell* Registry::FindOrCreate(const std::string& name, ValueType type, MetricKind kind) {
  std::lock_guard<std::mutex> l(mu_);
  auto it = cells_.find(name);
  if (it != cells_.end()) return it->second.get();
  auto cell = std::make_unique<Cell>();
  cell->descriptor.name = name;
  cell->descriptor.value_type = type;
  cell->descriptor.metric_kind = kind;
  if (type == ValueType::kHistogram) {
    cell->limits = ExponentialBuckets(1.0, 2.0, 32);
    cell->buckets.reset(new std::atomic<uint64_t>[cell->limits.size()]());
  }
  Cell* raw = cell.get();
  cells_[name] = std::move(cell);
  return raw;
}

void Registry::DefineCounter(const std::string& name, const std::string& description) {
  FindOrCreate(name, ValueType::kInt64, MetricKind::kCumulative)->descriptor.description = description;
}
void Registry::DefineGauge(const std::string& name, const std::string& description) {
  FindOrCreate(name, ValueType::kInt64, MetricKind::kGauge)->descriptor.description = description;
}
void Registry::DefineHistogram(const std::string& name, const std::string& description,
                               const std::vector<double>& bucket_limits) {
  Cell* c = FindOrCreate(name, ValueType::kHistogram, MetricKind::kCumulative);
  std::lock_guard<std::mutex> l(mu_);
  c->descriptor.description = description;
  if (!bucket_limits.empty() && c->num.load() == 0) {
    c->limits = bucket_limits;
    if (c->limits.back() < DBL_MAX) c->limits.push_back(DBL_MAX);
    c->buckets.reset(new std::atomic<uint64_t>[c->limits.size()]());
  }
}

void Registry::IncrementCounter(const std::string& name, int64_t delta) {
  Cell* c = Find(name);
  if (!c) c = FindOrCreate(name, ValueType::kInt64, MetricKind::kCumulative);
  c->value.fetch_add(delta, std::memory_order_relaxed);
}
void Registry::SetGauge(const std::string& name, int64_t value) {
  Cell* c = Find(name);
  if (!c) c = FindOrCreate(name, ValueType::kInt64, MetricKind::kGauge);
  c->value.store(value, std::memory_order_relaxed);
}
void Registry::Observe(const std::string& name, double value) {
  Cell* c = Find(name);
  if (!c) c = FindOrCreate(name, ValueType::kHistogram, MetricKind::kCumulative);
  if (c->descriptor.value_type != ValueType::kHistogram) return;
  size_t b = 0;
  while (b + 1 < c->limits.size() && value >= c->limits[b]) ++b;
  c->buckets[b].fetch_add(1, std::memory_order_relaxed);
  c->num.fetch_add(1, std::memory_order_relaxed);
  AtomicAddDouble(&c->sum_bits, value);
  AtomicAddDouble(&c->sumsq_bits, value * value);
}

void Registry::CollectMetrics(PointSetMap* points, DescriptorMap* descriptors) const {
  std::lock_guard<std::mutex> l(mu_);
  const uint64_t now = NowMillis();
  for (const auto& kv : cells_) {
    const Cell& c = *kv.second;
    PointSet ps;
    ps.metric_name = kv.first;
    Point p;
    p.value_type = c.descriptor.value_type;
    p.start_timestamp_millis = c.created_ms;
    p.end_timestamp_millis = now;
    if (p.value_type == ValueType::kHistogram) {
      Histogram& h = p.histogram_value;
      h.num = static_cast<double>(c.num.load(std::memory_order_relaxed));
      h.sum = LoadDouble(c.sum_bits);
      h.sum_squares = LoadDouble(c.sumsq_bits);
      h.bucket_limit = c.limits;
      for (size_t i = 0; i < c.limits.size(); ++i)
        h.bucket.push_back(static_cast<double>(c.buckets[i].load(std::memory_order_relaxed)));
    } else {
      p.int64_value = c.value.load(std::memory_order_relaxed);
    }
    ps.points.push_back(std::move(p));
    (*points)[kv.first] = std::move(ps);
    if (descriptors) (*descriptors)[kv.first] = c.descriptor;
  }
}

void Registry::Reset() {
  std::lock_guard<std::mutex> l(mu_);
  cells_.clear();
}

}  // namespace cb2mon
