// C++ unit test (reference: src/cpp/monitoring/stackdriver_client_test.cc: mock stub injected
// through the public constructor, golden request documents).  Same golden cases: an int64 point
// 33 and a 2-bucket histogram holding {-1.0, 1.0} at 12345 ms -> count 2, mean 0, SSD 2,
// bounds [0], counts [1, 1], end_time 12 s + 345 000 000 ns; plus config / filter / exporter
// behaviour the reference leaves untested.
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <chrono>
#include <thread>

#include "monitoring.h"

using namespace cb2mon;

static int failures = 0;
#define CHECK(cond)                                                         \
  do {                                                                      \
    if (!(cond)) { printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); ++failures; } \
  } while (0)
#define CHECK_HAS(hay, needle)                                              \
  do {                                                                      \
    if ((hay).find(needle) == std::string::npos) {                          \
      printf("FAIL %s:%d: missing %s in\n  %s\n", __FILE__, __LINE__, needle, (hay).c_str()); ++failures; } \
  } while (0)

static PointSetMap GoldenPoints() {
  PointSetMap m;
  PointSet a;
  a.metric_name = "/metric_1";
  Point p;
  p.value_type = ValueType::kInt64;
  p.int64_value = 33;
  p.end_timestamp_millis = 12345;
  a.points.push_back(p);
  m["/metric_1"] = a;
  PointSet b;
  b.metric_name = "/metric_2";
  Point h;
  h.value_type = ValueType::kHistogram;
  h.end_timestamp_millis = 12345;
  h.histogram_value.num = 2;
  h.histogram_value.sum = 0;            // {-1.0, 1.0}
  h.histogram_value.sum_squares = 2;
  h.histogram_value.bucket_limit = {0.0, DBL_MAX};
  h.histogram_value.bucket = {1, 1};
  b.points.push_back(h);
  m["/metric_2"] = b;
  return m;
}

int main() {
  // ---- CreateTimeSeries golden -----------------------------------------------------------------
  auto sink = std::make_shared<MemorySink>();
  SinkClient client({"test_project"}, sink);
  CHECK(client.CreateTimeSeries(GoldenPoints()).ok());
  CHECK(sink->time_series.size() == 1);
  const std::string& r = sink->time_series[0];
  CHECK_HAS(r, "\"name\": \"projects/test_project\"");
  CHECK_HAS(r, "\"type\": \"custom.googleapis.com/metric_1\"");
  CHECK_HAS(r, "\"resource\": {\"type\": \"global\"}");
  CHECK_HAS(r, "\"end_time\": {\"seconds\": 12, \"nanos\": 345000000}");
  CHECK_HAS(r, "\"int64_value\": 33");
  CHECK_HAS(r, "\"count\": 2, \"mean\": 0.0, \"sum_of_squared_deviation\": 2.0");
  CHECK_HAS(r, "\"bounds\": [0.0]");
  CHECK_HAS(r, "\"bucket_counts\": [1, 1]");
  CHECK(client.CreateTimeSeries(PointSetMap()).code == Code::kCancelled);

  // ---- CreateMetricDescriptor golden -------------------------------------------------------------
  MetricDescriptor d;
  d.name = "/metric_1";
  d.description = "A test metric";
  d.label_names = {"label_1", "label_2"};
  d.metric_kind = MetricKind::kCumulative;
  d.value_type = ValueType::kHistogram;
  CHECK(client.CreateMetricDescriptor(d).ok());
  const std::string& dr = sink->descriptors[0];
  CHECK_HAS(dr, "\"type\": \"custom.googleapis.com/metric_1\"");
  CHECK_HAS(dr, "{\"key\": \"label_1\", \"value_type\": \"STRING\"}");
  CHECK_HAS(dr, "\"metric_kind\": \"CUMULATIVE\", \"value_type\": \"DISTRIBUTION\"");

  // ---- config -------------------------------------------------------------------------------------
  Config dflt("");
  CHECK(dflt.IsAllowlisted("/cloud_b200/train/step_time_usecs"));
  CHECK(!dflt.IsAllowlisted("/something/else"));
  Config custom(" /a , /b,/c ");
  CHECK(custom.IsAllowlisted("/a") && custom.IsAllowlisted("/b") && custom.IsAllowlisted("/c"));
  CHECK(!custom.IsAllowlisted("/cloud_b200/train/step_time_usecs"));
  CHECK(custom.DebugString() == "allowlist: [/a, /b, /c]");

  // ---- ShouldExport ---------------------------------------------------------------------------------
  Config cfg("/metric_1,/metric_2,/zero");
  PointSetMap g = GoldenPoints();
  CHECK(ShouldExport(cfg, g["/metric_1"]) && ShouldExport(cfg, g["/metric_2"]));
  PointSet zero = g["/metric_1"];
  zero.metric_name = "/zero";
  zero.points[0].int64_value = 0;
  CHECK(!ShouldExport(cfg, zero));                       // empty int64
  PointSet other = g["/metric_1"];
  other.metric_name = "/not_listed";
  CHECK(!ShouldExport(cfg, other));                      // not allow-listed
  PointSet empty_hist = g["/metric_2"];
  empty_hist.points[0].histogram_value.num = 0;
  CHECK(!ShouldExport(cfg, empty_hist));

  // ---- registry + exporter --------------------------------------------------------------------------
  setenv("TFC_MONITORING_ALLOWLIST", "/cloud_b200/train/step_time_usecs,/cloud_b200/comm/bytes_reduced", 1);
  setenv("TFC_MONITORING_SINK", "memory", 1);
  Registry* reg = Registry::Default();
  reg->DefineHistogram("/cloud_b200/train/step_time_usecs", "step time", {100, 1000, 10000});
  reg->Observe("/cloud_b200/train/step_time_usecs", 50);
  reg->Observe("/cloud_b200/train/step_time_usecs", 5000);
  reg->IncrementCounter("/cloud_b200/comm/bytes_reduced", 4096);
  reg->IncrementCounter("/cloud_b200/ignored", 7);
  reg->DefineCounter("/cloud_b200/comm/bucket_launches", "never incremented");
  Exporter* ex = Exporter::Get();
  CHECK(!ex->PeriodicallyExportMetrics());               // disabled by default
  ex->ExportMetrics();
  ex->ExportMetrics();
  auto* mem = dynamic_cast<MemorySink*>(SinkClient::Get()->sink().get());
  CHECK(mem != nullptr);
  if (mem) {
    CHECK(mem->time_series.size() == 2);
    CHECK(mem->descriptors.size() == 2);                 // once per metric, not once per export
    CHECK_HAS(mem->time_series[0], "/cloud_b200/comm/bytes_reduced");
    CHECK_HAS(mem->time_series[0], "\"int64_value\": 4096");
    CHECK_HAS(mem->time_series[0], "\"bounds\": [100.0, 1000.0, 10000.0]");
    CHECK_HAS(mem->time_series[0], "\"bucket_counts\": [1, 0, 1, 0]");
    CHECK(mem->time_series[0].find("/cloud_b200/ignored") == std::string::npos);
  }
  setenv("TFC_MONITORING_ENABLED", "true", 1);
  setenv("TFC_MONITORING_INTERVAL_MS", "20", 1);
  CHECK(ex->PeriodicallyExportMetrics());
  CHECK(ex->PeriodicallyExportMetrics());                // idempotent
  const int before = ex->exports();
  for (int i = 0; i < 200 && ex->exports() < before + 2; ++i)
    std::this_thread::sleep_for(std::chrono::milliseconds(10));
  CHECK(ex->exports() >= before + 2);
  ex->Stop();

  if (failures == 0) printf("ALL MONITORING TESTS PASSED\n");
  return failures ? 1 : 0;
}
