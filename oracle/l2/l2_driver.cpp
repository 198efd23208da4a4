// oracle/l2/l2_driver.cpp -- TEST INFRASTRUCTURE ONLY (SURVEY 8(c) "L2": the whole reference, driven through its own
// public API).
//
// The reference's benchmark_runner cannot be configured in this tree (third_party/sqlite/sqlite3.c is a stripped blob,
// needed by BUILD_UNITTESTS which benchmark/ requires, CMakeLists.txt:817-823), so the ~40-line benchmark bodies are
// re-hosted here, statement for statement, on duckdb.hpp:
//   zipf        benchmark/micro/succinct/zipf_distribution.cpp:13-61   (Succinct / SuccinctPadded / NonSuccinct)
//   sequential  benchmark/micro/succinct/sequential.cpp:11-46
//   skews       benchmark/micro/succinct/zipf_distribution_diff_skews.cpp:13-111 (adaptive / non_adaptive / non_succinct,
//               the three arms of run_diff_zipfian.sh:12-14; prints (queries, memory) once per second like :79-96)
//   tpch        CALL dbgen(sf) (extension/tpch), CompactAllSegments(), Q1 + Q6 (extension/tpch/dbgen/queries), exports
//               lineitem's integer columns as flat little-endian files -> golden inputs + reference answers
//   nulls       a nullable INTEGER column through append / compaction / filter / SUM
// Every mode prints ONE JSON line with wall times, ColumnSegmentCatalog::GetTotalDataSize() and an answer checksum.
//
// The same source is linked twice by oracle/l2/build_drivers.sh:
//   oracle/_ref/l2/l2_driver_ref  against the UNMODIFIED reference library           (mints expected answers / timings)
//   oracle/_ref/l2/l2_driver_gpu  against the reference + integration/duckdb/reference.patch + succinct_gpu.cpp +
//                                 libadacom_gpu.so                                   (the drop-in run, needs a B200)
// The only difference from the reference bodies: std::mt19937 is seeded with a fixed number (the reference seeds from
// std::random_device, zipf_distribution.cpp:29-30), so both arms draw identical look-up keys.
#include <algorithm>
#include <atomic>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <vector>
#define private public // debugging aid (L2_DUMP_SEGMENTS): read-only walk over the catalog's segment map
#include "duckdb/catalog/catalog_entry/column_segment_catalog.hpp"
#undef private
#include "duckdb/storage/table/column_segment.hpp"
#include "duckdb.hpp"
#include "duckdb/catalog/catalog.hpp"
#include "duckdb/main/appender.hpp"
#include "duckdb/main/database_manager.hpp"
#include "duckdb/storage/buffer_manager.hpp"
#include "tpch-extension.hpp"
#include "zipf.cpp" // benchmark/micro/succinct/zipf.cpp (the reference's rejection-inversion sampler, header-style)

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <random>
#include <sstream>
#include <string>
#include <vector>

using namespace duckdb;

namespace {

double Now() {
	return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

ColumnSegmentCatalog *SegmentCatalog(DuckDB &db) {
	return db.instance->GetDatabaseManager().GetSystemCatalog().GetColumnSegmentCatalog();
}

//! L2_DUMP_SEGMENTS=<file>: one line per data segment (type, start, count, size, compacted, function), sorted by the caller
void DumpSegments(DuckDB &db) {
	const char *path = getenv("L2_DUMP_SEGMENTS");
	if (!path) {
		return;
	}
	std::ofstream out(path);
	for (auto &entry : SegmentCatalog(db)->statistics) {
		auto seg = entry.first;
		out << seg->type.ToString() << " start=" << seg->start << " count=" << seg->count << " size=" << seg->GetDataSize()
		    << " segment_size=" << seg->SegmentSize() << " compacted=" << seg->IsBitCompressed()
		    << " fn=" << (int)seg->function->type << " min=" << seg->GetMinFactor() << " max=" << seg->GetMax() << "\n";
	}
}

void MustSucceed(unique_ptr<MaterializedQueryResult> &result, const std::string &what) {
	if (result->HasError()) {
		fprintf(stderr, "l2_driver: %s failed: %s\n", what.c_str(), result->GetError().c_str());
		exit(3);
	}
}

void ApplyVariant(DuckDB &db, const std::string &variant) {
	auto &config = db.instance->config;
	if (variant == "succinct") {
		// defaults (config.hpp:189-197)
	} else if (variant == "padded") {
		config.succinct_padded_to_next_byte_enabled = true; // zipf_distribution.cpp:65
	} else if (variant == "nonsuccinct" || variant == "non_succinct") {
		config.succinct_enabled = false; // zipf_distribution.cpp:109
	} else if (variant == "adaptive") {
		config.adaptive_succinct_compression_enabled = true; // zipf_distribution_diff_skews.cpp:20
	} else if (variant == "non_adaptive") {
		config.succinct_enabled = true; // zipf_distribution_diff_skews.cpp:115
	} else {
		fprintf(stderr, "l2_driver: unknown variant %s\n", variant.c_str());
		exit(2);
	}
}

template <class T>
void LoadSequential(Connection &conn, const char *sql_type, size_t n) {
	auto r = conn.Query(std::string("CREATE TABLE t1(i ") + sql_type + ");");
	MustSucceed(r, "create");
	Appender appender(conn, "t1");
	for (size_t i = 0; i < n; i++) {
		appender.BeginRow();
		appender.Append<T>((T)i);
		appender.EndRow();
	}
	appender.Close();
}

//! FNV-1a over the 64-bit values of a result (order-sensitive)
struct Checksum {
	uint64_t h = 1469598103934665603ull;
	uint64_t rows = 0;
	void Add(uint64_t v) {
		for (int b = 0; b < 8; b++) {
			h ^= (v >> (8 * b)) & 0xff;
			h *= 1099511628211ull;
		}
		rows++;
	}
	void AddResult(MaterializedQueryResult &result) {
		for (idx_t r = 0; r < result.RowCount(); r++) {
			for (idx_t c = 0; c < result.ColumnCount(); c++) {
				auto v = result.GetValue(c, r);
				Add(v.IsNull() ? 0x9e3779b97f4a7c15ull : (uint64_t)v.GetValue<int64_t>());
			}
		}
	}
};

// ---- zipf_distribution.cpp:13-61 ----
int RunZipf(size_t n_inserts, size_t n_lookups, uint32_t seed, const std::string &variant, double skew) {
	DuckDB db(nullptr);
	Connection conn(db);
	ApplyVariant(db, variant);
	double t0 = Now();
	LoadSequential<uint32_t>(conn, "UINTEGER", n_inserts);
	std::mt19937 gen {seed};
	Zipf<uint32_t, double> zipf(n_inserts, skew);
	std::vector<uint32_t> keys;
	for (size_t i = 0; i < n_lookups; ++i) {
		keys.push_back(uint32_t(std::round(zipf(gen))));
	}
	uint64_t size_before = SegmentCatalog(db)->GetTotalDataSize();
	if (variant == "succinct") {
		SegmentCatalog(db)->CompactAllSegments(); // zipf_distribution.cpp:35-36
	}
	double t1 = Now();
	uint64_t size_loaded = SegmentCatalog(db)->GetTotalDataSize();
	Checksum sum;
	conn.Query("BEGIN TRANSACTION");
	for (size_t i = 0; i < n_lookups; ++i) {
		auto result = conn.Query("SELECT i FROM t1 where i == " + std::to_string(keys[i]));
		MustSucceed(result, "lookup");
		sum.AddResult(*result);
	}
	conn.Query("COMMIT");
	double t2 = Now();
	uint64_t size_after = SegmentCatalog(db)->GetTotalDataSize();
	printf("{\"mode\": \"zipf\", \"variant\": \"%s\", \"n_inserts\": %zu, \"n_lookups\": %zu, \"seed\": %u, \"skew\": %g, "
	       "\"load_s\": %.6f, \"run_s\": %.6f, \"lookups_per_s\": %.1f, \"size_before_compact\": %llu, \"size_loaded\": %llu, "
	       "\"size_after\": %llu, \"result_rows\": %llu, \"checksum\": \"%016llx\"}\n",
	       variant.c_str(), n_inserts, n_lookups, seed, skew, t1 - t0, t2 - t1, n_lookups / (t2 - t1),
	       (unsigned long long)size_before, (unsigned long long)size_loaded, (unsigned long long)size_after,
	       (unsigned long long)sum.rows, (unsigned long long)sum.h);
	return 0;
}

// ---- sequential.cpp:11-46 ----
int RunSequential(size_t n_inserts, size_t n_scans, const std::string &variant) {
	DuckDB db(nullptr);
	Connection conn(db);
	ApplyVariant(db, variant);
	double t0 = Now();
	LoadSequential<int32_t>(conn, "INTEGER", n_inserts);
	double t1 = Now();
	uint64_t size_loaded = SegmentCatalog(db)->GetTotalDataSize();
	Checksum sum;
	conn.Query("BEGIN TRANSACTION");
	for (size_t i = 0; i < n_scans; ++i) {
		auto result = conn.Query("SELECT * FROM t1");
		MustSucceed(result, "scan");
		// fold the chunks without GetValue (10^6+ rows): the raw int32 payload
		while (true) {
			auto chunk = result->Fetch();
			if (!chunk || chunk->size() == 0) {
				break;
			}
			auto data = FlatVector::GetData<int32_t>(chunk->data[0]);
			for (idx_t r = 0; r < chunk->size(); r++) {
				sum.Add((uint64_t)(int64_t)data[r]);
			}
		}
	}
	conn.Query("COMMIT");
	double t2 = Now();
	uint64_t size_after = SegmentCatalog(db)->GetTotalDataSize();
	printf("{\"mode\": \"sequential\", \"variant\": \"%s\", \"n_inserts\": %zu, \"n_scans\": %zu, \"load_s\": %.6f, "
	       "\"run_s\": %.6f, \"rows_per_s\": %.1f, \"size_loaded\": %llu, \"size_after\": %llu, \"result_rows\": %llu, "
	       "\"checksum\": \"%016llx\"}\n",
	       variant.c_str(), n_inserts, n_scans, t1 - t0, t2 - t1, (double)n_inserts * n_scans / (t2 - t1),
	       (unsigned long long)size_loaded, (unsigned long long)size_after, (unsigned long long)sum.rows,
	       (unsigned long long)sum.h);
	return 0;
}

// ---- zipf_distribution_diff_skews.cpp:13-111: look-ups for `duration` seconds, (queries, memory) per second ----
int RunSkews(size_t n_inserts, size_t n_lookups, uint32_t seed, const std::string &variant, double skew,
             double duration, bool background) {
	DuckDB db(nullptr);
	Connection conn(db);
	ApplyVariant(db, variant);
	double t0 = Now();
	LoadSequential<uint32_t>(conn, "UINTEGER", n_inserts);
	std::mt19937 gen {seed};
	Zipf<uint32_t, double> zipf(n_inserts / 2, skew);
	std::vector<uint32_t> keys;
	for (size_t i = 0; i < n_lookups; ++i) {
		keys.push_back(uint32_t(std::round(zipf(gen))));
	}
	double t1 = Now();
	if (background && variant == "adaptive") {
		// the reference starts the policy thread from Catalog::Initialize when adaptive mode is configured before the
		// database is opened; the benchmark sets the flag afterwards, so the thread is started explicitly here
		SegmentCatalog(db)->EnableBackgroundThreadCompaction();
	}
	auto &buffer_manager = db.instance->GetBufferManager();
	Checksum sum;
	std::ostringstream series;
	size_t i = 0, transaction_count = 0, total = 0;
	double start = Now(), qps_start = start;
	bool first = true;
	while (Now() - start < duration) {
		conn.Query("BEGIN TRANSACTION");
		auto result = conn.Query("SELECT t1.i FROM t1 where i == " + std::to_string(keys[i]));
		MustSucceed(result, "lookup");
		sum.AddResult(*result);
		conn.Query("COMMIT");
		i++;
		transaction_count++;
		total++;
		if (Now() - qps_start >= 1.0) {
			series << (first ? "" : ", ") << "[" << transaction_count << ", " << buffer_manager.GetUsedMemory() << ", "
			       << SegmentCatalog(db)->GetTotalDataSize() << "]";
			first = false;
			transaction_count = 0;
			qps_start = Now();
		}
		if (i >= keys.size()) {
			i = 0;
		}
	}
	double t2 = Now();
	printf("{\"mode\": \"skews\", \"variant\": \"%s\", \"n_inserts\": %zu, \"n_lookups\": %zu, \"seed\": %u, \"skew\": %g, "
	       "\"load_s\": %.6f, \"run_s\": %.6f, \"queries\": %zu, \"queries_per_s\": %.1f, \"result_rows\": %llu, "
	       "\"series_qps_usedmem_datasize\": [%s]}\n",
	       variant.c_str(), n_inserts, n_lookups, seed, skew, t1 - t0, t2 - t1, total, total / (t2 - t1),
	       (unsigned long long)sum.rows, series.str().c_str());
	return 0;
}

// ---- TPC-H ----
std::string DecimalDigits(const std::string &text) {
	// "11803420.2534" -> "118034202534": the integer at the DECIMAL's own scale
	std::string out;
	for (char c : text) {
		if (c != '.') {
			out.push_back(c);
		}
	}
	return out;
}

template <class T>
void ExportColumn(Connection &conn, const std::string &expr, const std::string &path) {
	auto result = conn.Query("SELECT " + expr + " FROM lineitem");
	MustSucceed(result, "export " + expr);
	std::ofstream out(path, std::ios::binary);
	while (true) {
		auto chunk = result->Fetch();
		if (!chunk || chunk->size() == 0) {
			break;
		}
		chunk->data[0].Flatten(chunk->size());
		auto data = FlatVector::GetData<T>(chunk->data[0]);
		out.write((const char *)data, chunk->size() * sizeof(T));
	}
}

int RunTpch(double sf, const std::string &outdir, int threads, int reps) {
	DuckDB db(nullptr);
	db.LoadExtension<TPCHExtension>();
	Connection conn(db);
	auto r = conn.Query("PRAGMA threads=" + std::to_string(threads));
	MustSucceed(r, "threads");
	double t0 = Now();
	r = conn.Query("CALL dbgen(sf=" + std::to_string(sf) + ")");
	MustSucceed(r, "dbgen");
	double t1 = Now();
	uint64_t size_before = SegmentCatalog(db)->GetTotalDataSize();
	SegmentCatalog(db)->CompactAllSegments();
	uint64_t size_after = SegmentCatalog(db)->GetTotalDataSize();
	double t2 = Now();
	DumpSegments(db);
	std::string q6_text = TPCHExtension::GetQuery(6), q1_text = TPCHExtension::GetQuery(1);
	double q6_best = 1e30, q1_best = 1e30;
	unique_ptr<MaterializedQueryResult> q6, q1;
	for (int k = 0; k < reps; k++) {
		double a = Now();
		q6 = conn.Query(q6_text);
		double b = Now();
		q1 = conn.Query(q1_text);
		double c = Now();
		MustSucceed(q6, "Q6");
		MustSucceed(q1, "Q1");
		q6_best = std::min(q6_best, b - a);
		q1_best = std::min(q1_best, c - b);
	}
	auto rows = conn.Query("SELECT count(*) FROM lineitem");
	MustSucceed(rows, "count");
	std::ostringstream q1_json;
	for (idx_t g = 0; g < q1->RowCount(); g++) {
		q1_json << (g ? ", " : "") << "{\"returnflag\": \"" << q1->GetValue(0, g).ToString() << "\", \"linestatus\": \""
		        << q1->GetValue(1, g).ToString() << "\", \"sum_qty\": \"" << DecimalDigits(q1->GetValue(2, g).ToString())
		        << "\", \"sum_base_price\": \"" << DecimalDigits(q1->GetValue(3, g).ToString())
		        << "\", \"sum_disc_price\": \"" << DecimalDigits(q1->GetValue(4, g).ToString()) << "\", \"sum_charge\": \""
		        << DecimalDigits(q1->GetValue(5, g).ToString()) << "\", \"avg_qty\": " << q1->GetValue(6, g).ToString()
		        << ", \"avg_price\": " << q1->GetValue(7, g).ToString() << ", \"avg_disc\": " << q1->GetValue(8, g).ToString()
		        << ", \"count_order\": " << q1->GetValue(9, g).ToString() << "}";
	}
	if (outdir != "-") {
		ExportColumn<int32_t>(conn, "l_orderkey", outdir + "/l_orderkey.i32");
		ExportColumn<int32_t>(conn, "l_partkey", outdir + "/l_partkey.i32");
		ExportColumn<int32_t>(conn, "l_suppkey", outdir + "/l_suppkey.i32");
		ExportColumn<int32_t>(conn, "l_linenumber", outdir + "/l_linenumber.i32");
		ExportColumn<int32_t>(conn, "l_quantity", outdir + "/l_quantity.i32");
		// DECIMAL(15,2) is an INT64 scaled by 100 physically (extension/tpch/dbgen/dbgen.cpp:402-412)
		ExportColumn<int64_t>(conn, "CAST(l_extendedprice * 100 AS BIGINT)", outdir + "/l_extendedprice.i64");
		ExportColumn<int64_t>(conn, "CAST(l_discount * 100 AS BIGINT)", outdir + "/l_discount.i64");
		ExportColumn<int64_t>(conn, "CAST(l_tax * 100 AS BIGINT)", outdir + "/l_tax.i64");
		// DATE is an INT32 of days since 1970-01-01 physically
		ExportColumn<int32_t>(conn, "CAST(l_shipdate - DATE '1970-01-01' AS INTEGER)", outdir + "/l_shipdate.i32");
		ExportColumn<int32_t>(conn, "CAST(l_commitdate - DATE '1970-01-01' AS INTEGER)", outdir + "/l_commitdate.i32");
		ExportColumn<int32_t>(conn, "CAST(l_receiptdate - DATE '1970-01-01' AS INTEGER)", outdir + "/l_receiptdate.i32");
		// VARCHAR in the reference, not succinct-eligible: supplied to the GPU scan as one UTINYINT group code
		// 0:A/F 1:N/F 2:N/O 3:R/F (documented deviation)
		ExportColumn<uint8_t>(conn,
		                      "CAST(CASE WHEN l_returnflag='A' THEN 0 WHEN l_returnflag='R' THEN 3 WHEN l_linestatus='F' "
		                      "THEN 1 ELSE 2 END AS UTINYINT)",
		                      outdir + "/l_rfls.u8");
	}
	printf("{\"mode\": \"tpch\", \"sf\": %g, \"threads\": %d, \"rows\": %s, \"dbgen_s\": %.4f, \"compact_s\": %.4f, "
	       "\"size_before_compact\": %llu, \"size_after_compact\": %llu, \"q6_s\": %.6f, \"q1_s\": %.6f, "
	       "\"q6_revenue\": \"%s\", \"q6_revenue_scaled\": \"%s\", \"q1\": [%s]}\n",
	       sf, threads, rows->GetValue(0, 0).ToString().c_str(), t1 - t0, t2 - t1, (unsigned long long)size_before,
	       (unsigned long long)size_after, q6_best, q1_best, q6->GetValue(0, 0).ToString().c_str(),
	       DecimalDigits(q6->GetValue(0, 0).ToString()).c_str(), q1_json.str().c_str());
	return 0;
}

// ---- nullable column: append with NULL gaps, compaction, filter pushdown, SUM ----
int RunNulls(size_t n, const std::string &variant) {
	DuckDB db(nullptr);
	Connection conn(db);
	ApplyVariant(db, variant == "zeros" ? "succinct" : variant == "zeros_plain" ? "nonsuccinct" : variant);
	auto r = conn.Query("CREATE TABLE t1(i INTEGER);");
	MustSucceed(r, "create");
	{
		Appender appender(conn, "t1");
		for (size_t i = 0; i < n; i++) {
			appender.BeginRow();
			if (variant == "zeros" || variant == "zeros_plain") {
				appender.Append<int32_t>(0);
			} else if (i % 7 == 3) {
				appender.Append<std::nullptr_t>(nullptr);
			} else {
				appender.Append<int32_t>((int32_t)(1000 + (i * 2654435761u) % 5000));
			}
			appender.EndRow();
		}
		appender.Close();
	}
	SegmentCatalog(db)->CompactAllSegments();
	uint64_t size = SegmentCatalog(db)->GetTotalDataSize();
	DumpSegments(db);
	Checksum sum;
	for (const char *sql : {"SELECT sum(i), count(i), count(*), min(i), max(i) FROM t1",
	                        "SELECT sum(i), count(*) FROM t1 WHERE i < 3000",
	                        "SELECT count(*) FROM t1 WHERE i IS NULL",
	                        "SELECT sum(i), count(*) FROM t1 WHERE i >= 2000 AND i <= 2500",
	                        "SELECT sum(i), count(*) FROM t1 WHERE i = 1234 OR i > 5900"}) {
		auto result = conn.Query(sql);
		MustSucceed(result, sql);
		sum.AddResult(*result);
	}
	printf("{\"mode\": \"nulls\", \"variant\": \"%s\", \"n\": %zu, \"size\": %llu, \"checksum\": \"%016llx\"}\n",
	       variant.c_str(), n, (unsigned long long)size, (unsigned long long)sum.h);
	return 0;
}

} // namespace

int main(int argc, char **argv) {
	if (argc < 2) {
		fprintf(stderr,
		        "usage: %s zipf <n_inserts> <n_lookups> <seed> <succinct|padded|nonsuccinct> [skew=1]\n"
		        "       %s sequential <n_inserts> <n_scans> <succinct|padded|nonsuccinct>\n"
		        "       %s skews <n_inserts> <n_lookups> <seed> <adaptive|non_adaptive|non_succinct> <skew> <seconds> [background=1]\n"
		        "       %s tpch <sf> <outdir|-> [threads=1] [reps=3]\n"
		        "       %s nulls <n> <variant>\n"
		        "       %s zipfkeys <n> <skew> <count> <seed>\n",
		        argv[0], argv[0], argv[0], argv[0], argv[0], argv[0]);
		return 2;
	}
	std::string mode = argv[1];
	try {
		if (mode == "zipf" && argc >= 6) {
			return RunZipf(strtoull(argv[2], nullptr, 10), strtoull(argv[3], nullptr, 10), (uint32_t)atoi(argv[4]), argv[5],
			               argc > 6 ? atof(argv[6]) : 1.0);
		}
		if (mode == "sequential" && argc >= 5) {
			return RunSequential(strtoull(argv[2], nullptr, 10), strtoull(argv[3], nullptr, 10), argv[4]);
		}
		if (mode == "skews" && argc >= 8) {
			return RunSkews(strtoull(argv[2], nullptr, 10), strtoull(argv[3], nullptr, 10), (uint32_t)atoi(argv[4]), argv[5],
			                atof(argv[6]), atof(argv[7]), argc > 8 ? atoi(argv[8]) != 0 : true);
		}
		if (mode == "tpch" && argc >= 4) {
			return RunTpch(atof(argv[2]), argv[3], argc > 4 ? atoi(argv[4]) : 1, argc > 5 ? atoi(argv[5]) : 3);
		}
		if (mode == "zipfkeys" && argc >= 6) {
			// the reference's sampler (benchmark/micro/succinct/zipf.cpp) driven exactly like zipf_distribution.cpp:29-33
			std::mt19937 gen {(uint32_t)atoi(argv[5])};
			Zipf<uint32_t, double> zipf(strtoull(argv[2], nullptr, 10), atof(argv[3]));
			printf("[");
			for (size_t i = 0, n = strtoull(argv[4], nullptr, 10); i < n; i++) {
				printf("%s%u", i ? ", " : "", uint32_t(std::round(zipf(gen))));
			}
			printf("]\n");
			return 0;
		}
		if (mode == "nulls" && argc >= 4) {
			return RunNulls(strtoull(argv[2], nullptr, 10), argv[3]);
		}
	} catch (std::exception &ex) {
		fprintf(stderr, "l2_driver: exception: %s\n", ex.what());
		return 3;
	}
	fprintf(stderr, "l2_driver: bad arguments\n");
	return 2;
}
