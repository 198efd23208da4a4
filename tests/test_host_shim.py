"""The C++ host shim above the C ABI (duckdb_adaptive_compression_b200/host): the reference's
CompressionFunction / ColumnSegment / ColumnSegmentCatalog surface re-hosted on the GPU codec."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "duckdb_adaptive_compression_b200", "host")


def _build():
    subprocess.check_call(["make", "-C", HOST], stdout=subprocess.DEVNULL)


def test_policy_decision_rule_cpu():
    """90/10 split with the reference's float arithmetic, counter reset, accounting, thread start/stop."""
    _build()
    out = subprocess.check_output([os.path.join(HOST, "build", "policy_test")], text=True, timeout=60)
    assert "POLICY_TESTS_OK" in out


def test_filter_pushdown_stages_cpu():
    """host logic of ColumnData::FilterAggregate (how a pushed TableFilter becomes the chain of adacom_scan_select calls),
    no GPU touched"""
    _build()
    out = subprocess.check_output([os.path.join(HOST, "build", "filter_stage_test")], text=True, timeout=60)
    assert "FILTER_STAGE_TESTS_OK" in out


def test_hot_vector_cache_cpu():
    """ColumnData's hot-vector cache on host-resident (adaptive, never uploaded) segments: second-touch fill, hits, scans
    that continue behind a cached vector and across segment borders, the growing tail; no GPU touched"""
    _build()
    out = subprocess.check_output([os.path.join(HOST, "build", "cache_test")], text=True, timeout=60)
    assert "CACHE_TESTS_OK" in out


def test_column_data_over_host_resident_segments_cpu():
    """ColumnData driven like DuckDB drives the reference over segments that live on the host (adaptive mode, nothing
    uploaded): appends with NULLs and selection vectors, scans across segment borders, FetchRow / FetchRows, catalog
    accounting; no GPU touched"""
    _build()
    out = subprocess.check_output([os.path.join(HOST, "build", "host_resident_test")], text=True, timeout=120)
    assert "HOST_RESIDENT_TESTS_OK" in out


def test_host_shim_under_thread_sanitizer_cpu(tmp_path):
    """readers (point reads through the hot-vector cache, FetchRow, full scans with zero-copy vectors, catalog read
    counters) against a writer that appends to the same ColumnData and opens new segments, built with -fsanitize=thread
    over the shim's own sources; any report fails.  (Found and fixed with it: the segment vector could move under a
    reader while an append opened a segment; count and version are atomics like SegmentBase::count in the reference.)"""
    exe = str(tmp_path / "race_tsan")
    probe = subprocess.run(["g++", "-fsanitize=thread", "-x", "c++", "-", "-o", str(tmp_path / "probe")],
                           input="int main(){return 0;}", text=True, capture_output=True)
    if probe.returncode != 0:
        pytest.skip("no ThreadSanitizer runtime in this toolchain")
    csrc = os.path.join(os.path.dirname(HOST), "csrc")
    subprocess.check_call(["g++", "-O1", "-g", "-std=c++14", "-pthread", "-fsanitize=thread", "-o", exe,
                           os.path.join(HOST, "tests", "race_test.cpp"), os.path.join(HOST, "succinct_gpu.cpp"),
                           os.path.join(HOST, "column_data.cpp"), os.path.join(HOST, "column_segment_catalog.cpp"),
                           "-L" + csrc, "-ladacom_gpu", "-Wl,-rpath," + csrc])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, TSAN_OPTIONS="halt_on_error=0 exitcode=66"))
    assert r.returncode == 0 and "RACE_TEST_OK" in r.stdout, (r.stdout[-500:], r.stderr[-2000:])
    assert "ThreadSanitizer" not in r.stderr, r.stderr[-2000:]


@pytest.mark.parametrize("name", ["cache_test", "host_resident_test"])
def test_host_shim_under_address_sanitizer_cpu(tmp_path, name):
    """the CPU drivers of the shim again, built with -fsanitize=address,undefined over the shim's own sources (leak
    detection on): ColumnData over host-resident segments and the hot-vector cache"""
    probe = subprocess.run(["g++", "-fsanitize=address,undefined", "-x", "c++", "-", "-o", str(tmp_path / "probe")],
                           input="int main(){return 0;}", text=True, capture_output=True)
    if probe.returncode != 0:
        pytest.skip("no AddressSanitizer runtime in this toolchain")
    exe = str(tmp_path / (name + "_asan"))
    csrc = os.path.join(os.path.dirname(HOST), "csrc")
    subprocess.check_call(["g++", "-O1", "-g", "-std=c++14", "-pthread", "-fsanitize=address,undefined",
                           "-fno-sanitize-recover=undefined", "-o", exe, os.path.join(HOST, "tests", name + ".cpp"),
                           os.path.join(HOST, "succinct_gpu.cpp"), os.path.join(HOST, "column_data.cpp"),
                           os.path.join(HOST, "column_segment_catalog.cpp"), "-L" + csrc, "-ladacom_gpu",
                           "-Wl,-rpath," + csrc])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, ASAN_OPTIONS="detect_leaks=1:protect_shadow_gap=0"))
    assert r.returncode == 0 and "_OK" in r.stdout, (r.stdout[-500:], r.stderr[-2000:])


def test_shim_exports_reference_surface():
    _build()
    out = subprocess.check_output(["nm", "-DC", "--defined-only", os.path.join(HOST, "libadacom_host.so")], text=True)
    for sym in ["adacom_host::SuccinctGpuFun::GetFunction", "adacom_host::SuccinctGpuFun::TypeIsSupported",
                "adacom_host::ColumnSegment::CreateTransientSegment", "adacom_host::ColumnSegment::Compact",
                "adacom_host::ColumnSegment::Uncompact", "adacom_host::ColumnSegment::GetDataSize",
                "adacom_host::ColumnSegment::Scan", "adacom_host::ColumnSegment::FetchRow",
                "adacom_host::ColumnSegment::Append", "adacom_host::ColumnSegmentCatalog::AddReadAccess",
                "adacom_host::ColumnSegmentCatalog::CompactAllSegments",
                "adacom_host::ColumnSegmentCatalog::GetTotalDataSize",
                "adacom_host::ColumnSegmentCatalog::EnableBackgroundThreadCompaction",
                "adacom_host::ColumnSegmentCatalog::GetDataSizePerDevice",
                "adacom_host::ColumnData::AppendData", "adacom_host::ColumnData::ScanVector",
                "adacom_host::ColumnData::InitializeScanWithOffset", "adacom_host::ColumnData::FetchRow",
                "adacom_host::ColumnData::FetchRows", "adacom_host::ColumnData::CompactPending",
                "adacom_host::ColumnData::CompactMany", "adacom_host::ColumnData::UncompactMany",
                "adacom_host::ColumnData::FilterAggregate", "adacom_host::ColumnData::BuildFilterStages"]:
        assert sym in out, sym


@pytest.mark.gpu
def test_shim_driven_like_duckdb_gpu():
    exe = os.path.join(HOST, "build", "host_test")
    if not os.path.exists(exe):
        _build()
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "HOST_TESTS_OK" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_row_groups_dealt_over_devices_gpu():
    """One process, all visible GPUs: row group g -> device g mod G, per-device scan threads, host-side combination,
    key -> owner routing, per-device resident bytes (host/tests/multi_gpu_test.cpp)."""
    exe = os.path.join(HOST, "build", "multi_gpu_test")
    if not os.path.exists(exe):
        _build()
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "MULTI_GPU_TEST_OK" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_vtable_bench_small_gpu():
    """The drop-in bench driver on a small column: answers checked inside the binary (scan_ok / fetch_ok)."""
    import json
    exe = os.path.join(HOST, "build", "vtable_bench")
    if not os.path.exists(exe):
        _build()
    r = subprocess.run([exe, "22", "13", "2", "16"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["scan_ok"] and out["fetch_ok"] and out["device_columns"] >= 2
    assert out["zero_copy_vectors"] + 2 * out["segments"] >= out["vectors"]
