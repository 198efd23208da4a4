"""CPU check of the host-side filter canonicalisation (csrc/filter_range.hpp): the conjunction of a FilterSelection as ONE
value range or its complement.  csrc/tests/filter_range_test.cpp enumerates every conjunction of up to three children
(6 comparisons x 12 constants, IS NOT NULL) under AND / OR for a signed and an unsigned 8-bit type -- 1.6 million
conjunctions -- and compares the range with TemplatedFilterSelection's row-by-row answer on all 256 values; every accepted
range of up to two children is then clipped to every linear 8-bit segment by clip_range_to_segment (the function
filter_words_kernel runs per segment, same header) and the field-domain test is compared with the value test."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "duckdb_adaptive_compression_b200", "csrc")


def test_canonical_filter_range_exhaustive(tmp_path):
    exe = str(tmp_path / "filter_range_test")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-Werror", "-o", exe,
                           os.path.join(CSRC, "tests", "filter_range_test.cpp")])
    out = subprocess.check_output([exe], text=True, timeout=120)
    assert "FILTER_RANGE_OK" in out, out
    fields = dict(kv.split("=") for kv in out.split()[1:])
    assert int(fields["checked"]) > 1_500_000 and int(fields["accepted"]) > int(fields["refused"])
    assert int(fields["clipped"]) > 200_000_000


def test_swar8_byte_packing_exhaustive(tmp_path):
    """csrc/swar8.hpp (the word-parallel pack / unpack of one-byte slots inside pack_staged_kernel<1> and
    decode_kernel<1>) against the scalar definition of the codec: every width, every frame of reference, every slot value
    in every byte lane."""
    exe = str(tmp_path / "swar8_test")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-Werror", "-o", exe,
                           os.path.join(CSRC, "tests", "swar8_test.cpp")])
    out = subprocess.check_output([exe], text=True, timeout=120)
    assert "SWAR8_OK" in out, out
