"""llama_kv_cache_paged (llama.cpp_test_b200/host/src/llama_kv_cache_paged.h): the reference's llama_kv_cache method surface
(src/llama-kv-cache.h:17-26, :93-111) on the block manager — SURVEY.md §8(f) N2.  The checks live in C++
(tests/cpp/kv_cache_paged_check.cpp: a cell-per-token model of llama_kv_cache_unified driven by the same seeded call stream,
with the adapter's work lists applied to an array of token ids the way the device applies them to K/V rows); this test
compiles that program against the header + the built library and runs it."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "llama.cpp_test_b200", "host")


def test_llama_kv_cache_paged_against_the_cell_model(tmp_path):
    exe = str(tmp_path / "kv_cache_paged_check")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-I", os.path.join(HOST, "src"), "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "kv_cache_paged_check.cpp"),
           os.path.join(HOST, "llama_kv_cache_paged.cpp"), os.path.join(HOST, "paged_kv_cache.cpp"), "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    run = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0, (run.stdout[-1500:], run.stderr[-1500:])
    lines = run.stdout.strip().splitlines()
    assert lines[-1] == "OK" and len(lines) == 25
    # the random streams really exercised the interesting paths
    assert sum(int(l.split(",")[1].split()[0]) for l in lines[:-1]) > 50          # context shifts


def test_library_exports_the_adapter():
    """the class is part of libb200attn.so (a C++ host links it like the fork's classes)"""
    lib = os.path.join(ROOT, "llama.cpp_test_b200", "libb200attn.so")
    out = subprocess.run(["nm", "-DC", "--defined-only", lib], capture_output=True, text=True).stdout
    for m in ("llama_kv_cache_paged::seq_rm", "llama_kv_cache_paged::seq_add", "llama_kv_cache_paged::update", "llama_kv_cache_paged::find_slot"):
        assert m in out, m


def test_paged_cache_stress_under_sanitizers(tmp_path):
    """tests/cpp/paged_cache_stress.cpp: 12 seeded streams of 1500 calls over PagedKVCache (append, prepare_batch, seq_cp, seq_rm,
    seq_keep, context_shift, defrag) in small pools, built with AddressSanitizer + UBSan where the toolchain has them; after
    every call the reference counts are recomputed from the block tables and the free count must close the books"""
    exe = str(tmp_path / "paged_cache_stress")
    base = ["g++", "-std=c++17", "-g", "-O1", "-Wall", "-Wextra", "-I", os.path.join(HOST, "src"),
            os.path.join(ROOT, "tests", "cpp", "paged_cache_stress.cpp"), os.path.join(HOST, "paged_kv_cache.cpp"), "-o", exe]
    r = subprocess.run(base + ["-fsanitize=address,undefined", "-fno-omit-frame-pointer"], capture_output=True, text=True)
    if r.returncode != 0:                                   # no sanitizer runtime on this machine: the invariants still run
        r = subprocess.run(base, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    run = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert run.returncode == 0, (run.stdout[-1500:], run.stderr[-3000:])
    lines = run.stdout.strip().splitlines()
    assert lines[-1] == "OK" and len(lines) == 13
    assert sum(int(l.split(",")[2].split()[0]) for l in lines[:-1]) > 500         # the refusal paths ran
