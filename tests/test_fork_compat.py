"""The fork's block-KV prototype classes, restated in
llama.cpp_test_b200/host/fork_compat.cpp behind the fork's own header names.

(a) When the reference tree is present (this container), the reference's OWN test programs
    (tests/test_block_manager.cpp, test_block_kv_cache.cpp, test_kv_cache.cpp, test_block_kv_attention.cpp,
    test_paged_attention.cpp) are compiled UNCHANGED, where they lie, against our headers + fork_compat.cpp and
    their stdout/stderr must equal tests/golden/fork/* — the output of the reference's prebuilt binaries.
(b) Everywhere: the allocator traces recovered in SURVEY.md §2.3, asserted through the C ABI (b200_host.h).
"""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
HOST = os.path.join(ROOT, "llama.cpp_test_b200", "host")
GOLD = os.path.join(ROOT, "tests", "golden", "fork")

# test_block_manager: the prebuilt binary predates the "edge tests" half of tests/test_block_manager.cpp (:28-62).
# test_kv_cache: the prebuilt binary dies with glibc "double free or corruption (out)" after its 4th line (a heap
# overrun in the prototype's insertKVBlock once the manager grows a second PhysicalBlock); the restatement follows the
# recovered semantics (SURVEY.md §2.3 expected trace) and runs to completion.
PREFIX_ONLY = {"test_block_manager", "test_kv_cache"}

PROGRAMS = ["test_block_manager", "test_block_kv_cache", "test_kv_cache", "test_block_kv_attention", "test_paged_attention"]


@pytest.mark.parametrize("name", PROGRAMS)
def test_reference_test_program_output_matches_reference_binary(name, tmp_path):
    src = os.path.join(REF, "tests", name + ".cpp")
    if not os.path.exists(src):
        pytest.skip("reference tree not present (GPU box): the committed goldens are checked in this container")
    exe = str(tmp_path / name)
    # the reference tests include "block_manager.h", "../src/block_kv_cache.h", "../src/block_kv_attention.h"
    # relative to tests/: -iquote a directory whose ../src is ours
    shim = tmp_path / "tests"
    shim.mkdir()
    os.symlink(os.path.join(HOST, "src"), tmp_path / "src")
    cmd = ["g++", "-std=c++17", "-O1", "-w", "-I", os.path.join(HOST, "src"), "-iquote", str(shim), "-I", str(shim),
           src, os.path.join(HOST, "fork_compat.cpp"), "-o", exe]
    # "../src/x.h" is resolved relative to the including file's directory first; that directory is read-only
    # reference territory, so give the preprocessor the same relative path from a directory we own
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=str(shim))
    if r.returncode != 0:
        # second try: compile a one-line wrapper that sits in tmp/tests/ and includes the reference file
        wrap = shim / (name + "_wrap.cpp")
        wrap.write_text(f'#include "{src}"\n')
        cmd[cmd.index(src)] = str(wrap)
        r = subprocess.run(cmd, capture_output=True, text=True, cwd=str(shim))
    assert r.returncode == 0, r.stderr[-2000:]
    run = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert run.returncode == 0
    with open(os.path.join(GOLD, name + ".stdout")) as f:
        gold_out = f.read()
    with open(os.path.join(GOLD, name + ".stderr")) as f:
        gold_err = f.read().splitlines()
    if name in PREFIX_ONLY:
        # the prebuilt binary stops early (see PREFIX_ONLY): everything it printed must be what we print first
        assert run.stdout.startswith(gold_out)
        gold_err = [l for l in gold_err if not l.startswith("double free") and "Destructor called" not in l]
        ours = [l for l in run.stderr.splitlines() if "Destructor called" not in l]
        assert ours[:len(gold_err)] == gold_err
    else:
        assert run.stdout == gold_out
        assert run.stderr.splitlines() == gold_err


def test_apply_attention_matches_the_reference_binary(tmp_path):
    """F6 (blockKVAttentionRealFwd / applyAttention): the reference ships ./test_apply_attention without its source; its
    main() is restated in tests/cpp/apply_attention_main.cpp from the binary's output, linked with fork_compat.cpp, and
    stdout must equal what the reference binary printed (tests/golden/fork/test_apply_attention.stdout) — both the
    usePaged = true branch (chunked dummy attention) and the 4-key scalar softmax."""
    exe = str(tmp_path / "apply_attention")
    cmd = ["g++", "-std=c++17", "-O1", "-w", "-I", os.path.join(HOST, "src"), os.path.join(ROOT, "tests", "cpp", "apply_attention_main.cpp"),
           os.path.join(HOST, "fork_compat.cpp"), "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    run = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert run.returncode == 0
    with open(os.path.join(GOLD, "test_apply_attention.stdout")) as f:
        assert run.stdout == f.read()


def test_block_manager_trace_3_blocks(pkg):
    """tests/test_block_manager.cpp:8-26: alloc 0,1,2; 4th alloc evicts the LOWEST allocated index (0) and
    re-allocates it; free(1); alloc -> 1; all allocated at the end"""
    pkg.lib.b200_fork_trace(0)
    bm = pkg.BlockManager(3)
    assert [bm.allocateBlock() for _ in range(3)] == [0, 1, 2]
    assert bm.allocateBlock() == 0
    bm.freeBlock(1)
    assert bm.allocateBlock() == 1
    assert bm.getBlockStatus() == [True, True, True]


def test_block_manager_trace_2_blocks_edge_cases(pkg):
    """tests/test_block_manager.cpp:30-62: freeing a never-allocated block and a double free are silent no-ops;
    out-of-range indices are ignored; evict on an empty manager returns -1"""
    pkg.lib.b200_fork_trace(0)
    bm = pkg.BlockManager(2)
    assert bm.evictBlock() == -1
    bm.freeBlock(1)                      # never allocated
    bm.freeBlock(-1); bm.freeBlock(2)    # out of range
    assert [bm.allocateBlock() for _ in range(3)] == [0, 1, 0]
    bm.freeBlock(1); bm.freeBlock(1)
    assert bm.getBlockStatus() == [True, False]
    assert bm.evictBlock() == 0 and bm.evictBlock() == -1
    assert bm.getBlockStatus() == [False, False]


def test_block_manager_zero_blocks(pkg):
    pkg.lib.b200_fork_trace(0)
    bm = pkg.BlockManager(0)
    assert bm.allocateBlock() == -1
    assert bm.getBlockStatus() == []
