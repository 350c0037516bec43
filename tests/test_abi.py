"""The drop-in boundary: libb200attn.so loads, exports every symbol include/*.h declares (and nothing else with
the b200_ prefix), mirrors struct ggml_tensor byte for byte, and rejects malformed calls with the documented
status codes.  No compute is launched here (no GPU in this container)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "include")
REF = "/root/reference"


def _declared(header):
    with open(os.path.join(INC, header)) as f:
        text = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"B200_API[^;(]*?\b(b200_\w+)\s*\(", text)))


@pytest.mark.parametrize("header", ["b200_paged_attn.h", "b200_ggml_abi.h", "b200_host.h"])
def test_every_declared_symbol_is_exported(pkg, header):
    names = _declared(header)
    assert names == sorted(pkg.EXPORTS[header]), "driver binding table out of sync with the header"
    for n in names:
        assert hasattr(pkg.lib, n), f"{n} declared in {header} but not exported"


def test_no_undeclared_b200_exports(pkg):
    out = subprocess.run(["nm", "-D", "--defined-only", pkg.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l and l.split()[-1].startswith("b200_")}
    declared = {n for h in pkg.EXPORTS for n in pkg.EXPORTS[h]}
    assert exported == declared


def test_headers_compile_as_c_and_cxx(tmp_path):
    src = tmp_path / "probe.c"
    src.write_text('#include "b200_paged_attn.h"\n#include "b200_ggml_abi.h"\n#include "b200_host.h"\nint main(void){return 0;}\n')
    for cc, std in (("gcc", "-std=c99"), ("g++", "-std=c++11")):
        r = subprocess.run([cc, std, "-Wall", "-Werror", "-pedantic", "-I", INC, "-x", "c" if cc == "gcc" else "c++", str(src),
                            "-o", str(tmp_path / "probe")], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr


def test_struct_sizes_match_ctypes(pkg, tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "b200_ggml_abi.h"\nint main(void){printf("%zu %zu %zu\\n", '
                   'sizeof(b200_attn_desc), sizeof(b200_kv_write_desc), sizeof(struct b200_ggml_tensor));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", INC, str(src), "-o", str(exe)], check=True)
    a, k, t = map(int, subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split())
    assert (a, k, t) == (C.sizeof(pkg.AttnDesc), C.sizeof(pkg.KVWriteDesc), C.sizeof(pkg.GgmlTensor))
    assert t == 336


def test_ggml_tensor_mirror_matches_reference_header(tmp_path):
    """sizeof / offsetof / enum values of the mirror against the reference's own ggml.h (only where it exists)"""
    hdr = os.path.join(REF, "ggml", "include", "ggml.h")
    if not os.path.exists(hdr):
        pytest.skip("reference tree not present")
    fields = ["type", "buffer", "ne", "nb", "op", "op_params", "flags", "src", "view_src", "view_offs", "data", "name", "extra"]
    checks = "".join(f'  if (offsetof(struct ggml_tensor, {f}) != offsetof(struct b200_ggml_tensor, {f})) {{ printf("offset {f}\\n"); bad = 1; }}\n'
                     for f in fields)
    src = tmp_path / "probe.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "ggml.h"\n#include "b200_ggml_abi.h"\nint main(void){ int bad = 0;\n'
                   + checks +
                   '  if (sizeof(struct ggml_tensor) != sizeof(struct b200_ggml_tensor)) { printf("size\\n"); bad = 1; }\n'
                   '  if (GGML_OP_CPY != B200_GGML_OP_CPY || GGML_OP_DUP != B200_GGML_OP_DUP || GGML_OP_CONT != B200_GGML_OP_CONT ||\n'
                   '      GGML_OP_FLASH_ATTN_EXT != B200_GGML_OP_FLASH_ATTN_EXT) { printf("op enum\\n"); bad = 1; }\n'
                   '  if (GGML_TYPE_F32 != B200_TYPE_F32 || GGML_TYPE_F16 != B200_TYPE_F16 || GGML_TYPE_Q4_0 != B200_TYPE_Q4_0 ||\n'
                   '      GGML_TYPE_Q8_0 != B200_TYPE_Q8_0 || GGML_TYPE_I32 != B200_GGML_TYPE_I32) { printf("type enum\\n"); bad = 1; }\n'
                   '  if (GGML_MAX_DIMS != B200_GGML_MAX_DIMS || GGML_MAX_SRC != B200_GGML_MAX_SRC || GGML_MAX_OP_PARAMS != B200_GGML_MAX_OP_PARAMS ||\n'
                   '      GGML_MAX_NAME != B200_GGML_MAX_NAME || GGML_PREC_F32 != B200_GGML_PREC_F32) { printf("limits\\n"); bad = 1; }\n'
                   '  if (GGML_STATUS_SUCCESS != B200_OK || GGML_STATUS_FAILED != B200_ERR_FAILED || GGML_STATUS_ALLOC_FAILED != B200_ERR_ALLOC) { printf("status\\n"); bad = 1; }\n'
                   '  return bad; }\n')
    exe = tmp_path / "probe"
    r = subprocess.run(["gcc", "-I", os.path.dirname(hdr), "-I", INC, str(src), "-o", str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    run = subprocess.run([str(exe)], capture_output=True, text=True)
    assert run.returncode == 0, run.stdout


def test_malformed_calls_return_status_not_crash(pkg):
    lib = pkg.lib
    assert lib.b200_paged_attention(None, None) == -4                    # B200_ERR_INVALID
    assert b"NULL descriptor" in lib.b200_last_error()
    d = pkg.AttnDesc()
    assert lib.b200_paged_attention(C.byref(d), None) == -4              # NULL tensors
    assert lib.b200_supports_attention(C.byref(d)) == 0
    assert lib.b200_paged_attention_workspace(C.byref(d)) == 0
    w = pkg.KVWriteDesc()
    assert lib.b200_kv_write(None, None) == -4
    assert lib.b200_kv_write(C.byref(w), None) == -4                     # bad shape
    assert lib.b200_block_copy(None, 1, 16, 1, None, 1, None) == -4
    assert lib.b200_flash_attn_ext(None, None, 0, None) == -4
    assert lib.b200_cpy(None, None, None) == -4
    assert lib.b200_supports_op(None) == 0
    assert lib.b200_version().startswith(b"b200attn")


def test_supports_op_mirrors_reference_rules(pkg):
    """ggml_backend_cuda_device_supports_op (ggml-cuda.cu:3227-3245, :3080-3133): FLASH_ATTN_EXT node with F16 /
    Q8_0 / Q4_0 K,V at head size 128 is taken; a non-F32 q or a BF16 cache is not; CPY F32 -> cache types is taken"""
    T = pkg.GgmlTensor

    def tensor(type_id, ne, nb=None, op=0):
        t = T()
        t.type, t.op = type_id, op
        for i in range(4):
            t.ne[i] = ne[i] if i < len(ne) else 1
        rb = pkg.row_bytes(type_id, ne[0]) if type_id in (0, 1, 2, 8) else ne[0] * 4
        stride = [rb // max(ne[0], 1) * (32 if type_id in (2, 8) else 1) if False else 0] * 4
        t.nb[0] = {0: 4, 1: 2, 2: 18, 8: 34}.get(type_id, 4)
        t.nb[1] = rb
        for i in range(2, 4):
            t.nb[i] = t.nb[i - 1] * t.ne[i - 1]
        if nb:
            for i, v in enumerate(nb):
                t.nb[i] = v
        t.data = 0x1000
        return t

    for tkv in (pkg.TYPE_F16, pkg.TYPE_Q8_0, pkg.TYPE_Q4_0):
        q = tensor(0, [128, 8, 32])
        k = tensor(tkv, [128, 256, 8]); v = tensor(tkv, [128, 256, 8])
        m = tensor(1, [256, 64])
        dst = tensor(0, [128, 32, 8], op=pkg.GGML_OP_FLASH_ATTN_EXT)
        for i, s in enumerate((q, k, v, m)):
            dst.src[i] = C.pointer(s)
        assert pkg.lib.b200_supports_op(C.byref(dst)) == 1
        q.type = 1                                                        # F16 q: launch_fattn asserts F32 (fattn-common.cuh:691)
        assert pkg.lib.b200_supports_op(C.byref(dst)) == 0
        q.type = 0
        k.type = 30                                                       # BF16: excluded (ggml-cuda.cu:3233-3236)
        assert pkg.lib.b200_supports_op(C.byref(dst)) == 0
    src = tensor(0, [1024, 16])
    for tkv in (pkg.TYPE_F16, pkg.TYPE_Q8_0, pkg.TYPE_Q4_0, pkg.TYPE_F32):
        dstc = tensor(tkv, [1024, 16], op=pkg.GGML_OP_CPY)
        dstc.src[0] = C.pointer(src); dstc.src[1] = C.pointer(dstc)
        assert pkg.lib.b200_supports_op(C.byref(dstc)) == 1
    odd = tensor(0, [1000, 16])                                           # rows not a multiple of 32 -> quantised dst refused
    dstc = tensor(pkg.TYPE_Q8_0, [1024, 16], op=pkg.GGML_OP_CPY)
    dstc.src[0] = C.pointer(odd); dstc.src[1] = C.pointer(dstc)
    assert pkg.lib.b200_supports_op(C.byref(dstc)) == 0


def test_sass_of_the_hot_kernels_is_sm100a_tma_and_tcgen05(pkg, tmp_path):
    """The built library carries sm_100a SASS only, the decode kernel gathers with TMA (UTMALDG) and computes with HMMA,
    the prefill tile kernel issues tcgen05 MMAs (UTCHMMA) with TMEM loads (the mnemonics
    /opt/skills/guides/B200_PROFILING.md names as proof) — a build-level guard against a silent fallback path"""
    import shutil
    import subprocess
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        pytest.skip("cuobjdump not available")
    elf = subprocess.run([exe, "-lelf", pkg.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in elf and not any(a in elf for a in ("sm_90", "sm_80", "sm_100.")), elf
    sass = subprocess.run([exe, "-sass", pkg.LIB_PATH], capture_output=True, text=True).stdout
    funcs, cur = {}, None
    for line in sass.splitlines():
        if "Function :" in line:
            cur = line.split("Function :")[1].strip()
            funcs[cur] = set()
        elif cur:
            for m in ("UTMALDG", "UTCHMMA", "HMMA.16816", "LDTM", "STTM", "SYNCS"):
                if m in line:
                    funcs[cur].add(m)
    sk = [f for f in funcs if "sk_main_kernel" in f]
    assert len(sk) == 12 and all({"UTMALDG", "HMMA.16816", "SYNCS"} <= funcs[f] for f in sk), {f: funcs[f] for f in sk}   # 9 K/V type pairs + 3 dense-mask forms
    for name in ("pf_main_kernel",):
        fs = [f for f in funcs if name in f]
        assert fs and all({"UTMALDG", "UTCHMMA", "LDTM"} <= funcs[f] for f in fs), {f: funcs[f] for f in fs}


def test_product_sources_never_reach_into_the_oracle():
    """oracle/ is the checker: nothing under llama.cpp_test_b200/ or include/ may import, include, dlopen or exec it (only tests/,
    __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do)"""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    bad = []
    for base in ("llama.cpp_test_b200", "include"):
        for dirpath, dirnames, files in os.walk(os.path.join(root, base)):
            dirnames[:] = [d for d in dirnames if d not in ("build", "__pycache__")]
            for f in files:
                if not f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".c", "Makefile")):
                    continue
                text = open(os.path.join(dirpath, f), errors="replace").read()
                for m in re.finditer(r"(import\s+oracle|from\s+oracle|oracle/|liboracle|libref_|refkv)", text):
                    line = text[text.rfind("\n", 0, m.start()) + 1:text.find("\n", m.end())]
                    if not line.lstrip().startswith(("//", "#", "*", "/*")) and '"""' not in line:
                        bad.append((f, line.strip()))
    assert not bad, bad
    # and bench.py only touches it in the two CPU legs
    bench = open(os.path.join(root, "bench.py")).read()
    for m in re.finditer(r"from oracle|import oracle", bench):
        ctx = bench[max(0, bench.rfind("\ndef ", 0, m.start())):m.start()]
        assert re.search(r"def \w*(cpu|reference)\w*\(", ctx), ctx[:80]         # cpu_layer_op & co: the CPU-baseline / reference legs
