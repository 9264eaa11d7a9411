"""No-GPU checks of the drop-in boundary: the library loads and exports every symbol include/shifu_b200.h declares;
the ctypes structs match the C layout; creating a handle without a GPU fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from shifu_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "shifu_b200.h")


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    g.build()
    return capi.load_library()


def declared_symbols():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"SB_API\s+[A-Za-z_0-9\s\*]+?\b(sbh?_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_all_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), "libshifu_b200.so does not export %s" % s
    assert sorted(capi.EXPORTS) == syms, "capi.EXPORTS and the header disagree"


def test_struct_sizes_match_c(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "shifu_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(sb_config),sizeof(sb_node),sizeof(sb_tree),sizeof(sb_todo_node),sizeof(sb_iter_in),'
                   'sizeof(sb_iter_out),sizeof(sb_gain));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)])
    sizes = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    ours = [C.sizeof(t) for t in (capi.SbConfig, capi.SbNode, capi.SbTree, capi.SbTodoNode, capi.SbIterIn,
                                  capi.SbIterOut, capi.SbGain)]
    assert sizes == ours


def test_header_is_plain_c(tmp_path):
    src = tmp_path / "c.c"
    src.write_text('#include "shifu_b200.h"\nint main(void){return SB_OK;}\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                           "-c", "-o", str(tmp_path / "c.o"), str(src)])


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    cfg = capi.ConfigArrays([8, 8])
    h = C.c_void_p()
    rc = lib.sb_create(C.byref(cfg.struct), C.byref(h))
    assert rc == -2 and b"no CPU fallback" in lib.sb_last_error()


def test_product_never_imports_oracle():
    """The oracle is a checker; the product path must not import, include, link or dlopen it."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "shifu_b200")):
        for f in files:
            if not f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                continue
            for line in open(os.path.join(dirpath, f)):
                low = line.lower()
                if "oracle" in low:
                    assert not any(k in low for k in ("import", "#include", "cdll", "dlopen", "liboracle")), (f, line)


def test_cpp_host_program_on_the_bare_c_abi():
    """tools/train_gbt.cpp uses only include/shifu_b200.h.  Without a GPU it must stop with the library's own error
    (no CPU fallback); with one it grows its trees."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "tools", "train_gbt")
    if not os.path.exists(exe):
        subprocess.check_call(["bash", os.path.join(root, "build.sh")])
    p = subprocess.run([exe, "20000", "12", "16", "2", "4"], capture_output=True, text=True, timeout=300)
    if p.returncode == 0:
        assert "2 trees grown" in p.stdout
    else:
        assert "no CUDA device: shifu_b200 has no CPU fallback" in p.stderr
