"""The C-ABI shared library loads and exports every symbol include/synthpop_cuda.h declares (no GPU needed)."""
import ctypes as C
import os

from synthpop import lib as splib


def test_library_exports_every_declared_symbol():
    assert os.path.exists(splib.LIB_PATH), "libsynthpop_cuda.so is not built (run __graft_entry__.build())"
    dll = C.CDLL(splib.LIB_PATH)
    names = splib.declared_symbols()
    assert len(names) >= 15
    missing = [n for n in names if not hasattr(dll, n)]
    assert not missing, missing


def test_binding_declares_prototypes():
    lib = splib.load()
    assert lib.sp_version().startswith(b"synthpop-cuda")


def test_struct_layouts_of_the_bindings_match_the_header(tmp_path):
    """The argument structs cross the C ABI by layout: the ctypes mirror (synthpop/lib.py) and the byte offsets the Java FFM
    shims write at (population-sv-simulator_b200/java/*.java) must be what gcc makes of include/synthpop_cuda.h."""
    import re
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    structs = {"sp_build_model_args": splib.BuildModelArgs, "sp_annotate_args": splib.AnnotateArgs, "sp_build_bins_args": splib.BuildBinsArgs}
    src = ['#include <stdio.h>', '#include <stddef.h>', '#include "include/synthpop_cuda.h"', 'int main(void){']
    for name, cls in structs.items():
        src.append(f'printf("{name} size %zu\\n", sizeof({name}));')
        for f, _ in cls._fields_:
            src.append(f'printf("{name} {f} %zu\\n", offsetof({name}, {f}));')
    src.append("return 0;}")
    (tmp_path / "off.c").write_text("\n".join(src))
    subprocess.run(["gcc", "-I", root, "-o", str(tmp_path / "off"), str(tmp_path / "off.c")], check=True)
    want = {}
    for ln in subprocess.run([str(tmp_path / "off")], check=True, capture_output=True, text=True).stdout.splitlines():
        s, f, v = ln.split()
        want[(s, f)] = int(v)
    for name, cls in structs.items():
        assert C.sizeof(cls) == want[(name, "size")], name
        for f, _ in cls._fields_:
            assert getattr(cls, f).offset == want[(name, f)], (name, f)
    # the Java shims set fields by literal offsets, in the order of the struct
    java = os.path.join(root, "population-sv-simulator_b200", "java")
    for cls_file, name in (("BuildModel.java", "sp_build_model_args"), ("AnnotateVCF.java", "sp_annotate_args"), ("BuildReadLengthBins.java", "sp_build_bins_args")):
        offs = [int(x) for x in re.findall(r"a\.set\([A-Z_]+, (\d+),", open(os.path.join(java, cls_file)).read())]
        assert offs == [want[(name, f)] for f, _ in structs[name]._fields_], cls_file
