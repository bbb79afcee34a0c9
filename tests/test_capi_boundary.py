"""The C-ABI library loads and exports every symbol include/gravomg_b200.h declares; without a GPU
the product path fails loudly instead of falling back to anything. CPU only, no compute calls."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    with open(os.path.join(ROOT, "include", "gravomg_b200.h")) as f:
        text = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(gmg_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from gravomg_b200 import capi
    lib = capi.load_library()
    names = declared_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), f"{name} is declared in include/gravomg_b200.h but not exported"


def test_no_silent_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from gravomg_b200.capi import Context, GmgError
    with pytest.raises(GmgError) as e:
        Context(0)
    assert e.value.status == 1   # GMG_ERR_CUDA


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "gravo-mg_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                with open(os.path.join(dirpath, fn)) as f:
                    src = f.read()
                assert "pyoracle" not in src and "liboracle" not in src and "oracle_api" not in src, fn


def test_meshgen_is_deterministic_and_symmetric():
    import numpy as np
    from gravomg_b200 import meshgen
    a = meshgen.make_config("C2", 1 / 1024)
    b = meshgen.make_config("C2", 1 / 1024)
    assert np.array_equal(a["V"], b["V"]) and np.array_equal(a["A"].indices, b["A"].indices)
    A = a["A"]
    pat = A.copy(); pat.data[:] = 1
    assert (pat != pat.T).nnz == 0
    assert np.all(A.diagonal() != 0)
    assert A.indices.dtype == np.int32 and A.has_sorted_indices
    u = meshgen.uniform_pm1(42, np.arange(5, dtype=np.uint64), 1)
    assert np.all((u >= -1) & (u < 1))
    assert np.array_equal(u, meshgen.uniform_pm1(42, np.arange(5, dtype=np.uint64), 1))


def test_ctypes_structs_match_the_c_header(tmp_path):
    """The ctypes mirrors of the C structs must have the header's layout (a stale mirror reads garbage past the
    end of a smaller struct): compile a C program against include/gravomg_b200.h and compare sizes and offsets."""
    import ctypes as C
    import os
    import subprocess
    from gravomg_b200 import capi
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "layout.c"
    src.write_text('''#include <stdio.h>
#include <stddef.h>
#include "gravomg_b200.h"
int main(void) {
    printf("%zu %zu %zu %zu %zu\\n", sizeof(gmg_hierarchy_options), offsetof(gmg_hierarchy_options, radius_override),
           offsetof(gmg_hierarchy_options, build_galerkin), offsetof(gmg_hierarchy_options, on_level),
           offsetof(gmg_hierarchy_options, user));
    printf("%zu %zu %zu\\n", sizeof(gmg_level_info), offsetof(gmg_level_info, radius), offsetof(gmg_level_info, stage_ms));
    printf("%zu\\n", sizeof(gmg_prolongation_stats));
    return 0;
}
''')
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c11", "-I", os.path.join(root, "include"), str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()
    got = [int(v) for v in out]
    O, L = capi.HierarchyOptions, capi.LevelInfo
    want = [C.sizeof(O), O.radius_override.offset, O.build_galerkin.offset, O.on_level.offset, O.user.offset,
            C.sizeof(L), L.radius.offset, L.stage_ms.offset, C.sizeof(capi.ProlongationStats)]
    assert got == want, (got, want)
