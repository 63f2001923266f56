"""CPU: the C-ABI library loads and exports every symbol include/kssd_b200.h declares (no compute calls)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "kssd_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(kssd_b200_\w+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    from rabbitkssd_b200 import _lib
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
        assert n in _lib.SYMBOLS, f"{n} declared in the header but not bound in _lib.SYMBOLS"
    for n in _lib.SYMBOLS:
        assert n in names, f"{n} bound but not declared in include/kssd_b200.h"


def test_init_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from rabbitkssd_b200 import _lib, api
    with pytest.raises(_lib.KssdError) as e:
        api.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_shuffle_table_equals_reference_shuf_file():
    """Host-only entry point (no GPU needed): kssd_b200_shuffle_table == the table inside the .shuf file the unmodified
    reference binary wrote when the golden fixtures were made (SHA-256 pinned in tests/golden/golden.json)."""
    import numpy as np
    import util
    from rabbitkssd_b200 import _lib, api
    seen = set()
    for case, m in util.golden_meta().items():
        c = m["recipe"]
        key = (c["k"], c["s"], c["l"])
        if key in seen or c["s"] > 6:  # the s = 7 table is 1 GiB: covered through the oracle's copy in the GPU tests
            continue
        seen.add(key)
        dim = api.shuffle_table(*key)
        assert util.shuf_sha256(*key, dim) == m["shuf_sha256"], case
    assert len(seen) >= 3
    with pytest.raises(_lib.KssdError):
        api.shuffle_table(10, 8, 3)
    with pytest.raises(_lib.KssdError):
        np.asarray(api.shuffle_table(4, 6, 3))
