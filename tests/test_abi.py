"""The C-ABI library loads and exports every symbol include/dmhot.h declares (no compute without a GPU),
the ctypes structures match the header's field order, and the product fails loudly without a CUDA device."""
import ctypes as C
import os
import re

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header():
    with open(os.path.join(REPO, "include", "dmhot.h")) as f:
        return f.read()


def test_every_declared_symbol_is_exported():
    from deepmimic_mujoco_b200 import _native
    decl = set(re.findall(r"^(?:int|const char\*)\s+(dm_\w+)\s*\(", _header(), flags=re.M))
    assert len(decl) >= 20
    assert decl == set(_native.SYMBOLS), decl ^ set(_native.SYMBOLS)
    L = _native.lib()
    for s in decl:
        assert hasattr(L, s), s
    assert L.dm_abi_version() == int(re.search(r"#define DM_ABI_VERSION (\d+)", _header()).group(1))


def _struct_fields(name):
    h = _header()
    end = h.index("} %s;" % name)
    body = h[h.rindex("typedef struct {", 0, end) + len("typedef struct {"):end]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    out = []
    for stmt in body.split(";"):
        stmt = stmt.strip()
        if not stmt:
            continue
        names = [re.sub(r"[\*\s]", "", x.split()[-1]) for x in stmt.split(",")]
        out += names
    return out


@pytest.mark.parametrize("cname,pyname", [("dm_step_args", "StepArgs"), ("dm_reset_args", "ResetArgs"), ("dm_clip_t", "Clip"),
                                          ("dm_dims_t", "Dims"), ("dm_raw_clip_t", "RawClip"), ("dm_policy_t", "Policy"),
                                          ("dm_policy_args", "PolicyArgs")])
def test_ctypes_structs_follow_the_header(cname, pyname):
    from deepmimic_mujoco_b200 import _native
    assert [f[0] for f in getattr(_native, pyname)._fields_] == _struct_fields(cname)


def test_signatures_have_no_torch_types():
    h = _header()
    assert "torch" not in h and "at::" not in h and 'extern "C"' in h


def test_no_cpu_fallback():
    """Without a CUDA device dm_create fails with DM_E_CUDA and HotPath raises; with one this test is skipped."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a machine without a GPU")
    from deepmimic_mujoco_b200 import _native
    from deepmimic_mujoco_b200.hotpath import HotPath
    h = C.c_void_p()
    assert _native.lib().dm_create(0, 0, 0, C.byref(h)) == -2
    assert b"no CUDA device" in _native.lib().dm_last_error(None)
    with pytest.raises(_native.DmError):
        HotPath("humanoid3d")


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(REPO, "deepmimic_mujoco_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(root, f)) as fh:
                    src = fh.read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f


def test_ctypes_struct_sizes_match_a_c_compiler(tmp_path):
    """sizeof / offsetof of every argument block as gcc lays them out == ctypes' layout (padding included)."""
    import subprocess
    from deepmimic_mujoco_b200 import _native
    pairs = [("dm_step_args", "StepArgs"), ("dm_reset_args", "ResetArgs"), ("dm_clip_t", "Clip"), ("dm_dims_t", "Dims"),
             ("dm_raw_clip_t", "RawClip"), ("dm_policy_t", "Policy"), ("dm_policy_args", "PolicyArgs")]
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "dmhot.h"', 'int main(void) {']
    for cname, pyname in pairs:
        lines.append('printf("%s %%zu", sizeof(%s));' % (cname, cname))
        for f, _ in getattr(_native, pyname)._fields_:
            lines.append('printf(" %%zu", offsetof(%s, %s));' % (cname, f))
        lines.append('printf("\\n");')
    lines += ["return 0;", "}"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(REPO, "include"), "-o", str(exe), str(src)])
    out = subprocess.check_output([str(exe)], text=True).strip().splitlines()
    for (cname, pyname), line in zip(pairs, out):
        nums = [int(x) for x in line.split()[1:]]
        cls = getattr(_native, pyname)
        assert nums[0] == C.sizeof(cls), cname
        assert nums[1:] == [getattr(cls, f).offset for f, _ in cls._fields_], cname
