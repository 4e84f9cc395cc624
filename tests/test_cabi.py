"""CPU-side checks of the drop-in boundary: the library loads without a GPU, exports every symbol
include/volrender.h declares, and refuses to work (loudly) when there is no CUDA device."""
from __future__ import annotations

import ctypes as C
import os
import re

import pytest

from conftest import ROOT, load_pkg_module

HEADER = os.path.join(ROOT, "include", "volrender.h")


@pytest.fixture(scope="module")
def vr():
    mod = load_pkg_module("volrender")
    mod.build()
    return mod


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vr_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported(vr):
    L = vr.lib()
    decl = _declared_symbols()
    assert len(decl) >= 30
    for name in decl:
        assert hasattr(L, name), f"{name} declared in include/volrender.h but not exported"
    assert sorted(vr.SYMBOLS) == decl


def test_struct_layouts_match_header(vr):
    # vr_render_params: 15 int32/float fields + 4 reserved; vr_render_stats: 2 doubles + 5 u64 + 2 u32 + 6 u64 (executed work)
    assert C.sizeof(vr.RenderParams) == 19 * 4
    assert C.sizeof(vr.RenderStats) == 2 * 8 + 5 * 8 + 2 * 4 + 6 * 8
    # the C compiler's view of the same structs
    import subprocess, tempfile
    src = '#include "volrender.h"\n#include <stdio.h>\nint main(void){printf("%zu %zu\\n", sizeof(vr_render_params), sizeof(vr_render_stats));return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "s.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), os.path.join(d, "s.c"), "-o", os.path.join(d, "s")])
        a, b = subprocess.check_output([os.path.join(d, "s")], text=True).split()
    assert int(a) == C.sizeof(vr.RenderParams) and int(b) == C.sizeof(vr.RenderStats)


def test_no_cpu_fallback(vr):
    L = vr.lib()
    assert L.vr_version().decode().startswith("volrender_b200")
    if L.vr_device_count() == 0:
        with pytest.raises(vr.VolRenderError) as e:
            vr.Context(0)
        assert e.value.code == 3  # VR_ERR_NO_DEVICE
    else:
        ctx = vr.Context(0)
        with pytest.raises(vr.VolRenderError):
            ctx.render(vr.Context.params(vr.INTEGRATOR_MS, 8, 8, 4))  # no scene / camera yet
        ctx.close()


def test_product_does_not_touch_the_oracle():
    """The oracle is test infrastructure: nothing under the package may import, link or exec it."""
    pkg = os.path.join(ROOT, "3dg-vol-renderer_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liboracle" not in text and "oracle_binding" not in text and "ref_driver" not in text, f
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f


def test_documented_options_exist_in_the_library():
    """every option of INTEGRATION.md section 4 is a name vr_set_option accepts (csrc/vr_api.cu), and every name it accepts is
    documented there or named as a tuning / debug option below"""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    doc = open(os.path.join(root, "INTEGRATION.md")).read()
    sec = doc[doc.index("## 4. Options"):]
    sec = sec[:sec.index("\n## ", 5)] if "\n## " in sec[5:] else sec
    documented = set(re.findall(r"^\| `([a-z0-9_]+)` \|", sec, flags=re.M)) | set(re.findall(r"`([a-z0-9_]+_min_samples|grid3d_res)`", sec))
    src = open(os.path.join(root, "3dg-vol-renderer_b200", "csrc", "vr_api.cu")).read()
    accepted = set(re.findall(r'strcmp\(name, "([a-z0-9_]+)"\)', src))
    assert documented, "no options found in INTEGRATION.md"
    missing = documented - accepted
    assert not missing, f"documented but not accepted by vr_set_option: {sorted(missing)}"
    tuning = {"max_pass_samples", "blocks_per_sm"}   # include/volrender.h documents these beside vr_set_option
    undocumented = accepted - documented - tuning
    assert not undocumented, f"accepted by vr_set_option but not in INTEGRATION.md section 4: {sorted(undocumented)}"
