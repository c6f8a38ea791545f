"""CPU checks of the drop-in boundary: the C-ABI library builds, loads without a GPU, exports every symbol that
include/svtme_cuda.h declares, and the ctypes mirror of its POD structs has the C layout.  No compute calls here."""
import ctypes as C
import os
import re
import subprocess

import pytest

from svtme import abi, api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "svtme_cuda.h")


@pytest.fixture(scope="module")
def lib():
    api.build()
    return api.load_library()


def test_every_declared_symbol_is_exported(lib):
    text = open(HEADER).read()
    declared = set(re.findall(r"^\s*(?:int32_t|uint32_t|uint64_t|void|const char \*)\s*\*?\s*(svtme_\w+)\s*\(", text, re.M))
    assert len(declared) >= 20, declared
    for name in sorted(declared):
        assert hasattr(lib, name), "libsvtme_cuda.so does not export %s" % name
    assert declared == set(api.EXPORTS), declared ^ set(api.EXPORTS)


def test_struct_layouts_match_the_header(tmp_path):
    """sizeof/offsetof of the POD structs as gcc sees the header == the ctypes mirror."""
    fields = {"SvtMeKnobs": abi.Knobs, "SvtMePresetCfg": abi.PresetCfg, "SvtMePictureResults": abi.PictureResults,
              "SvtMeDeviceCfg": abi.DeviceCfg, "SvtMeTfCfg": abi.TfCfg, "SvtMeDgMetrics": abi.DgMetrics, "SvtMePicStats": abi.PicStats}
    src = ['#include <stdio.h>', '#include <stddef.h>', '#include "svtme_cuda.h"', "int main(void) {"]
    for cname, cls in fields.items():
        src.append('printf("%s %%zu\\n", sizeof(%s));' % (cname, cname))
        for f, _ in cls._fields_:
            src.append('printf("%s.%s %%zu\\n", offsetof(%s, %s));' % (cname, f, cname, f))
    src.append("return 0; }")
    c = tmp_path / "layout.c"
    c.write_text("\n".join(src))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(c)])
    got = dict(line.split() for line in subprocess.check_output([str(exe)]).decode().splitlines())
    for cname, cls in fields.items():
        assert int(got[cname]) == C.sizeof(cls), cname
        for f, _ in cls._fields_:
            assert int(got["%s.%s" % (cname, f)]) == getattr(cls, f).offset, "%s.%s" % (cname, f)


def test_header_is_plain_c(tmp_path):
    c = tmp_path / "inc.c"
    c.write_text('#include "svtme_cuda.h"\nint main(void){return sizeof(SvtMeKnobs) > 0 ? 0 : 1;}\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c", str(c),
                           "-o", str(tmp_path / "inc.o")])


def test_no_cpu_fallback_without_a_device(lib):
    """On a box without CUDA the handle cannot be created: the product path fails loudly instead of falling back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(api.SvtMeError):
        api.MeDevice(device=0)


def test_product_does_not_touch_the_oracle():
    """Nothing under the package includes, imports, links or loads anything from oracle/ or the emulation backend."""
    pkg = os.path.join(ROOT, "svt-av1-mirror_b200")
    bad = re.compile(r'#include\s+"[^"]*(oracle|host_emul)|import\s+refbind|from\s+refbind|svtorc_|svtref_|svtemu_|'
                     r'libsvtme_oracle|libsvtref|libsvtme_emul')
    n = 0
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")) or f == "Makefile":
                n += 1
                m = bad.search(open(os.path.join(d, f), errors="replace").read())
                assert m is None, "%s: %s" % (f, m.group(0))
    assert n >= 8
