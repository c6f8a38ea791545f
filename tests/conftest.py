import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "svt-av1-mirror_b200"), os.path.join(ROOT, "oracle"), ROOT, os.path.dirname(__file__)):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """The plain-C restatement (oracle/libsvtme_oracle.so); built on demand with gcc."""
    import refbind
    so = os.path.join(ROOT, "oracle", "libsvtme_oracle.so")
    src = os.path.join(ROOT, "oracle", "svtme_oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"])
    return refbind.Oracle()


@pytest.fixture(scope="session")
def reference():
    """The compiled reference (oracle/_ref/libsvtref.so).  Built here when /root/reference exists; the GPU box uses
    the prebuilt library that travels with the snapshot."""
    import refbind
    if os.path.exists("/root/reference/Source/Lib/Codec/motion_estimation.c"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], stdout=subprocess.DEVNULL)
    if not refbind.have_reference():
        pytest.skip("compiled reference (oracle/_ref/libsvtref.so) not available")
    return refbind.Reference()
