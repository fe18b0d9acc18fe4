"""The Mitsuba-side plugin sources (integration/src/integrators: b200phase / b200range / b200time + the shared flattening header;
SURVEY.md 8(f) row 1) type-check: `g++ -fsyntax-only` against integration/shim, a stand-in that declares the reference's API with
its own names and signatures (the real headers need Enoki, which the reference archive lacks).  Each plugin is explicitly
instantiated by the shim's MTS_EXPORT_PLUGIN, so every member the sources touch is checked."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "integration", "src", "integrators")


@pytest.mark.parametrize("plugin", ["b200phase", "b200range", "b200time"])
def test_plugin_source_type_checks(plugin):
    out = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "integration", "shim"),
                          "-I" + os.path.join(ROOT, "include"), "-I" + SRC, os.path.join(SRC, plugin + ".cpp")], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr


def test_every_abi_call_of_the_plugins_is_declared():
    """The plugin sources only call entry points include/beifong_b200/abi.h declares (and the library exports)."""
    import re
    from beifong_b200 import _native
    text = "".join(open(os.path.join(SRC, f)).read() for f in os.listdir(SRC))
    called = {c for c in re.findall(r"\b(bf_[a-z_]+)\s*\(", text) if not c.endswith("_desc")}      # "bf_bsdf_desc (...)" in a comment is a type
    assert called and called <= set(_native.PROTOTYPES), called - set(_native.PROTOTYPES)
    assert {"bf_scene_add_mesh", "bf_scene_add_rect", "bf_scene_add_transmitter", "bf_scene_set_receiver", "bf_scene_set_adc", "bf_receive", "bf_profile_read"} <= called
