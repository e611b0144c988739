"""The C ABI: structure layouts agree with the C compiler, libcrt.so loads on a
machine without a GPU and exports every symbol include/crt.h declares, and the
product path fails loudly (no CPU fallback) when no device is present."""
import ctypes as C
import os
import re
import subprocess

import pytest

from _util import ROOT, scene_path
from c_raytracer_b200 import _abi, device, host


def test_struct_sizes_match_c(tmp_path):
    src = tmp_path / "sizes.c"
    src.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "crt.h"\n'
        'int main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(crt_material), sizeof(crt_object),'
        ' sizeof(crt_scene_desc), sizeof(crt_render_opts), sizeof(crt_stats), sizeof(crt_scene_info),'
        ' offsetof(crt_render_opts, seed), offsetof(crt_stats, kernel_ms), sizeof(crt_post_opts), sizeof(crt_post_stats),'
        ' offsetof(crt_post_opts, mist_color), offsetof(crt_post_stats, dof_max_radius)); return 0;}\n')
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    assert got == [C.sizeof(_abi.Material), C.sizeof(_abi.Object), C.sizeof(_abi.SceneDesc), C.sizeof(_abi.RenderOpts),
                   C.sizeof(_abi.Stats), C.sizeof(_abi.SceneInfo), _abi.RenderOpts.seed.offset, _abi.Stats.kernel_ms.offset,
                   C.sizeof(_abi.PostOpts), C.sizeof(_abi.PostStats), _abi.PostOpts.mist_color.offset,
                   _abi.PostStats.dof_max_radius.offset]
    assert got[0] == 128 and got[1] == 64


def test_header_symbols_are_exported():
    if not os.path.exists(device.LIB_PATH):
        pytest.skip("libcrt.so not built (run __graft_entry__.build())")
    header = open(os.path.join(ROOT, "include", "crt.h")).read()
    declared = set(re.findall(r"\b(crt_[a-z0-9_]+)\s*\(", header))
    assert declared == set(device.SYMBOLS)
    L = device.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert L.crt_abi_version() == _abi.CRT_ABI_VERSION


def test_no_cpu_fallback_without_gpu():
    if not os.path.exists(device.LIB_PATH):
        pytest.skip("libcrt.so not built")
    if device.device_count() > 0:
        pytest.skip("a GPU is present")
    sc = host.Scene(scene_path("scene1"))
    with pytest.raises(device.CrtError, match="no CUDA device"):
        device.DeviceScene(sc, 0)


def test_product_does_not_reference_oracle():
    """Nothing under the package, src/ or include/ may name the oracle."""
    bad = []
    for base in ("c-raytracer_b200", "src", "include", "lib"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".so", ".o", ".pyc")):
                    continue
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"liboracle|oracle_render|from oracle|import oracle|oracle/", text):
                    bad.append(os.path.join(dirpath, f))
    # crt_rng.h documents the shared stream and mentions the shim's file name only in a comment
    bad = [b for b in bad if not b.endswith("crt_rng.h")]
    assert not bad, bad


def test_scene_coordinates_are_validated_before_any_device_work():
    """crt_scene_create refuses non-finite or huge coordinates (the tree builders form FP32 surfaces and Morton codes;
    surfaces that overflowed once overran a stack buffer in the 8-wide collapse): checked on the host, so the
    answer is the same with and without a GPU."""
    import ctypes as C

    import numpy as np
    if not os.path.exists(device.LIB_PATH):
        pytest.skip("libcrt.so not built")
    sc = host.Scene(scene_path("scene1"))
    objs = np.array(sc.objects)  # a copy of the 64-byte records
    for field, value in (("a", 1e13), ("a", np.inf), ("b", np.nan), ("a", -3e38)):
        bad = objs.copy()
        k = int(np.flatnonzero(bad["type"] == _abi.CRT_OBJ_SPHERE)[0])
        bad[field][k, 0] = value
        desc = _abi.SceneDesc.from_buffer_copy(sc.desc)
        desc.objects = C.cast(bad.ctypes.data, type(desc.objects))
        h = C.c_void_p()
        rc = device.lib().crt_scene_create(C.addressof(desc), 0, C.byref(h))
        assert rc == -5 and h.value is None, (field, value, rc)  # CRT_ERR_UNSUPPORTED
        assert b"coordinate" in device.lib().crt_last_error()


def test_dram_traffic_record_belongs_to_this_build():
    """profiles/dram_traffic.json (bench.py's roofline.traffic) names the device sources it was captured from, and any
    later sources that compile to the same device code; the SASS of the library built here is the recorded one."""
    import hashlib
    import json
    import shutil
    t = json.load(open(os.path.join(ROOT, "profiles", "dram_traffic.json")))
    same = [t["device_sources_sha256"]] + [e["device_sources_sha256"] for e in t.get("same_device_code", [])]
    if device.source_fingerprint() not in same:
        pytest.skip("the device sources changed after the capture: bench.py reports traffic null")
    if not os.path.exists(device.LIB_PATH) or not shutil.which("cuobjdump"):
        pytest.skip("libcrt.so or cuobjdump missing")
    sass = subprocess.run(["cuobjdump", "-sass", device.LIB_PATH], capture_output=True, check=True).stdout
    assert hashlib.sha256(sass).hexdigest() == t["sass_sha256"]
