"""CPU tests of the C-ABI boundary: the library builds for sm_100a, loads, exports every symbol
include/merl_b200.h declares, and refuses to compute without a GPU (no fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "merl_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(merl_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported(built_lib):
    from mitsuba_customization_b200 import _cabi
    declared = _declared_symbols()
    assert declared, "no symbols parsed from the header"
    for name in declared:
        assert hasattr(built_lib, name), f"{name} declared in merl_b200.h but not exported"
    assert sorted(_cabi.SYMBOLS) == declared, "the ctypes binding must track the header"


def test_library_is_sm100a_only(built_lib):
    import subprocess
    from mitsuba_customization_b200 import _cabi
    out = subprocess.run(["cuobjdump", "-lelf", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out and not re.search(r"sm_(?!100a)\d+", out), out


def test_abi_version_and_defaults(built_lib):
    from mitsuba_customization_b200 import _cabi
    assert built_lib.merl_abi_version() == 2
    p = _cabi.default_params()
    assert (p.nh, p.nd, p.np, p.theta_h_sqrt, p.ni) == (90, 90, 180, 1, 32)
    np.testing.assert_allclose(list(p.scale), [1 / 1500, 1.15 / 1500, 1.66 / 1500])


def test_struct_layouts_match_header(built_lib):
    """ctypes mirrors are laid out exactly as the C structs (sizes from a tiny C probe)."""
    import subprocess
    import tempfile
    src = ('#include "merl_b200.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu %zu\\n", '
           'sizeof(merl_table_params_t), sizeof(merl_table_info_t), sizeof(merl_render_params_t), '
           'sizeof(merl_peerbuf_info_t), sizeof(merl_file_report_t));return 0;}\n')
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "p.c"), "w").write(src)
        subprocess.run(["/usr/bin/gcc", "-I", os.path.join(ROOT, "include"), os.path.join(d, "p.c"), "-o",
                        os.path.join(d, "p")], check=True)
        a, b, c, e, f = (int(x) for x in subprocess.run([os.path.join(d, "p")], capture_output=True, text=True).stdout.split())
    from mitsuba_customization_b200 import _cabi, render
    assert C.sizeof(_cabi.TableParams) == a and C.sizeof(_cabi.TableInfo) == b
    assert C.sizeof(render._Params) == c and C.sizeof(_cabi.PeerbufInfo) == e and C.sizeof(_cabi.FileReport) == f


def test_argument_errors_without_gpu(built_lib):
    from mitsuba_customization_b200 import _cabi
    h = C.c_void_p()
    rc = built_lib.merl_table_create_from_file(b"/nonexistent/file.binary", None, 0, C.byref(h))
    assert rc == -2 and b"cannot open" in built_lib.merl_last_error()
    assert built_lib.merl_set_sample_search(7) == -1
    assert built_lib.merl_peerbuf_create(None, 16, 0, C.byref(h)) == -1
    assert built_lib.merl_peerbuf_allreduce(None, 16, None, None) == -1
    assert built_lib.merl_grad_apply_mask(None, None, None) == -1
    assert built_lib.merl_bins(None, None, None, None, None, None, None, None, 4, None, None) == -1
    with pytest.raises(_cabi.MerlError):
        _cabi.check(-1)


def test_no_cpu_fallback():
    """Without a CUDA device the plugin refuses to construct; nothing routes through the oracle."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mitsuba_customization_b200 import bsdf, synth
    with pytest.raises(bsdf.MerlError):
        bsdf.MERL(data=synth.synthetic_table(0))
    # the product package never imports the oracle
    pkg = os.path.join(ROOT, "mitsuba_customization_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                assert "oracle" not in open(os.path.join(dirpath, f)).read().replace("the oracle", "").replace(
                    "oracle:", "").replace("oracle's", "").replace("CPU oracle", ""), f


def test_header_is_plain_c(tmp_path):
    """The boundary is a C ABI: the header must compile as strict C99 (what a C / cgo / JNI stub includes),
    and a C program must link against the library without a C++ runtime in its own link line."""
    import shutil
    import subprocess
    gcc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else shutil.which("gcc")
    src = tmp_path / "c99.c"
    src.write_text('#include "merl_b200.h"\n'
                   'int main(void) { merl_table_params_t p; merl_table_params_default(&p);\n'
                   '  return (p.nh == 90 && merl_abi_version() > 0 && merl_last_error() != 0) ? 0 : 1; }\n')
    inc = os.path.join(ROOT, "include")
    r = subprocess.run([gcc, "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-I", inc, "-fsyntax-only", str(src)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    libdir = os.path.join(ROOT, "mitsuba_customization_b200")
    exe = tmp_path / "c99"
    r = subprocess.run([gcc, "-std=c99", "-I", inc, str(src), "-o", str(exe), "-L", libdir, "-lmerl_b200",
                        f"-Wl,-rpath,{libdir}"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert subprocess.run([str(exe)]).returncode == 0        # defaults and diagnostics work without a GPU


def test_corrupt_file_header_is_an_error_not_an_exception(built_lib, tmp_path):
    """A header that multiplies to an absurd cell count must come back as a status code before anything
    is allocated for it (bad_alloc used to escape the extern "C" boundary)."""
    h = C.c_void_p()
    from mitsuba_customization_b200 import _cabi
    p = _cabi.default_params()
    p.nh = p.nd = p.np = 0                      # customized_measurment: resolution from the header
    bad = tmp_path / "huge.binary"
    bad.write_bytes(np.array([1 << 20, 1 << 20, 1 << 20], dtype=np.int32).tobytes() + b"\0" * 64)
    assert built_lib.merl_table_create_from_file(str(bad).encode(), C.byref(p), 0, C.byref(h)) in (-2, -5)
    neg = tmp_path / "neg.binary"
    neg.write_bytes(np.array([-3, 90, 180], dtype=np.int32).tobytes())
    assert built_lib.merl_table_create_from_file(str(neg).encode(), C.byref(p), 0, C.byref(h)) == -2
    short = tmp_path / "short.binary"            # plausible header, payload missing
    short.write_bytes(np.array([64, 96, 192], dtype=np.int32).tobytes() + b"\0" * 1000)
    rc = built_lib.merl_table_create_from_file(str(short).encode(), C.byref(p), 0, C.byref(h))
    assert rc == -2 and b"needs" in built_lib.merl_last_error()


def test_python_constants_match_header():
    """The flag / transport / error values the ctypes binding hard-codes are the header's."""
    import re
    from mitsuba_customization_b200 import _cabi
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "merl_b200.h")).read()

    def define(name):
        m = re.search(r"#define\s+%s\s+(-?\d+)u?\b" % name, hdr)
        assert m, f"{name} not #defined in merl_b200.h"
        return int(m.group(1))
    assert define("MERL_BWD_WARP_AGGREGATE") == _cabi.BWD_WARP_AGGREGATE
    assert define("MERL_BWD_DEFER_MASK") == _cabi.BWD_DEFER_MASK
    assert define("MERL_BWD_CTA_AGGREGATE") == _cabi.BWD_CTA_AGGREGATE
    assert (define("MERL_XCHG_PEER"), define("MERL_XCHG_NCCL"), define("MERL_XCHG_NVLS")) == \
        (_cabi.XCHG_PEER, _cabi.XCHG_NCCL, _cabi.XCHG_NVLS)
    assert define("MERL_STATIC_STAGE_MAX_QUERIES") == _cabi.STATIC_STAGE_MAX_QUERIES
    m = re.search(r"MERL_ERR_UNSUPPORTED\s*=\s*(-?\d+)", hdr)
    assert m and int(m.group(1)) == _cabi.ERR_UNSUPPORTED


def test_unknown_exchange_transport_is_rejected(built_lib):
    """merl_peerbuf_create checks its arguments before touching NCCL or the device."""
    import ctypes as C
    from mitsuba_customization_b200 import _cabi
    h = C.c_void_p()
    assert _cabi.lib().merl_peerbuf_create(None, 16, _cabi.XCHG_PEER, C.byref(h)) != 0      # no communicator
    assert b"argument" in _cabi.lib().merl_last_error().lower()
