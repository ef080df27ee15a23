"""CPU: the C-ABI library builds, loads and exports every symbol include/*.h declares."""
import ctypes
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def _declared_symbols():
    names = []
    for h in sorted((ROOT / "include").glob("*.h")):
        text = re.sub(r"/\*.*?\*/", "", h.read_text(), flags=re.S)
        names += re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_library_exports_every_declared_symbol():
    from wenet_celoss_b200 import build
    so = build.build_library()
    lib = ctypes.CDLL(str(so))
    names = _declared_symbols()
    assert len(names) >= 8
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ but not exported"
    lib.b200_version.restype = ctypes.c_int
    assert lib.b200_version() >= 100
    lib.b200_compiled_arch.restype = ctypes.c_int
    assert lib.b200_compiled_arch() == 100


def test_library_is_sm100a_only():
    import subprocess
    from wenet_celoss_b200 import build
    so = build.build_library()
    out = subprocess.run(["cuobjdump", "-lelf", str(so)], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_argument_errors_do_not_need_a_gpu():
    from wenet_celoss_b200 import _lib
    L = _lib.lib()
    L.b200_rnnt_lattice.restype = ctypes.c_int
    rc = L.b200_rnnt_lattice(None, None, None, None, 1, 1, 1, None, None, None, None, None)
    assert rc == 1
    assert b"null pointer" in L.b200_last_error()


def test_product_does_not_import_oracle():
    pkg = ROOT / "wenet_celoss_b200"
    for f in pkg.rglob("*.py"):
        src = f.read_text()
        assert "oracle" not in re.sub(r'""".*?"""', "", src, flags=re.S).replace("build_oracle", "") \
            or f.name == "build.py", f"{f} references the oracle"


def test_round2_entry_points_validate_arguments_without_a_gpu():
    """The entry points added in round 2 (no-recompute backward, rescoring, fp32 twin) reject null /
    inconsistent arguments with status 1 and a message, before touching the device."""
    from wenet_celoss_b200 import _lib
    L = _lib.lib()
    c = ctypes
    cases = [
        ("b200_joint_fwd_p", (None,) * 7 + (1, 1, 1, 8, 64, 0) + (None,) * 5 + (64, None, 16, 0, None), b"null pointer"),
        ("b200_joint_rowscale", (None,) * 5 + (1, 1, 1, 8, 64, 0, None, 16, None, 64, None, None, None), b"null pointer"),
        ("b200_joint_dw_db", (None, 16, None, 64, c.c_longlong(4), 8, 128, None, None, None, 0, None), b"null pointer"),
        ("b200_joint_dh_scaled", (None, 16, None, None, 64, None, 1, 1, 1, 8, 256, None, None, None, None, 0, None),
         b"null pointer"),
        ("b200_attn_rescore", (None, None, None, None, 1, 1, 0, 8, 0, c.c_double(0), c.c_double(1), None,
                               c.c_double(0), None, c.c_double(0), None, None, None, None), b"null pointer"),
        ("b200_joint_h", (None, None, 1, 1, 1, 64, None, 64, None), b"null pointer"),
        ("b200_joint_rowscale_split", (None,) * 3 + (1, 1, 1, 8, 64, 0, None, None, c.c_longlong(16), None, None, 64,
                                                      None, None, None, None), b"null pointer"),
        ("b200_colsum_split_w", (None, None, c.c_longlong(1), 8, c.c_longlong(8), None, None, None), b"null"),
    ]
    for name, args, needle in cases:
        fn = getattr(L, name)
        fn.restype = ctypes.c_int
        rc = fn(*args)
        assert rc == 1, (name, rc)
        assert needle in L.b200_last_error(), (name, L.b200_last_error())
