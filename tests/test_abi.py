"""The C-ABI library builds, loads and exports exactly what include/adlm_head.h declares (no compute:
this runs without a GPU)."""
import ctypes
import os
import re
import subprocess

import pytest

from adlm_inherently_interpretable_segmentation_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "adlm_head.h")


def declared():
    """name -> number of parameters, parsed from the header"""
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"ADLM_API\s+[\w\s\*]+?\b(adlm_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else len(args.split(","))
    return out


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _lib.load()


def test_header_declares_the_binding_table():
    d = declared()
    assert set(d) == set(_lib.SIGNATURES), (sorted(d), sorted(_lib.SIGNATURES))
    for name, n_args in d.items():
        assert n_args == len(_lib.SIGNATURES[name][1]), name


def test_library_exports_every_declared_symbol(lib):
    for name in declared():
        assert hasattr(lib, name), name
    nm = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in nm.splitlines() if " T " in l}
    assert {e for e in exported if e.startswith("adlm_")} == set(declared())


def test_abi_version_and_constants(lib):
    src = open(HEADER).read()
    ver = int(re.search(r"#define ADLM_ABI_VERSION (\d+)", src).group(1))
    assert lib.adlm_abi_version() == ver == _lib.ABI_VERSION
    for macro, val in (("ADLM_PREC_FP32", _lib.PREC_FP32), ("ADLM_PREC_TF32", _lib.PREC_TF32),
                       ("ADLM_PREC_TF32X3", _lib.PREC_TF32X3), ("ADLM_ACT_LOG", _lib.ACT_LOG),
                       ("ADLM_ACT_LINEAR", _lib.ACT_LINEAR), ("ADLM_MAX_PROTOS_PER_CLASS", _lib.MAX_PROTOS_PER_CLASS),
                       ("ADLM_MAX_PUSH_CLASSES", _lib.MAX_PUSH_CLASSES), ("ADLM_STEP_NSCALARS", _lib.STEP_NSCALARS)):
        assert int(re.search(rf"#define {macro} (-?\d+)", src).group(1)) == val, macro
    assert int(re.search(r"#define ADLM_PUSH_NONE (0x[0-9a-fA-F]+)ULL", src).group(1), 16) == _lib.PUSH_NONE


def test_argument_errors_are_reported_without_a_gpu(lib):
    # null pointers / bad sizes are rejected before any CUDA call
    rc = lib.adlm_head_fwd(None, 1, 64, 16, None, None, 10, 2, ctypes.c_float(1e-4), 0, 0, None, None, None, None, None)
    assert rc == _lib.E_INVALID and b"non-null" in lib.adlm_last_error()
    rc = lib.adlm_head_fwd(None, 1, 1024, 16, None, None, 10, 2, ctypes.c_float(1e-4), 0, 0, None, None, None, None, None)
    assert rc == _lib.E_UNSUPPORTED
    with pytest.raises(NotImplementedError):
        _lib.check(rc, "adlm_head_fwd")
    assert lib.adlm_step_workspace_bytes(16, 64, 4225, 190, 19) > 16 * 4225 * 16 * 4


def test_dp_exchange_layout_is_consistent(lib):
    """host-only entry points of the peer-memory data-parallel step (no GPU needed)"""
    assert int(re.search(r"#define ADLM_DP_MAX_WORLD (\d+)", open(HEADER).read()).group(1)) == _lib.DP_MAX_WORLD
    K, N, C = 64, 190, 19
    for world_size in (1, 2, 8, _lib.DP_MAX_WORLD):
        out = (ctypes.c_size_t * 8)()
        assert lib.adlm_dp_exchange_layout(K, N, C, world_size, ctypes.cast(out, ctypes.c_void_p)) == 0
        cnt, cflag, acc, gflag, grad, total, n_grad_pad, n_blocks = (int(x) for x in out)
        assert total == lib.adlm_dp_exchange_bytes(K, N, C, world_size)
        assert 0 == cnt < cflag < acc < gflag < grad < total and all(o % 256 == 0 for o in (cflag, acc, gflag, grad, total))
        assert n_grad_pad >= N * K + C * N and n_grad_pad % 64 == 0
        assert n_blocks * 64 >= N * K + C * N                        # 64 outputs per 256-thread block
        # regions hold [2 parities][world][...] without overlap
        assert cflag - cnt >= 2 * world_size * 2 * 4 and acc - cflag >= 2 * world_size * 4
        assert gflag - acc >= 2 * world_size * 4 * 8 and grad - gflag >= 2 * world_size * n_blocks * 4
        assert total - grad >= 2 * world_size * n_grad_pad * 4
    assert lib.adlm_dp_exchange_bytes(K, N, C, _lib.DP_MAX_WORLD + 1) == 0
    assert lib.adlm_dp_exchange_layout(K, N, C, 0, None) == _lib.E_INVALID
    comm = _lib.DpComm(0, 1, 0, None, 0)
    rc = lib.adlm_step_fwd_bwd_dp(None, 1, 64, 16, None, None, 10, 2, ctypes.c_float(1e-4), 1, None, None, 1,
                                  ctypes.c_float(1), ctypes.c_float(1), ctypes.c_float(1), ctypes.c_float(1),
                                  None, None, None, None, None, None, None, 0, ctypes.byref(comm), None)
    assert rc == _lib.E_INVALID
