"""The drop-in boundary without a GPU: `libb2seg.so` loads, exports every function
`include/b2seg.h` declares (and nothing the header does not know about), reports the ABI version
the Python loader expects, and rejects bad arguments with a message instead of crashing.  No
compute calls: those are the `-m gpu` tests."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b2seg.h")
LIB = os.path.join(ROOT, "image-segmentation-project_b200", "b2seg", "libb2seg.so")


def _declared():
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    return sorted(set(re.findall(r"\b(b2seg_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        pytest.fail(f"{LIB} is missing: run `python __graft_entry__.py` (build()) first")
    return C.CDLL(LIB)


def test_header_declares_the_entry_points_the_docs_name():
    names = _declared()
    for must in ("b2seg_abi_version", "b2seg_last_error", "b2seg_seg_loss_fwd", "b2seg_seg_loss_bwd",
                 "b2seg_seg_loss_finish", "b2seg_seg_metrics_from_probs", "b2seg_aug_permute_u8",
                 "b2seg_aug_warp_affine_u8", "b2seg_aug_remap_u8", "b2seg_tapgemm_create", "b2seg_wgrad_create",
                 "b2seg_ew_create", "b2seg_adam_step", "b2seg_seq_create", "b2seg_seq_append_lane",
                 "b2seg_seq_join", "b2seg_seq_run", "b2seg_op_run"):
        assert must in names, must


def test_library_exports_every_declared_symbol(lib):
    missing = [n for n in _declared() if not hasattr(lib, n)]
    assert not missing, missing


def test_no_undeclared_public_symbols():
    out = subprocess.run(["nm", "-D", "--defined-only", LIB], capture_output=True, text=True, check=True).stdout
    exported = sorted({ln.split()[-1] for ln in out.splitlines() if " T " in ln and ln.split()[-1].startswith("b2seg_")})
    undeclared = [n for n in exported if n not in _declared()]
    assert not undeclared, undeclared


def test_abi_version_matches_the_python_loader(lib):
    import sys
    sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200"))
    from b2seg import _lib as L
    assert lib.b2seg_abi_version() == L.ABI_VERSION
    src = open(HEADER).read()
    assert int(re.search(r"#define\s+B2SEG_ABI_VERSION\s+(\d+)", src).group(1)) == L.ABI_VERSION


def test_bad_arguments_return_an_error_code_and_a_message(lib):
    lib.b2seg_last_error.restype = C.c_char_p
    lib.b2seg_seg_reduce_workspace_bytes.restype = C.c_size_t
    assert lib.b2seg_seg_reduce_workspace_bytes() > 0
    # argument validation happens on the host before any CUDA call
    rc = lib.b2seg_op_run(None, None)
    assert rc != 0 and b"null" in lib.b2seg_last_error()
    rc = lib.b2seg_tapgemm_create(None, None)
    assert rc != 0 and lib.b2seg_last_error()
    rc = lib.b2seg_adam_step(None, 0, C.c_float(1e-3), C.c_float(0.9), C.c_float(0.999), C.c_float(1e-8),
                             C.c_float(0.0), C.c_double(1.0), None)
    assert rc != 0 and b"adam_step" in lib.b2seg_last_error()
