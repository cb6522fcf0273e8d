"""CPU: the C-ABI library loads and exports exactly what include/dllm_b200.h declares."""
import ctypes
import os
import re
import subprocess
import tempfile

import pytest
import torch

from src.engine import cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "dllm_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    return sorted(set(re.findall(r"DLLM_API[^;(]*?\b(dllm_\w+)\s*\(", text)))


def test_header_declares_symbols():
    syms = declared_symbols()
    assert "dllm_shard_forward" in syms and "dllm_last_error" in syms and len(syms) >= 15


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(cabi.LIB_PATH)
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} declared in dllm_b200.h but not exported"


def test_binding_covers_every_declared_symbol():
    assert sorted(cabi.SIGNATURES) == declared_symbols()


def test_no_stray_exports():
    out = subprocess.run(["nm", "-D", "--defined-only", cabi.LIB_PATH], capture_output=True, text=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l and "dllm_" in l.split()[-1])
    assert exported == declared_symbols()


def test_abi_version_and_error_string():
    lib = cabi.lib()
    assert lib.dllm_abi_version() == 2
    assert isinstance(lib.dllm_last_error(), bytes)
    assert lib.dllm_kernel_launch_count() >= 0


def test_header_compiles_as_c():
    r = subprocess.run(["gcc", "-std=c99", "-fsyntax-only", "-x", "c", HEADER], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_struct_layouts_match_header():
    # 11 int32 + float + 7 int32; 17 and 7 pointers.  The compiler's view of the header is the reference point.
    src = '#include <stdio.h>\n#include "dllm_b200.h"\nint main(void){printf("%zu %zu %zu", sizeof(dllm_shard_config), sizeof(dllm_layer_weights), sizeof(dllm_io_weights));return 0;}'
    exe = os.path.join(tempfile.mkdtemp(), "sz")
    r = subprocess.run(["gcc", "-std=c99", "-I", os.path.dirname(HEADER), "-x", "c", "-", "-o", exe], input=src, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    sizes = [int(v) for v in subprocess.run([exe], capture_output=True, text=True).stdout.split()]
    assert sizes == [ctypes.sizeof(cabi.ShardConfig), ctypes.sizeof(cabi.LayerWeights), ctypes.sizeof(cabi.IoWeights)]
    assert ctypes.sizeof(cabi.ShardConfig) == 19 * 4
    assert ctypes.sizeof(cabi.LayerWeights) == 17 * 8
    assert ctypes.sizeof(cabi.IoWeights) == 7 * 8


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the loud failure on a machine without CUDA")
def test_compute_fails_loudly_without_gpu():
    lib = cabi.lib()
    assert lib.dllm_device_count() == 0
    rc = lib.dllm_op_gemv(0, 1, 1, None, 1, 1, 8, 8, 0, None)
    assert rc != 0 and b"no CUDA device" in lib.dllm_last_error()
    cfg = cabi.ShardConfig(arch=1, dtype=0, hidden=64, n_heads=2, n_kv_heads=2, head_dim=32, ffn=64, vocab=16,
                           n_layers=0, has_embed=0, has_head=0, norm_eps=1e-5, max_rows=1, max_seqs=1,
                           max_context=8, page_size=8, num_pages=0)
    h = ctypes.c_void_p()
    rc = lib.dllm_shard_create(ctypes.byref(cfg), None, ctypes.byref(cabi.IoWeights()), ctypes.byref(h))
    assert rc != 0 and b"no CUDA device" in lib.dllm_last_error() and not h.value


def test_mailboxes_fail_loudly_without_gpu_and_reject_bad_shapes():
    """The fused hand-off has no host stand-in either: no device -> error; shapes are checked before any CUDA call."""
    import pytest
    from src.network.handoff import Mailbox
    lib = cabi.lib()
    h = ctypes.c_void_p()
    assert lib.dllm_mailbox_create(0, 0, 1, 64, ctypes.byref(h)) != 0 and b"bad shape" in lib.dllm_last_error()
    assert lib.dllm_mailbox_create(0, 1, 17, 64, ctypes.byref(h)) != 0      # more rows than the persistent kernel takes
    assert lib.dllm_mailbox_create(0, 1, 1, 60, ctypes.byref(h)) != 0       # hidden must be a multiple of 8
    assert lib.dllm_mailbox_create(0, 2, 1, 64, ctypes.byref(h)) != 0 and b"no CUDA device" in lib.dllm_last_error()
    assert not h.value
    with pytest.raises(cabi.DllmError):
        Mailbox.create(0, 2, 1, 64)
    assert lib.dllm_shard_decode_pipelined(None, None, None, 1, 1, 1, 0, None, None, None, None) != 0
    assert lib.dllm_mailbox_destroy(None) == 0


def test_create_rejects_bad_geometry():
    lib = cabi.lib()
    cfg = cabi.ShardConfig(arch=1, dtype=0, hidden=64, n_heads=3, n_kv_heads=2, head_dim=48, ffn=64, vocab=16,
                           n_layers=0, has_embed=0, has_head=0, norm_eps=1e-5, max_rows=1, max_seqs=1,
                           max_context=8, page_size=8, num_pages=0)
    h = ctypes.c_void_p()
    rc = lib.dllm_shard_create(ctypes.byref(cfg), None, ctypes.byref(cabi.IoWeights()), ctypes.byref(h))
    assert rc != 0 and b"head_dim" in lib.dllm_last_error()
