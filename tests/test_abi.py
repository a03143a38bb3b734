"""CPU-side checks of the C-ABI boundary: the library builds, loads, and exports every symbol
include/b200cnn.h declares; argument validation works without a GPU (no compute calls)."""
import ctypes as C

import pytest

import b200cnn
from b200cnn import _lib


def test_every_declared_symbol_is_exported():
    lib = _lib.load()
    names = [n for _, n, _ in _lib.declared_functions()]
    assert len(names) >= 25 and len(set(names)) == len(names)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200cnn.h but not exported"
    assert lib.b200_version() >= 100


def test_create_validates_config_without_gpu():
    lib = _lib.load()
    h = C.c_void_p()
    bad = _lib.Config(0, 32, 32, 24, 24, 3, 64, 64, 384, 192, 10, 0)
    assert lib.b200_create(C.byref(bad), C.byref(h)) == _lib.E_INVALID
    assert b"batch" in lib.b200_last_error()
    bad = _lib.Config(8, 32, 32, 40, 24, 3, 64, 64, 384, 192, 10, 0)
    assert lib.b200_create(C.byref(bad), C.byref(h)) == _lib.E_INVALID
    bad = _lib.Config(8, 32, 32, 24, 24, 3, 64, 64, 384, 192, 17, 0)
    assert lib.b200_create(C.byref(bad), C.byref(h)) == _lib.E_INVALID


def test_layout_matches_reference_variable_sizes():
    """cifar10cnn.py:106-144 -> 1,068,298 parameters; each tensor starts on a 64-float boundary."""
    lib = _lib.load()
    h = C.c_void_p()
    cfg = _lib.Config(128, 32, 32, 24, 24, 3, 64, 64, 384, 192, 10, 0)
    assert lib.b200_create(C.byref(cfg), C.byref(h)) == 0
    offs, sizes, total = (C.c_int64 * 10)(), (C.c_int64 * 10)(), C.c_int64()
    assert lib.b200_param_layout(h, offs, sizes, C.byref(total)) == 0
    assert list(sizes) == [4800, 64, 102400, 64, 884736, 384, 73728, 192, 1920, 10]
    assert sum(sizes) == 1068298
    assert all(o % 64 == 0 for o in offs) and total.value % 64 == 0
    nb = C.c_int64()
    assert lib.b200_workspace_bytes(h, C.byref(nb)) == 0 and nb.value > 0
    # calls before binding buffers fail with a state error, not a crash
    assert lib.b200_backward(h, None) == _lib.E_STATE
    assert lib.b200_bind_params(h, 12345) == _lib.E_INVALID      # misaligned pointer
    assert lib.b200_destroy(h) == 0


def test_engine_refuses_to_run_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        b200cnn.TrainStepEngine()


def test_trace_entry_point_reports_a_plain_build():
    """b200_debug_trace only works in a B200_TRACE=1 diagnostics build; the shipped library says so."""
    lib = _lib.load()
    assert lib.b200_debug_trace(None) == _lib.E_STATE
    assert b"B200_TRACE" in lib.b200_last_error()


def test_p2p_step_validates_its_tables_without_gpu():
    """b200_train_step_p2p (cifar10cnn.py:162-163 + :195-196 for world > 1) rejects missing peer tables before it
    touches the device."""
    lib = _lib.load()
    h = C.c_void_p()
    cfg = _lib.Config(128, 32, 32, 24, 24, 3, 64, 64, 384, 192, 10, 1)
    assert lib.b200_create(C.byref(cfg), C.byref(h)) == 0
    rc = lib.b200_train_step_p2p(h, None, None, None, None, None, None, None, None, 0, 2, 0.1, 1.0, None)
    assert rc != 0
    lib.b200_destroy(h)


def test_p2p_inbox_setter_validates_its_arguments_without_gpu():
    """b200_p2p_set_inbox (inbox of the one-shot conv-gradient exchange) stores pointers only: null handle, bad world
    size, missing peer table and a null peer entry are rejected; a null inbox switches the two-barrier kernel back on."""
    lib = _lib.load()
    assert lib.b200_p2p_set_inbox(None, None, None, None, 2, 0) != 0
    h = C.c_void_p()
    cfg = _lib.Config(128, 32, 32, 24, 24, 3, 64, 64, 384, 192, 10, 1)
    assert lib.b200_create(C.byref(cfg), C.byref(h)) == 0
    peers = (C.c_void_p * 2)(4096, 8192)
    assert lib.b200_p2p_set_inbox(h, 4096, None, peers, 2, 1 << 20) == 0
    assert lib.b200_p2p_set_inbox(h, None, None, None, 0, 0) == 0             # back to the two-barrier kernel
    assert lib.b200_p2p_set_inbox(h, 4096, None, None, 2, 1 << 20) == _lib.E_INVALID
    assert lib.b200_p2p_set_inbox(h, 4096, None, peers, 9, 1 << 20) == _lib.E_INVALID
    assert lib.b200_p2p_set_inbox(h, 4100, None, peers, 2, 1 << 20) == _lib.E_INVALID   # misaligned
    peers[1] = None
    assert lib.b200_p2p_set_inbox(h, 4096, None, peers, 2, 1 << 20) == _lib.E_INVALID
    assert b"rank 1" in lib.b200_last_error()
    lib.b200_destroy(h)
