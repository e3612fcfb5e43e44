"""CPU-side checks of the C ABI: the library loads, exports every symbol the header
declares, and rejects bad arguments without touching a GPU."""
import ctypes as C
import os

import pytest

import alpha_zero_b200 as az


def test_library_exports_every_declared_symbol():
    if not os.path.exists(az.lib.SO_PATH):
        import __graft_entry__
        __graft_entry__.build()
    L = az.lib.load()
    names = az.lib.declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), n
    assert b"sm_100a" in L.az_version()


def test_bad_arguments_are_rejected_before_any_cuda_call():
    L = az.lib.load()
    h = C.c_void_p()
    assert L.az_forest_create(7, 3, 3, 4, 10, 1, C.byref(h)) == -1          # unknown game
    assert b"unknown game" in L.az_last_error()
    assert L.az_forest_create(1, 7, 7, 4, 10, 1, C.byref(h)) == -1          # connect4 is 6x7
    assert L.az_forest_create(2, 16, 16, 4, 10, 1, C.byref(h)) == -1        # gomoku n <= 15
    assert L.az_forest_create(0, 3, 3, 0, 10, 1, C.byref(h)) == -1          # n_trees > 0
    with pytest.raises(az.lib.AzError):
        az.lib.check(-1, "probe")


def test_host_bitboard_packing():
    import numpy as np
    from alpha_zero_b200.forest import planes_to_words, game_spec
    g = game_spec("gomoku", 15)
    p = np.zeros((2, 2, 15, 15), np.float32)
    p[0, 0, 0, 0] = 1
    p[0, 1, 14, 14] = 1
    p[1, 0, 4, 4] = 1          # cell 64 -> word 1 bit 0
    w = planes_to_words(g, p)
    assert w.shape == (2, 2, 4)
    assert w[0, 0, 0] == 1 and w[0, 1, 3] == (1 << (224 - 192)) and w[1, 0, 1] == 1
