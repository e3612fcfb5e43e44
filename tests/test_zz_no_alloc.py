"""az_rules_legal_noalloc (no counterpart in the reference): the variant of the
batched legal-move kernel whose scratch is a static __device__ array, for callers that may not call cudaMalloc.

Called through the C ABI with device buffers owned by the test, twice in a row on different batches (the second call
finds the first call's bytes in the shared scratch), and compared with the existing entry point az_rules_legal --
bit-exact, the bar of every integer / index test in this repository.  Collected last on purpose (file name).
Runs alone in a fresh process as well as after the other tests."""
import ctypes as C

import numpy as np
import pytest

import az_oracle as orc

pytestmark = pytest.mark.gpu


def _random_boards(g, n, rng, max_stones):
    """uint64 [n][2][nw] side bitboards with 0..max_stones stones placed at random (any position is a valid input of
    get_valid_moves; clustered and scattered stones exercise every CPython-set table size)."""
    ncell = g.rows * g.cols
    nw = 1 if ncell <= 64 else 4
    words = np.zeros((n, 2, nw), np.uint64)
    for t in range(n):
        k = int(rng.integers(0, max_stones + 1))
        cells = rng.permutation(ncell)[:k]
        if g.gid == 1:                                              # Connect4: stones stack from the bottom row
            heights = np.zeros(g.cols, int)
            cells = []
            for _ in range(k):
                c = int(rng.integers(g.cols))
                if heights[c] < g.rows:
                    cells.append((g.rows - 1 - heights[c]) * g.cols + c)
                    heights[c] += 1
        for i, c in enumerate(cells):
            words[t, i & 1, int(c) >> 6] |= np.uint64(1) << np.uint64(int(c) & 63)
    return words


def _reference(L, lib, g, words):
    n = words.shape[0]
    counts = np.zeros(n, np.int32)
    cells = np.zeros((n, g.rows * g.cols), np.uint8)
    lib.check(L.az_rules_legal(g.gid, g.rows, g.cols, n, words.ctypes.data_as(C.c_void_p), counts.ctypes.data_as(C.c_void_p),
                               cells.ctypes.data_as(C.c_void_p)), "az_rules_legal")
    return counts, cells


def _noalloc(L, lib, g, words, stream=None):
    import torch
    n = words.shape[0]
    w = torch.from_numpy(words.view(np.int64)).cuda()
    counts = torch.full((n,), -7, dtype=torch.int32, device="cuda")
    cells = torch.full((n, g.rows * g.cols), 255, dtype=torch.uint8, device="cuda")
    lib.check(L.az_rules_legal_noalloc(g.gid, g.rows, g.cols, n, w.data_ptr(), counts.data_ptr(), cells.data_ptr(), stream),
              "az_rules_legal_noalloc")
    torch.cuda.synchronize()
    return counts.cpu().numpy(), cells.cpu().numpy()


def _same(g, got, ref):
    (gc, gcell), (rc, rcell) = got, ref
    assert np.array_equal(gc, rc), "counts differ"
    for t in range(len(rc)):                                        # cells beyond the count are unspecified in both
        assert np.array_equal(gcell[t, :rc[t]], rcell[t, :rc[t]]), f"position {t}: legal moves / their order differ"


@pytest.mark.parametrize("name,rows,n1,n2,stones", [("gomoku", 7, 3000, 517, 30), ("gomoku", 15, 1500, 2048, 60), ("connect4", 6, 4096, 100, 30),
                                                    ("tictactoe", 3, 700, 900, 8), ("gomoku", 9, 1000, 1000, 40)])
def test_noalloc_matches_az_rules_legal_on_two_calls_in_a_row(name, rows, n1, n2, stones):
    import alpha_zero_b200 as az
    lib = az.lib
    L = lib.load()
    g = {"tictactoe": orc.TICTACTOE, "connect4": orc.CONNECT4}.get(name) or orc.gomoku_spec(rows)
    rng = np.random.default_rng(1000 + rows)
    first, second = _random_boards(g, n1, rng, stones), _random_boards(g, n2, rng, stones)
    _same(g, _noalloc(L, lib, g, first), _reference(L, lib, g, first))
    _same(g, _noalloc(L, lib, g, second), _reference(L, lib, g, second))      # the shared scratch holds the first call's bytes
    _same(g, _noalloc(L, lib, g, first), _reference(L, lib, g, first))


def test_noalloc_on_a_user_stream_and_back_to_back_without_sync():
    """Two calls queued on one stream (stream order serialises them) with different games: each output is its own."""
    import torch
    import alpha_zero_b200 as az
    lib = az.lib
    L = lib.load()
    rng = np.random.default_rng(5)
    ga, gb = orc.GOMOKU, orc.gomoku_spec(15)
    wa, wb = _random_boards(ga, 2000, rng, 25), _random_boards(gb, 1000, rng, 80)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        ta = torch.from_numpy(wa.view(np.int64)).cuda()
        tb = torch.from_numpy(wb.view(np.int64)).cuda()
        ca = torch.zeros(2000, dtype=torch.int32, device="cuda")
        cb = torch.zeros(1000, dtype=torch.int32, device="cuda")
        xa = torch.zeros((2000, 49), dtype=torch.uint8, device="cuda")
        xb = torch.zeros((1000, 225), dtype=torch.uint8, device="cuda")
    s.synchronize()
    lib.check(L.az_rules_legal_noalloc(ga.gid, 7, 7, 2000, ta.data_ptr(), ca.data_ptr(), xa.data_ptr(), s.cuda_stream), "noalloc a")
    lib.check(L.az_rules_legal_noalloc(gb.gid, 15, 15, 1000, tb.data_ptr(), cb.data_ptr(), xb.data_ptr(), s.cuda_stream), "noalloc b")
    s.synchronize()
    _same(ga, (ca.cpu().numpy(), xa.cpu().numpy()), _reference(L, lib, ga, wa))
    _same(gb, (cb.cpu().numpy(), xb.cpu().numpy()), _reference(L, lib, gb, wb))


def test_noalloc_rejects_batches_beyond_the_static_scratch():
    import torch
    import alpha_zero_b200 as az
    lib = az.lib
    L = lib.load()
    g = orc.gomoku_spec(15)
    n = 4089                                                        # limit for 15x15: 4 MiB / (2 * 513) = 4088 positions
    w = torch.zeros((n, 2, 4), dtype=torch.int64, device="cuda")
    counts = torch.zeros(n, dtype=torch.int32, device="cuda")
    cells = torch.zeros((n, 225), dtype=torch.uint8, device="cuda")
    rc = L.az_rules_legal_noalloc(g.gid, 15, 15, n, w.data_ptr(), counts.data_ptr(), cells.data_ptr(), None)
    assert rc == -1 and b"4088" in L.az_last_error() and b"static array" in L.az_last_error()
    assert L.az_rules_legal_noalloc(g.gid, 15, 15, n - 1, w.data_ptr(), counts.data_ptr(), cells.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert (counts[:n - 1].cpu().numpy() == 1).all()                # empty boards: the centre cell only (gomoku.py:104-105)
    assert (cells[:n - 1, 0].cpu().numpy() == 7 * 15 + 7).all()
    assert L.az_rules_legal_noalloc(g.gid, 15, 15, 4, None, counts.data_ptr(), cells.data_ptr(), None) == -1      # null input
