"""Synthetic root positions for benchmarks: n boards after 0..max_plies random legal plies
(legal as the reference defines it, e.g. Gomoku moves touch an existing stone).  Vectorised
NumPy; the boards are float32 planes exactly as the reference holds them."""
import numpy as np


def random_roots(Game, n, max_plies, rng):
    rows, cols = Game.rows, Game.cols
    C = Game.feature_dim
    stones = np.zeros((n, 2, rows, cols), dtype=bool)
    target = rng.integers(0, max_plies + 1, size=n)
    idx = np.arange(n)
    for ply in range(max_plies):
        side = ply & 1
        occ = stones[:, 0] | stones[:, 1]
        if Game.name == "connect4":
            legal = np.zeros((n, rows, cols), dtype=bool)
            height = occ.sum(axis=1)                                   # stones per column
            r = rows - 1 - height                                      # drop row
            ok = height < rows
            nn, cc = np.nonzero(ok)
            legal[nn, r[nn, cc], cc] = True
        elif Game.name == "gomoku":
            near = np.zeros_like(occ)
            pad = np.pad(occ, ((0, 0), (1, 1), (1, 1)))
            for dr in (0, 1, 2):
                for dc in (0, 1, 2):
                    near |= pad[:, dr:dr + rows, dc:dc + cols]
            legal = near & ~occ
            empty = ~occ.reshape(n, -1).any(axis=1)
            legal[empty, rows // 2, cols // 2] = True
        else:
            legal = ~occ
        score = rng.random((n, rows * cols)) * legal.reshape(n, -1)
        pick = score.argmax(axis=1)
        go = (target > ply) & legal.reshape(n, -1).any(axis=1)
        stones[idx[go], side, pick[go] // cols, pick[go] % cols] = True
        target = np.where(go, target, np.minimum(target, ply))
    planes = np.zeros((n, C, rows, cols), dtype=np.float32)
    planes[:, 0] = stones[:, 0]
    planes[:, 1] = stones[:, 1]
    players = (target & 1).astype(np.int32)
    if C == 3:
        planes[:, 2] = players[:, None, None]                          # plane 2 = side to move
    return planes, players, target.astype(np.int32)
