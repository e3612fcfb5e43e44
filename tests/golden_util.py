"""Helpers shared by the golden-vector tests (oracle on CPU, CUDA path on GPU)."""
import json
import os

import numpy as np

import az_oracle as orc

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_json(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def spec_of(case):
    if case["game"] == "gomoku":
        return orc.GOMOKU if case["rows"] == 7 else orc.gomoku_spec(case["rows"])
    return orc.SPECS[case["game"]]


def board_from_words(g, p0_hex, p1_hex, player_to_move=None):
    b = orc.new_board(g)
    for plane, words in ((0, p0_hex), (1, p1_hex)):
        for w, hx in enumerate(words):
            v = int(hx, 16)
            i = 0
            while v:
                if v & 1:
                    cell = w * 64 + i
                    b[plane, cell // g.cols, cell % g.cols] = 1
                v >>= 1
                i += 1
    if g.planes == 3 and player_to_move is not None:
        b[2, :, :] = player_to_move
    return b


def position_after(g, opening):
    b = orc.new_board(g)
    player = 0
    for mv in opening:
        player = orc.make_move(g, b, player, tuple(mv))
    return b, player, len(opening)


def fhex(x):
    return float.fromhex(x)
