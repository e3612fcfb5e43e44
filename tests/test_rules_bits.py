"""Shift-and-mask forms of the device rules against the cell walks they replace on one-word boards (host code, built with nvcc).

The win test of the device rules on one-word boards is a shift-and-mask expression (csrc/az_rules.cuh: az_is_win_bits); the
reference's check_winner walks the four lines through the last move (games/connect4.py:72-98).  tests/native/win_test.cu holds
both forms as host code and compares them on 2.4 M random positions (TicTacToe, Connect4, Gomoku 5x5 .. 8x8)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shift_and_mask_win_test_equals_line_walk(tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not found")
    exe = str(tmp_path / "win_test")
    subprocess.check_call([nvcc, "-O2", "-std=c++17", "-I", os.path.join(ROOT, "alpha-zero_b200", "csrc"), "-o", exe,
                           os.path.join(ROOT, "tests", "native", "win_test.cu")])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "identical on 2400000 positions" in out.stdout, out.stdout + out.stderr


def test_shift_and_mask_rollout_legal_sets_equal_cell_walk(tmp_path):
    """Vanilla-mode rollouts (ai/mcts.py:62-79) draw uniformly from a legal SET; tests/native/legal_set_test.cu compares the
    bitboard expression (az_rollout_legal_bits) with the cell walk on 1.8 M random positions."""
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not found")
    exe = str(tmp_path / "legal_set_test")
    subprocess.check_call([nvcc, "-O2", "-std=c++17", "-I", os.path.join(ROOT, "alpha-zero_b200", "csrc"), "-o", exe,
                           os.path.join(ROOT, "tests", "native", "legal_set_test.cu")])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "identical legal sets on 1800000 positions" in out.stdout, out.stdout + out.stderr
