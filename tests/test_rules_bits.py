"""The win test of the device rules on one-word boards is a shift-and-mask expression (csrc/az_rules.cuh: az_is_win_bits); the
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
