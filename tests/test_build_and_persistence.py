"""(1) A cold compile of a CUDA unit works here without a GPU and the build gate is by content, not mtime.
(2) Persistence formats round-trip with the REFERENCE's own classes (utils.save_model / load_model, utils.py:57-69;
ReplayBuffer.save_pickle / load_pickle, replay_buffer.py:37-65): files written by the unmodified reference load here and
files written here load there.  The reference runs in a subprocess from baseline/_ref (the scratch copy __graft_entry__.build()
makes; importing its `games` package writes logs/ next to it).  Skipped where that copy does not exist."""
import ctypes as C
import importlib.util
import os
import pickle
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")


def _build_mod():
    spec = importlib.util.spec_from_file_location("az_build", os.path.join(ROOT, "alpha-zero_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_cold_compile_of_one_unit_and_content_gate(tmp_path):
    mod = _build_mod()
    out = str(tmp_path / "libaz_cold.so")
    units = ["az_noalloc.cu", "az_forest.cu"]                       # az_noalloc needs az_fail / az_make_game from az_forest
    mod.build(force=True, units=units, out=out, obj_dir=str(tmp_path))
    assert os.path.getsize(out) > 100000
    L = C.CDLL(out, mode=os.RTLD_LAZY)                             # az_forest.o calls into az_net.cu (minutes to compile: not part of this test)
    for name in ("az_rules_legal_noalloc", "az_forest_create", "az_version"):
        assert hasattr(L, name)
    dump = subprocess.run(["cuobjdump", "-lelf", out], capture_output=True, text=True).stdout
    assert "sm_100a" in dump                                        # compiled for Blackwell, nothing else
    stamp = open(str(tmp_path / "az_noalloc.o.sha256")).read()
    t0 = os.path.getmtime(str(tmp_path / "az_noalloc.o"))
    os.utime(os.path.join(mod.CSRC, "az_noalloc.cu"))               # a touched file with the same content: nothing to do
    mod.build(units=units, out=out, obj_dir=str(tmp_path))
    assert os.path.getmtime(str(tmp_path / "az_noalloc.o")) == t0 and open(str(tmp_path / "az_noalloc.o.sha256")).read() == stamp


def _ref_copy(tmp_path):
    """A private copy of the reference tree: its save / load functions work relative to their own module directory
    (utils.py:59,67; replay_buffer.py:57), so models/ and replay_buffers/ land inside this copy."""
    import shutil
    dst = str(tmp_path / "ref")
    shutil.copytree(REF, dst, ignore=shutil.ignore_patterns("__pycache__", "logs", "models", "replay_buffers"))
    return dst


def _ref_run(code, refdir):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", PYTHONPATH="")
    return subprocess.run([sys.executable, "-c", "import sys; sys.path.insert(0, %r)\n" % refdir + code], cwd=refdir, env=env,
                          capture_output=True, text=True, timeout=300)


needs_ref = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "games")), reason="baseline/_ref (copy of the reference) is not present")


@needs_ref
def test_checkpoints_round_trip_with_the_reference(tmp_path):
    import torch
    from alpha_zero_b200 import utils
    from alpha_zero_b200.ai import Net
    ref = _ref_copy(tmp_path)
    r = _ref_run("""
import torch, numpy as np, utils
from ai.nn import Net
torch.manual_seed(5)
net = Net(7, 3, 32, 49, 4, 2, 2, 0.1).eval()
utils.save_model(net, 3)                                   # -> ./models/model_v3.pth
x = torch.from_numpy(np.random.RandomState(1).randint(0, 2, (6, 2, 7, 7)).astype(np.float32))
with torch.no_grad():
    l, v = net(x)
np.savez('ref_out.npz', x=x.numpy(), l=l.numpy(), v=v.numpy())
""", ref)
    assert r.returncode == 0, r.stderr[-2000:]
    arr = np.load(os.path.join(ref, "ref_out.npz"))
    net = Net(7, 3, 32, 49, 4, 2, 2, 0.1).eval()
    utils.load_model(net, "model_v3.pth", directory=os.path.join(ref, "models"))
    with torch.no_grad():
        l, v = net(torch.from_numpy(arr["x"]))
    assert np.allclose(l.numpy(), arr["l"], atol=1e-6) and np.allclose(v.numpy(), arr["v"], atol=1e-6)
    # and back: a checkpoint written here loads into the reference's Net with its load_model
    torch.manual_seed(6)
    mine = Net(7, 3, 32, 49, 4, 2, 2, 0.1).eval()
    utils.save_model(mine, 9, directory=os.path.join(ref, "models"))
    x = torch.from_numpy(arr["x"])
    with torch.no_grad():
        ml, mv = mine(x)
    np.savez(os.path.join(ref, "mine_out.npz"), l=ml.numpy(), v=mv.numpy())
    r = _ref_run("""
import torch, numpy as np, utils
from ai.nn import Net
net = Net(7, 3, 32, 49, 4, 2, 2, 0.1).eval()
utils.load_model(net, 'model_v9.pth')
a = np.load('ref_out.npz'); m = np.load('mine_out.npz')
with torch.no_grad():
    l, v = net(torch.from_numpy(a['x']))
assert np.allclose(l.numpy(), m['l'], atol=1e-6) and np.allclose(v.numpy(), m['v'], atol=1e-6)
print('ok')
""", ref)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-2000:]


@needs_ref
def test_replay_buffer_pickles_round_trip_with_the_reference(tmp_path):
    from alpha_zero_b200.replay_buffer import ReplayBuffer
    ref = _ref_copy(tmp_path)
    bufdir = os.path.join(ref, "replay_buffers")
    r = _ref_run("""
import os, numpy as np
from replay_buffer import ReplayBuffer
rs = np.random.RandomState(2)
b = ReplayBuffer(50)
for i in range(60):                                        # maxlen drops the 10 oldest
    b.add(rs.randint(0, 2, (2, 7, 7)).astype(np.float32), rs.dirichlet([1.0] * 49), [int(rs.randint(-1, 2))])
b.save_pickle(os.path.join(os.getcwd(), 'replay_buffers', 'replay_buffer_v1_i_0.pkl'))
""", ref)
    assert r.returncode == 0, r.stderr[-2000:]
    mine = ReplayBuffer(50)
    mine.load_pickle("replay_buffer_v1_i_0", directory=bufdir)
    assert mine.size() == 50
    rs = np.random.RandomState(2)
    want = [(rs.randint(0, 2, (2, 7, 7)).astype(np.float32), rs.dirichlet([1.0] * 49), [int(rs.randint(-1, 2))]) for _ in range(60)][10:]
    for (s, p, z), (ws, wp, wz) in zip(mine.buffer, want):
        assert s.dtype == np.float32 and np.array_equal(s, ws) and np.array_equal(p, wp) and list(z) == wz
    s, p, z = mine.sample(16)
    assert tuple(s.shape) == (16, 2, 7, 7) and tuple(p.shape) == (16, 49) and tuple(z.shape) == (16, 1)
    # and back: written here (older replay_buffer* files in the folder are removed, as the reference does), read there
    mine.save_pickle(os.path.join(bufdir, "replay_buffer_v2_i_0.pkl"))
    assert sorted(os.listdir(bufdir)) == ["replay_buffer_v2_i_0.pkl"]
    with open(os.path.join(bufdir, "replay_buffer_v2_i_0.pkl"), "rb") as f:
        assert type(pickle.load(f)).__name__ == "deque"
    r = _ref_run("""
import numpy as np
from replay_buffer import ReplayBuffer
b = ReplayBuffer(50)
b.load_pickle('replay_buffer_v2_i_0')
assert b.size() == 50
s, p, z = b.sample(8)
assert tuple(s.shape) == (8, 2, 7, 7) and str(s.dtype) == 'torch.float32'
print('ok')
""", ref)
    assert r.returncode == 0 and "ok" in r.stdout, (r.stdout[-500:], r.stderr[-2000:])
