"""ReplayBuffer with the reference's contract (replay_buffer.py:7-65): a bounded FIFO of
(state float32[C,H,W], policy float64[A], [z]) and uniform sampling without replacement into
three float32 tensors [B,C,H,W], [B,A], [B,1].  Pickle files hold the plain deque, so buffers
written by the reference load here and vice versa."""
import os
import pickle
from collections import deque

import numpy as np
import torch


class ReplayBuffer:
    def __init__(self, buffer_size):
        self.buffer_size = buffer_size
        self.buffer = deque(maxlen=buffer_size)

    def add(self, state, policy_distribution, reward):
        self.buffer.append((state, policy_distribution, reward))

    def extend(self, samples):
        self.buffer.extend(samples)

    def size(self):
        return len(self.buffer)

    def sample(self, batch_size):
        picks = np.random.choice(np.arange(len(self.buffer)), size=batch_size, replace=False)
        states = np.empty((batch_size,) + self.buffer[0][0].shape, dtype=np.float32)
        policies = np.empty((batch_size, len(self.buffer[0][1])), dtype=np.float32)
        rewards = np.empty((batch_size, 1), dtype=np.float32)
        for row, i in enumerate(picks):
            s, p, z = self.buffer[i]
            states[row], policies[row], rewards[row, 0] = s, p, z[0]
        return torch.from_numpy(states), torch.from_numpy(policies), torch.from_numpy(rewards)

    def head(self, size):
        for i in range(size):
            print(self.buffer[i])

    def tail(self, size):
        for i in range(1, size + 1):
            print(self.buffer[-i])

    def save_pickle(self, filename):
        folder = os.path.dirname(filename)
        if folder and os.path.exists(folder):
            for name in os.listdir(folder):             # keep only the newest buffer file
                path = os.path.join(folder, name)
                if name.startswith("replay_buffer") and os.path.isfile(path):
                    os.remove(path)
        elif folder:
            os.makedirs(folder)
        try:
            with open(filename, "wb") as f:
                pickle.dump(self.buffer, f)
            print(f"Replay Buffer saved to {filename} successfully.")
        except Exception as e:                          # noqa: BLE001 -- reference behaviour
            print("Failed to save Replay Buffer.", e)

    def load_pickle(self, filename, directory=None):
        directory = directory or os.path.join(os.getcwd(), "replay_buffers")
        try:
            with open(os.path.join(directory, f"{filename}.pkl"), "rb") as f:
                self.buffer = deque(pickle.load(f), maxlen=self.buffer_size)
            print(f"Replay Buffer loaded from {filename} successfully.")
        except Exception:                               # noqa: BLE001 -- reference behaviour
            print("Failed to load Replay Buffer.")


class DeviceReplayBuffer:
    """The same contract with the entries resident in HBM (csrc/az_replay.cu): a ring of `buffer_size` entries in the
    reference's order, filled straight from the self-play engine's trajectory arrays (add_trajectories: z assignment,
    canonicalisation and the 8 dihedral variants of train.save_data_to_buffer run as one kernel), sampled without
    replacement by a keyed permutation into three float32 CUDA tensors [B,C,H,W], [B,A], [B,1] -- no host copy on either
    side.  to_host() rebuilds the reference's deque (for save_pickle)."""

    def __init__(self, Game, buffer_size, seed=None):
        import ctypes as C
        from . import lib as _lib
        from .forest import game_spec
        self._lib, self._C = _lib, C
        self._L = _lib.load()
        self.g = game_spec(Game.name, Game.rows)
        self.buffer_size = int(buffer_size)
        h = C.c_void_p()
        _lib.check(self._L.az_replay_create(self.g.gid, self.g.rows, self.g.cols, self.buffer_size, C.byref(h)), "az_replay_create")
        self.h = h
        self.seed = int(np.random.randint(0, 2 ** 31 - 1)) if seed is None else int(seed)
        self.draws = 0

    def close(self):
        if getattr(self, "h", None):
            self._L.az_replay_destroy(self.h)
            self.h = None

    __del__ = close

    def size(self):
        return int(self._L.az_replay_size(self.h))

    def entry_bytes(self):
        return int(self._L.az_replay_entry_bytes(self.h))

    def clear(self):
        self._lib.check(self._L.az_replay_clear(self.h), "az_replay_clear")

    def add_trajectories(self, t):
        """t: dict with CUDA tensors plies / winner / first (int32 [G]), words (int64 [G,P,2,nw]), visits (int16 [G,P,A]) --
        selfplay.SelfPlayEngine.device_tensors() or parallel.gather_trajectories(); games are stored in array order.
        Returns the number of entries appended."""
        stream = torch.cuda.current_stream().cuda_stream
        arrs = [t[k].contiguous() for k in ("plies", "winner", "first", "words", "visits")]
        assert all(a.is_cuda for a in arrs) and arrs[0].dtype == torch.int32
        n = self._lib.check(self._L.az_replay_add(self.h, int(arrs[0].shape[0]), int(arrs[3].shape[1]), *[a.data_ptr() for a in arrs], stream),
                            "az_replay_add")
        return int(n)

    def _out(self, n):
        g = self.g
        dev = torch.device("cuda", torch.cuda.current_device())
        return (torch.empty((n, g.planes, g.rows, g.cols), dtype=torch.float32, device=dev),
                torch.empty((n, g.action_dim), dtype=torch.float32, device=dev), torch.empty((n, 1), dtype=torch.float32, device=dev))

    def sample(self, batch_size, return_indices=False):
        s, p, z = self._out(batch_size)
        idx = torch.empty(batch_size, dtype=torch.int64, device=s.device) if return_indices else None
        self.draws += 1
        self._lib.check(self._L.az_replay_sample(self.h, int(batch_size), self._C.c_uint64(self.seed), self._C.c_uint64(self.draws),
                                                 s.data_ptr(), p.data_ptr(), z.data_ptr(), idx.data_ptr() if return_indices else None,
                                                 torch.cuda.current_stream().cuda_stream), "az_replay_sample")
        return (s, p, z, idx) if return_indices else (s, p, z)

    def read(self, start=0, count=None):
        count = self.size() - start if count is None else count
        s, p, z = self._out(count)
        self._lib.check(self._L.az_replay_read(self.h, int(start), int(count), s.data_ptr(), p.data_ptr(), z.data_ptr(),
                                               torch.cuda.current_stream().cuda_stream), "az_replay_read")
        return s, p, z

    def to_host(self):
        """ReplayBuffer holding the same entries.  (The policy comes back as the float32 quotient widened to float64 --
        the reference's deque holds the float64 quotient; sample() rounds both to the same float32.)"""
        out = ReplayBuffer(self.buffer_size)
        if self.size():
            s, p, z = (x.cpu().numpy() for x in self.read())
            for i in range(len(s)):
                out.add(s[i], p[i].astype(np.float64), [int(z[i, 0])])
        return out
