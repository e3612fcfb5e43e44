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
